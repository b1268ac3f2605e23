"""The bench.py contract that can be checked without a GPU: the `--impl reference` arm (the CPU restatement of the
reference path, timed on the host cores) prints ONE JSON line with the agreed keys, on the metric / config of the
GPU arm; ranks other than 0 print nothing."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, *flags):
    env = dict(os.environ)
    env.update(extra_env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '1', '--warmup', '0',
                          '--cpu-batch', '1', '--cpu-n-z', '2'] + list(flags), cwd=ROOT, env=env, capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return [l for l in out.stdout.splitlines() if l.startswith('{')]


def test_reference_arm_prints_one_contract_line():
    lines = _run()
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j['impl'] == 'reference'
    assert j['metric'] == 'importance-sampled NLL images/sec (K=1000, CIFAR-shape)' and j['unit'] == 'images/s'
    for k in ('value', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling', 'vs_baseline', 'dtype',
              'data', 'config', 'cpu_baseline', 'e2e'):
        assert k in j, k
    assert j['higher_is_better'] is True and j['vs_baseline'] is None and j['data'] == 'synthetic'
    assert 'configs[1]' in j['config']['workload'] and j['config']['n_z'] == 1000
    cb = j['cpu_baseline']
    assert cb['kind'] == 'port' and cb['cores'] >= 1 and cb['value'] == j['value'] and cb['value'] > 0
    assert 'B=1 images x K=2' in cb['sample']
    assert j['e2e'] == {'value': j['value'], 'unit': 'images/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}


def test_reference_arm_other_ranks_exit_quietly_and_labels_follow_the_config():
    assert _run({'RANK': '1', 'WORLD_SIZE': '2', 'LOCAL_RANK': '1'}) == []
    j = json.loads(_run(None, '--arch', 'M', '--z-dim', '40', '--n-z', '100', '--batch', '64')[0])
    assert j['metric'] == 'importance-sampled NLL images/sec (K=100, MNIST-shape)'
    assert 'configs[0]' in j['config']['workload']
