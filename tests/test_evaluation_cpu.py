"""Host logic of the sharded Evaluator (no GPU): shard slicing, the metric-sum allreduce and
the log-Z (max, sum-exp) combine across ranks, over gloo with world_size 2; and that the C-ABI
library loads and exports every symbol include/apvae.h declares."""
import os
import re
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from apvae import _lib
    with open(os.path.join(ROOT, 'include', 'apvae.h')) as f:
        src = re.sub(r'/\*.*?\*/', '', f.read(), flags=re.S)
        declared = set(re.findall(r'\b(apvae_[a-z0-9_]+)\s*\(', src))
    assert declared, 'no declarations parsed'
    for name in sorted(declared):
        assert hasattr(_lib.lib, name), 'library does not export %s' % name
    assert declared == set(_lib.SYMBOL_NAMES), declared ^ set(_lib.SYMBOL_NAMES)


def test_no_gpu_means_loud_failure_not_fallback():
    """Without a GPU every compute entry point must fail (APVAE_ERR_NO_DEVICE): no CPU fallback."""
    if torch.cuda.is_available():
        return
    import ctypes as C
    from apvae import _lib, ApvaeError
    from apvae.model import ExpConfig
    h = C.c_void_p()
    c = ExpConfig('B', dataset='cifar10').to_c()
    st = _lib.lib.apvae_create(C.byref(c), 0, C.byref(h))
    assert st == _lib.ERR_NO_DEVICE
    assert 'no CPU fallback' in _lib.last_error(None)
    try:
        _lib.check(st)
    except ApvaeError as e:
        assert e.status == _lib.ERR_NO_DEVICE
    else:
        raise AssertionError('check() did not raise')


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'adversarial-prior-vae_b200')
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith(('.py', '.cu', '.cpp', '.h')):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r'^\s*(import|from)\s+oracle\b', src, re.M), fn


def test_shard_range_covers_everything_once():
    from apvae.evaluation import shard_range
    for n in (0, 1, 7, 8, 9, 10000):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            covered = [i for lo, hi in spans for i in range(lo, hi)]
            assert covered == list(range(n))
            assert max(hi - lo for lo, hi in spans) == -(-n // world) or n == 0


def test_metrics_from_accumulator_matches_evaluator_average():
    """Summed accumulators == Evaluator's np.average of per-batch means with batch-size weights
    (tfsnippet/trainer/evaluator.py:249-254)."""
    import oracle as O
    from apvae import _lib
    from apvae.evaluation import metrics_from_accumulator
    rs = np.random.RandomState(0)
    sizes = [64, 64, 17]
    per_image = [rs.normal(size=(9, b)) for b in sizes]
    acc = np.zeros(_lib.ACCUM_SLOTS)
    for p in per_image:
        acc[:9] += p.sum(axis=1)
        acc[_lib.ACCUM_WEIGHT] += p.shape[1]
    acc[_lib.ACCUM_BATCHES] = 0
    got = metrics_from_accumulator(acc)
    want = O.evaluator_average([-p[_lib.ROW_ADV_LL].mean() for p in per_image], sizes)
    assert abs(got['adv_test_nll'] - want) < 1e-12
    want = O.evaluator_average([p[_lib.ROW_REAL_E].mean() for p in per_image], sizes)
    assert abs(got['real_energy'] - want) < 1e-12


def _worker(rank, world, port, q):
    sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from apvae import _lib
    from apvae.evaluation import combine_accumulators, metrics_from_accumulator
    rs = np.random.RandomState(rank)
    acc = torch.zeros(_lib.ACCUM_SLOTS, dtype=torch.float64)
    acc[:9] = torch.tensor(rs.normal(size=9) * 100)
    acc[_lib.ACCUM_WEIGHT] = 10 + rank
    # per-rank list of batch log-Z values lz_b, stored as (max, sum exp(lz - max), count)
    lz = rs.normal(size=3 + rank) * 50 + (400 if rank else -300)     # far apart: exercises the rescale
    acc[_lib.ROW_LZ_MAX] = float(lz.max())
    acc[_lib.ROW_LZ_SUM] = float(np.exp(lz - lz.max()).sum())
    acc[_lib.ACCUM_BATCHES] = len(lz)
    out = combine_accumulators(acc)
    q.put((rank, acc.numpy().copy(), lz, out.numpy().copy(), metrics_from_accumulator(out, 50000)))
    dist.destroy_process_group()


def test_two_rank_combine_over_gloo():
    import oracle as O
    from apvae import _lib
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, a0, lz0, o0, m0), (_, a1, lz1, o1, m1) = res
    np.testing.assert_allclose(o0, o1)                               # every rank holds the same result
    for s in (0, 1, 2, 3, 4, 5, 8, _lib.ACCUM_WEIGHT, _lib.ACCUM_BATCHES):
        assert abs(o0[s] - (a0[s] + a1[s])) < 1e-9
    want_log_Z = O.combine_log_Z(np.concatenate([lz0, lz1])) + np.log(50000)   # dmm_cifar10.py:1221-1222 (+ :922-924)
    assert abs(m0['log_Z'] - want_log_Z) < 1e-9
    assert abs(m0['test_nll'] - (-(a0[0] + a1[0]) / 21.0)) < 1e-9
