#!/usr/bin/env python
"""bench.py — importance-sampled NLL evaluation throughput (images/s, K=1000, CIFAR shape).

One "step" = one Evaluator mini-batch of the test graph (dmm_cifar10.py:880-927): q_net once per
image, K reparameterised draws, G_theta + D_psi once per (image, sample), the likelihood and the
per-image log-mean-exp, for `--batch` images per GPU.  Workload = BASELINE.json configs[1]:
arch A (dmm_cifar10.py ResNet VAE + energy prior), 32x32x3, z_dim=128, K=1000, synthetic uint8
images, random-init weights (glorot-normal kernels, N(0,0.1) biases).

  python bench.py --gpus N --steps K --warmup W          our arm (under torchrun for N>1)
  python bench.py --impl reference ...                   the CPU restatement of the reference path

Prints ONE JSON line (see DESIGN.md "Measurement" for every key).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, 'adversarial-prior-vae_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = 'importance-sampled NLL images/sec (K=1000, CIFAR-shape)'
UNIT = 'images/s'
FALLBACK_PEAKS = {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=8)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--arch', default='A')
    ap.add_argument('--dataset', default=None)
    ap.add_argument('--z-dim', type=int, default=128)
    ap.add_argument('--n-z', type=int, default=1000, help='importance samples per image (K)')
    ap.add_argument('--batch', type=int, default=37,
                    help='images per step per GPU (37 x 1000 samples = 125 x 296: every chunk is a whole number of 148-SM waves)')
    ap.add_argument('--chunk', type=int, default=7400, help='(image,sample) pairs per decoder/energy launch')
    ap.add_argument('--cpu-batch', type=int, default=2)
    ap.add_argument('--cpu-n-z', type=int, default=8)
    ap.add_argument('--cpu-steps', type=int, default=2, help='timed runs of the CPU baseline sample (after --cpu-warmup untimed ones)')
    ap.add_argument('--cpu-warmup', type=int, default=1)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-op-times', action='store_true')
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        d['_source'] = 'measured'
        return d
    d = dict(FALLBACK_PEAKS)
    d['_source'] = 'fallback'
    return d


def synthetic_weights(schema, seed=1234):
    """Random-init weights of the architecture (no checkpoint offline): glorot-normal kernels —
    including the heads the reference zero-initialises, which would make the run degenerate —
    and N(0, 0.1^2) biases."""
    rng = np.random.RandomState(seed)
    out = {}
    for name, shp in schema:
        if name.endswith('/kernel'):
            rf = shp[0] * shp[1] if len(shp) == 4 else 1
            fan_in, fan_out = (shp[2] * rf, shp[3] * rf) if len(shp) == 4 else shp
            out[name] = rng.normal(0.0, np.sqrt(2.0 / (fan_in + fan_out)), size=shp).astype(np.float32)
        else:
            out[name] = rng.normal(0.0, 0.1, size=shp).astype(np.float32)
    return out


def synthetic_images(n, x_shape, seed=0):
    """uint8 ~ U{0..255} mapped like dmm_cifar10.py:1084-1085."""
    u8 = np.random.RandomState(seed).randint(0, 256, size=(n,) + tuple(x_shape))
    return ((u8 - 127.5) / 256.0 * 2).astype(np.float32)


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md).  The nvidia-smi process is
    started BEFORE the warm-up (its start-up enumerates every GPU of the box under driver-wide locks, which stalls
    kernel launches for tens of ms on an 8-GPU node); only rows that arrive inside the timed window are kept."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(',')]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for t, r in self.rows:
            if (t0 is not None and t < t0) or (t1 is not None and t > t1):
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[4:8]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def cpu_reference_throughput(args, steps, warmup):
    """The reference's path restated on the CPU (oracle/, float32 = the reference dtype, G_theta
    evaluated once per sample), timed on a bounded sample and scaled linearly in K to K=n_z
    (samples/s is K-independent).  TensorFlow 1.x itself is unavailable on this box."""
    import oracle as O
    kw = {'z_dim': args.z_dim}
    if args.arch == 'B':
        kw['dataset'] = args.dataset or 'cifar10'
    cfg = O.OracleConfig(args.arch, **kw)
    params = O.init_weights(cfg, dtype=np.float32)
    B, K = args.cpu_batch, args.cpu_n_z
    x = synthetic_images(B, cfg.x_shape).astype(np.float32)
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count()
    # all the host threads the BLAS can use, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1);
    # `cores` reports the thread count actually in effect
    limit = None
    try:
        from threadpoolctl import threadpool_limits, threadpool_info
        limit = threadpool_limits(limits=cores, user_api='blas')
        got = [p.get('num_threads') for p in threadpool_info() if p.get('user_api') == 'blas']
        if got:
            cores = int(max(got))
    except Exception:
        pass
    times = []
    for i in range(warmup + steps):
        eps = np.random.RandomState(100 + i).normal(size=(K, B, cfg.z_dim)).astype(np.float32)
        t0 = time.perf_counter()
        O.eval_batch(cfg, params, x, x, eps)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    if limit is not None:
        limit.restore_original_limits()
    t = float(np.mean(times))
    samples_per_s = B * K / t
    return {'value': samples_per_s / args.n_z, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': 'oracle eval_batch (numpy fp32, BLAS threads) on B=%d images x K=%d samples per step, '
                      '%.2f s/step, %.2f (image,sample)/s; images/s scaled linearly to K=%d'
                      % (B, K, t, samples_per_s, args.n_z),
            'ms_per_step': t * 1e3}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    steps, warmup = min(args.steps, 3), min(args.warmup, 1)
    r = cpu_reference_throughput(args, steps, warmup)
    line = {
        'impl': 'reference', 'metric': metric_name(args), 'value': r['value'], 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': steps, 'warmup': warmup, 'ms_per_step': r['ms_per_step'], 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(args, None),
        'cpu_baseline': {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'note': 'TensorFlow 1.x / ZhuSuan (the reference runtime) unavailable on this box: CPU restatement timed',
    }
    print(json.dumps(line))


def metric_name(args):
    """BASELINE.json's metric for the default workload; other shapes / K (the parity configs run as benches) say so."""
    ds = args.dataset or {'A': 'cifar10', 'B': 'cifar10', 'M': 'fashion'}[args.arch]
    if args.n_z == 1000 and ds == 'cifar10':
        return METRIC
    shape = {'cifar10': 'CIFAR', 'celebA': 'CelebA', 'fashion': 'MNIST'}.get(ds, ds)
    return 'importance-sampled NLL images/sec (K=%d, %s-shape)' % (args.n_z, shape)


def workload_config(args, world):
    ds = args.dataset or {'A': 'cifar10', 'B': 'cifar10', 'M': 'fashion'}[args.arch]
    which = 'configs[1]' if (args.arch == 'A' and args.n_z == 1000) else \
        'configs[0]' if (args.arch == 'M' and args.n_z == 100 and args.batch == 64) else \
        'configs[4] (z_dim sweep)' if (args.arch == 'B' and ds == 'cifar10') else 'custom'
    return {'workload': 'BASELINE %s: arch %s (%s-shape) ResNet VAE + energy prior, z_dim=%d, K=%d '
                        'importance-sampled NLL' % (which, args.arch, ds, args.z_dim, args.n_z),
            'arch': args.arch, 'z_dim': args.z_dim, 'n_z': args.n_z, 'images_per_step_per_gpu': args.batch,
            'chunk': args.chunk, 'parallelism': ('dp%d (image shards, one allreduce of metric sums)' % world) if world else 'cpu',
            'l2': 'per-step activation working set (>1 GB) exceeds the 126 MB L2; a 256 MB buffer is also '
                  'rewritten between timed steps'}


def main():
    args = parse_args()
    if args.impl == 'reference':
        return run_reference(args)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a B200: there is no CPU fallback (use --impl reference for the CPU baseline)')
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    from apvae import _lib
    from apvae.model import ExpConfig, Model, weight_schema
    from apvae.evaluation import Evaluator, combine_accumulators, metrics_from_accumulator

    kw = {'z_dim': args.z_dim}
    cfg = ExpConfig(args.arch, dataset=args.dataset or ('cifar10' if args.arch != 'M' else None),
                    max_batch=max(args.batch, 1), max_chunk=args.chunk, **kw)
    model = Model(cfg, synthetic_weights(weight_schema(cfg)), device=local)
    B, K, W, S = args.batch, args.n_z, args.warmup, args.steps
    ev = Evaluator(model, n_z=K, batch_size=B, seed=1 + 1000 * rank)

    # this rank's images for every step, resident in HBM before the timed region
    n_img = (W + S) * B
    x_host = torch.from_numpy(synthetic_images(n_img, cfg.x_shape, seed=rank)).pin_memory()
    x_dev = x_host.to(model.device)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=model.device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident throughput ("value")
    accum = torch.zeros(_lib.ACCUM_SLOTS, dtype=torch.float64, device=model.device)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(W):
        wa = ev.run_shard(x_dev[i * B:(i + 1) * B], accum=accum.clone())
        combine_accumulators(wa)                                 # warm-up of the collective path on a FILLED accumulator
                                                                 # (NCCL lazy init + first use of the torch kernels it runs)
    barrier()
    launches0 = model.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_region0 = time.perf_counter()
    e0.record()
    for i in range(W, W + S):
        flush.fill_(i & 0xFF)                                    # evict L2 between timed steps
        ev.run_shard(x_dev[i * B:(i + 1) * B], accum=accum, first_index=i * B)
    total = combine_accumulators(accum)                          # the one allreduce of metric sums
    e1.record()
    barrier()
    t_region1 = time.perf_counter()
    if os.environ.get('APVAE_BENCH_DEBUG'):
        sys.stderr.write('rank %d value-loop ms %.2f\n' % (rank, e0.elapsed_time(e1)))
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=model.device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    clocks = sampler.stop(t_region0, t_region1) if rank == 0 else None
    launches = model.launch_count() - launches0
    metrics = metrics_from_accumulator(total)
    value = S * B * world / (ms / 1e3)

    # ---------------- end to end through the public API with HOST buffers ("e2e")
    out_host = torch.empty((_lib.PER_IMAGE_ROWS, B), dtype=torch.float32).pin_memory()
    for i in range(min(W, 2)):
        r = model.is_nll(x_host[i * B:(i + 1) * B].to(model.device, non_blocking=True), K, seed=5 + i)
        out_host.copy_(r['per_image'])
    barrier()
    t0 = time.perf_counter()
    for i in range(W, W + S):
        xb = x_host[i * B:(i + 1) * B].to(model.device, non_blocking=True)      # H2D of the step's inputs
        r = model.is_nll(xb, K, seed=77 + i)
        out_host.copy_(r['per_image'])                                           # D2H of the step's result
        torch.cuda.synchronize()
    if os.environ.get('APVAE_BENCH_DEBUG'):
        sys.stderr.write('rank %d e2e-loop ms %.2f (before the closing barrier)\n' % (rank, 1e3 * (time.perf_counter() - t0)))
        for rep in range(2):                                 # the value loop again, without flush / collective
            barrier()
            e0.record()
            for i in range(W, W + S):
                ev.run_shard(x_dev[i * B:(i + 1) * B], accum=accum, first_index=i * B)
            e1.record()
            torch.cuda.synchronize()
            sys.stderr.write('rank %d value-loop again (no flush, no collective) ms %.2f\n' % (rank, e0.elapsed_time(e1)))
        for rep in range(2):
            barrier()
            e0.record()
            for i in range(W, W + S):
                flush.fill_(i & 0xFF)
                ev.run_shard(x_dev[i * B:(i + 1) * B], accum=accum, first_index=i * B)
            e1.record()
            torch.cuda.synchronize()
            sys.stderr.write('rank %d value-loop again (flush, no collective) ms %.2f\n' % (rank, e0.elapsed_time(e1)))
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=model.device)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = S * B * world / float(t_e2e)
    h2d = B * int(np.prod(cfg.x_shape)) * 4
    d2h = _lib.PER_IMAGE_ROWS * B * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel (tap-GEMM) + per-step tensor fraction
    peaks = load_peaks()
    mq, mg, md = model.macs()
    flops_per_step = 2.0 * (mq * B + (mg + md) * K * B + md * B)
    roofline, op_table = None, None
    if not args.no_op_times:
        n_units = min(args.chunk, K * B)
        rows = model.op_times(n_units, reps=5)
        tot = sum(r[3] * (1 if r[0] == 'q_net' else 1) for r in rows if r[0] != 'q_net')
        op_table = []
        for net, name, macs, t_ms, nn in rows:
            tf = 2.0 * macs * nn / (t_ms * 1e-3) / 1e12 if t_ms > 0 else 0.0
            row = {'op': name, 'units': nn, 'ms': round(t_ms, 4), 'tflops': round(tf, 1),
                   'share': round(t_ms / tot, 4) if net != 'q_net' else None}
            # the HBM-bound CUDA-core ops: algorithmic bytes per unit -> achieved GB/s against the copy peak
            H, Wd, Cx = cfg.x_shape
            by = None
            if '[heads' in name and args.arch == 'A':
                by = H * Wd * (96 * 2 + 2 * Cx * 4 + 2 * 16 * 2)           # trunk row in, fp32 heads + D_psi's raw / act rows out
            elif '[fused block' in name:
                by = H * Wd * (16 * 2) * 4                                # x in, r, raw and act out: 16-channel bf16 rows
            if by and t_ms > 0:
                row['hbm_gbs'] = round(by * nn / (t_ms * 1e-3) / 1e9, 1)
                row['hbm_frac'] = round(row['hbm_gbs'] / peaks['hbm_gbs'], 4)
            op_table.append(row)
        dom = max((r for r in rows if r[0] != 'q_net'), key=lambda r: r[3])
        ach = 2.0 * dom[2] * dom[4] / (dom[3] * 1e-3) / 1e12
        all_flops = sum(2.0 * r[2] * r[4] for r in rows if r[0] != 'q_net')
        traffic, traffic_src = None, None
        try:                                   # DRAM bytes per launch from the committed ncu --set full capture
            with open(os.path.join(ROOT, 'profiles', 'r1_traffic.json')) as f:
                tj = json.load(f)
            if args.chunk == tj.get('chunk') and args.arch == 'A' and args.z_dim == 128:
                traffic = tj['bytes_per_launch'].get(dom[1].split(' [')[0])
                traffic_src = tj['source']
        except (OSError, ValueError, KeyError):
            pass
        roofline = {'bound': 'tensor', 'kernel': 'tcgen05/TMEM/TMA implicit GEMM, op ' + dom[1],
                    'traffic_unit': 'bytes of DRAM per launch', 'traffic_source': traffic_src,
                    'achieved': round(ach, 1), 'peak': peaks['bf16_tflops'], 'unit': 'TFLOP/s',
                    'frac': round(ach / peaks['bf16_tflops'], 4), 'traffic': traffic,
                    'peak_source': peaks['_source'] + ' burst (kernel timed alone, CUDA events)',
                    'share_of_chunk_time': round(dom[3] / tot, 4),
                    'all_ops_tflops': round(all_flops / (tot * 1e-3) / 1e12, 1),
                    'all_ops_frac': round(all_flops / (tot * 1e-3) / 1e12 / peaks['bf16_tflops'], 4)}
    step_tflops = flops_per_step * S / (ms / 1e3) / 1e12
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_throughput(args, steps=max(1, args.cpu_steps), warmup=max(0, args.cpu_warmup))
        cpu.pop('ms_per_step')

    line = {
        'metric': metric_name(args), 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': S, 'warmup': W,
        'ms_per_step': ms / S, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'bf16', 'data': 'synthetic', 'config': workload_config(args, world),
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h},
        'gpu_launches': int(launches),
        'roofline': roofline,
        'step_tensor': {'algorithmic_tflops_per_gpu': round(step_tflops, 1),      # flops_per_step is per rank
                        'frac_of_sustained_peak': round(step_tflops / peaks['bf16_tflops_sustained'], 4),
                        'gflop_per_image': round(flops_per_step / B / 1e9, 1)},
        'cpu_baseline': cpu,
        'nll': {k: metrics.get(k) for k in ('test_nll', 'adv_test_nll')},
        'ops': op_table,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
