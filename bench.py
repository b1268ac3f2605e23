#!/usr/bin/env python
"""bench.py — importance-sampled NLL evaluation throughput (images/s, K=1000, CIFAR shape).

One "step" = one Evaluator mini-batch of the test graph (dmm_cifar10.py:880-927): q_net once per
image, K reparameterised draws, G_theta + D_psi once per (image, sample), the likelihood, the
per-image log-mean-exp, test_mse, and the test_pn_net prior branch (config.test_n_pz = 1000 prior
draws through G_theta and D_psi per mini-batch, :883,915-927), for `--batch` images per GPU.
Workload = BASELINE.json configs[1]: arch A (dmm_cifar10.py ResNet VAE + energy prior), 32x32x3,
z_dim=128, K=1000, synthetic uint8 images, random-init weights (glorot-normal kernels, N(0,0.1)
biases).

  python bench.py --gpus N --steps K --warmup W          our arm (under torchrun for N>1; weak scaling)
  python bench.py --images 10000 [--gpus N ...]          BASELINE configs[2]: a FIXED image set through Evaluator.run,
                                                         ragged contiguous shards over the ranks (strong scaling)
  python bench.py --impl reference ...                   the CPU restatement of the reference path (torch-CPU convs)

Prints ONE JSON line (see DESIGN.md "Measurement" for every key).  At N = 1 the line also carries
`mala` (BASELINE configs[3]: CelebA-shape latent MALA sampling) and `aux_kernels` (achieved HBM GB/s
of the CUDA-core kernels).
"""
import argparse
import ctypes as C
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


def parse_args(argv=None):
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
    ap.add_argument('--images', type=int, default=0,
                    help='strong scaling (BASELINE configs[2]): evaluate this many images in total, sharded over the ranks')
    ap.add_argument('--operand-dtype', default='fp16', choices=['fp16', 'bf16'])
    ap.add_argument('--tuning', default='', help='apvae_tuning overrides for A/B runs, e.g. patch_2cta_min_k=1500 (default: none)')
    ap.add_argument('--no-prior-branch', action='store_true',
                    help='leave out the test_pn_net branch (pd_energy / pn_energy / kl_adv_and_gaussian); the line says so')
    ap.add_argument('--cpu-batch', type=int, default=1)
    ap.add_argument('--cpu-n-z', type=int, default=64)
    ap.add_argument('--cpu-steps', type=int, default=2, help='timed runs of the CPU baseline sample (after --cpu-warmup untimed ones)')
    ap.add_argument('--cpu-warmup', type=int, default=1)
    ap.add_argument('--cpu-backend', default='torch', choices=['torch', 'numpy'],
                    help="conv backend of the CPU restatement: torch-CPU (oneDNN; default) or plain numpy")
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-op-times', action='store_true')
    ap.add_argument('--no-mala', action='store_true')
    ap.add_argument('--mala-chains', type=int, default=1024)
    ap.add_argument('--mala-steps', type=int, default=20)
    return ap.parse_args(argv)


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


# ----------------------------------------------------------------------------------------- CPU arm
def _host_threads():
    """All the host threads the CPU restatement may use, whatever OMP_NUM_THREADS says (torchrun exports 1)."""
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    return cores


class _CpuThreads(object):
    """Context: BLAS (numpy) and torch intra-op pools opened to `cores` threads; reports the count in effect."""

    def __init__(self, backend):
        self.backend, self.cores, self._limit, self._old_backend, self._old_torch = backend, _host_threads(), None, None, None

    def __enter__(self):
        import oracle as O
        try:
            from threadpoolctl import threadpool_limits
            self._limit = threadpool_limits(limits=self.cores, user_api='blas')
        except Exception:
            pass
        if self.backend == 'torch':
            import torch
            self._old_torch = torch.get_num_threads()
            torch.set_num_threads(self.cores)
            self.cores = torch.get_num_threads()
        self._old_backend = O.set_conv_backend(self.backend)
        return self

    def __exit__(self, *exc):
        import oracle as O
        O.set_conv_backend(self._old_backend)
        if self._old_torch is not None:
            import torch
            torch.set_num_threads(self._old_torch)
        if self._limit is not None:
            self._limit.restore_original_limits()


def cpu_reference_throughput(args, steps, warmup):
    """The reference's path restated on the CPU (oracle/, float32 = the reference dtype, G_theta evaluated once per
    sample, convolutions through torch-CPU / oneDNN — the closest available stand-in for TF-CPU's Eigen / MKL-DNN
    kernels), timed on a bounded sample: `--cpu-batch` images x `--cpu-n-z` samples per step, prior branch included in
    proportion.  `measured` is the sample's own (image,sample) rate; `value` scales it linearly in K to K = n_z
    (samples/s is K-independent).  TensorFlow 1.x itself is unavailable on this box."""
    import oracle as O
    kw = {'z_dim': args.z_dim}
    if args.arch == 'B':
        kw['dataset'] = args.dataset or 'cifar10'
    cfg = O.OracleConfig(args.arch, **kw)
    params = O.init_weights(cfg, dtype=np.float32)
    B, K = args.cpu_batch, args.cpu_n_z
    x = synthetic_images(B, cfg.x_shape).astype(np.float32)
    # the prior branch costs test_n_pz / (batch * n_z) of a GPU step; the CPU sample carries the same proportion
    Kp = 0 if args.no_prior_branch else max(1, int(round(1000.0 * B * K / (args.batch * args.n_z))))
    times = []
    with _CpuThreads(args.cpu_backend) as th:
        for i in range(warmup + steps):
            rs = np.random.RandomState(100 + i)
            eps = rs.normal(size=(K, B, cfg.z_dim)).astype(np.float32)
            eps_p = rs.normal(size=(Kp, 1, cfg.z_dim)).astype(np.float32)
            t0 = time.perf_counter()
            O.eval_batch(cfg, params, x, x, eps)
            if Kp:
                O.prior_branch(cfg, params, eps_p)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        cores = th.cores
    t = float(np.mean(times))
    samples_per_s = B * K / t
    return {'value': samples_per_s / args.n_z, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'backend': args.cpu_backend,
            'measured': {'images': B, 'n_z': K, 'prior_draws': Kp, 's_per_step': round(t, 3),
                         'image_samples_per_s': round(samples_per_s, 2),
                         'images_per_s_at_sample_n_z': round(B / t, 4)},
            'sample': 'oracle eval_batch (float32, %s convs, %d threads) on B=%d images x K=%d samples (+%d prior draws) per '
                      'step, %.2f s/step = %.2f (image,sample)/s measured; images/s scaled linearly to K=%d (extrapolated)'
                      % (args.cpu_backend, cores, B, K, Kp, t, samples_per_s, args.n_z),
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
        'cpu_baseline': {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'backend', 'measured', 'sample')},
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'note': 'TensorFlow 1.x / ZhuSuan (the reference runtime) unavailable on this box: CPU restatement timed '
                '(rank 0 only; it does not scale with --gpus)',
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
    which = ('configs[2] (%d images, strong scaling)' % args.images) if (args.images and args.arch == 'A' and args.n_z == 1000) else \
        'configs[1]' if (args.arch == 'A' and args.n_z == 1000) else \
        'configs[0]' if (args.arch == 'M' and args.n_z == 100 and args.batch == 64) else \
        'configs[4] (z_dim sweep)' if (args.arch == 'B' and ds == 'cifar10') else 'custom'
    return {'workload': 'BASELINE %s: arch %s (%s-shape) ResNet VAE + energy prior, z_dim=%d, K=%d '
                        'importance-sampled NLL' % (which, args.arch, ds, args.z_dim, args.n_z),
            'arch': args.arch, 'z_dim': args.z_dim, 'n_z': args.n_z, 'images_per_step_per_gpu': args.batch,
            'scalars': 'p_net/xi = -6 (pre-sigmoid), beta = 0, log_Z = (z_dim / 2) log 2 pi',
            'chunk': args.chunk,
            'operands': '%s operands, fp32 accumulation (tcgen05 kind::f16)' % args.operand_dtype,
            'prior_branch': (not args.no_prior_branch),
            'parallelism': ('dp%d (image shards, one all-gather of the metric accumulators)' % world) if world else 'cpu',
            'l2': 'per-step activation working set (>1 GB) exceeds the 126 MB L2; a 256 MB buffer is also '
                  'rewritten between timed steps'}


# ----------------------------------------------------------------------------------------- sampler (configs[3])
def mala_bench(args, local, peaks):
    """BASELINE configs[3]: CelebA-shape (64x64x3, arch B = dmm_for_sampling.py) latent MALA sampling of the energy prior
    at the reference's settings (pull_back_energy_weight = 1000, alpha = 5e-5; dmm_cifar10.py:179-252): `--mala-chains`
    chains x `--mala-steps` steps, in-kernel Philox noise, then decode to uint8.  Three modes of the same code, each with
    its mean acceptance ratio: split precision (precise=1: activations and kernels as hi + lo pairs, three tensor-core
    terms, fp32-class energies), activation split (precise=2: activations hi + lo, kernels single, two terms — removes the
    per-element activation rounding that differs between z and z'; the kernels' rounding is a fixed smooth perturbation
    that cancels in E(z) - E(z')) and the single-pass 16-bit forward."""
    import torch
    import oracle as O
    from apvae.model import ExpConfig, Model, weight_schema
    n, steps = args.mala_chains, args.mala_steps
    out = {'config': {'workload': 'BASELINE configs[3]: arch B CelebA-shape 64x64x3, z_dim=128, latent MALA on the energy prior',
                      'chains': n, 'steps': steps, 'alpha': 5e-5, 'pull_back_energy_weight': 1000.0,
                      'noise': 'in-kernel Philox', 'flops': '4 (M_G + M_D) per chain-step (1 forward + 1 input-gradient pass)'},
           'unit': 'chain-steps/s'}
    z0_host = torch.from_numpy(np.random.RandomState(2).normal(size=(n, 128)).astype(np.float32)).pin_memory()
    for mode, precise in (('split_precision', 1), ('activation_split', 2), ('single_pass', 0)):
        cfg = ExpConfig('B', dataset='celebA', z_dim=128, max_batch=8, max_chunk=min(n, 1024), enable_grad=True,
                        precise=precise, operand_dtype=args.operand_dtype)
        m = Model(cfg, synthetic_weights(weight_schema(cfg)), device=local)
        z0 = z0_host.to(m.device)
        m.mala(z0, 2, seed=1, history=False)                      # warm-up (first-launch costs, state buffers)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = m.launch_count()
        e0.record()
        r = m.mala(z0, steps, seed=7)
        e1.record()
        torch.cuda.synchronize()
        sec = e0.elapsed_time(e1) * 1e-3
        launches = m.launch_count() - l0
        # end to end: z0 from pinned host memory in, uint8 images out to the host, inside the timed region
        u8_host = torch.empty((n, 64, 64, 3), dtype=torch.uint8).pin_memory()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        zz = z0_host.to(m.device, non_blocking=True)
        rr = m.mala(zz, steps, seed=8, history=False)
        u8_host.copy_(m.sample_images(rr['z']))
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
        mq, mg, md = m.macs()
        tflops = 4.0 * (mg + md) * n * steps / sec / 1e12
        out[mode] = {'value': n * steps / sec, 'ms_per_step': sec / steps * 1e3,
                     'accept_mean': float(r['history_ratio'].mean()),
                     'algorithmic_tflops': round(tflops, 1),
                     'frac_of_sustained_peak': round(tflops / peaks['bf16_tflops_sustained'], 4),
                     'gpu_launches': int(launches),
                     'e2e': {'value': n * steps / t_e2e, 'unit': 'chain-steps/s', 'h2d_bytes': n * 128 * 4,
                             'd2h_bytes': n * 64 * 64 * 3}}
        m.close()
        del m
        torch.cuda.empty_cache()
    # the headline: the fastest mode whose mean acceptance is within 0.02 of the full split-precision (fp32-class) chain's
    ref_acc = out['split_precision']['accept_mean']
    ok = [k for k in ('split_precision', 'activation_split', 'single_pass') if abs(out[k]['accept_mean'] - ref_acc) <= 0.02]
    out['headline_mode'] = max(ok, key=lambda k: out[k]['value'])
    out['value'] = out[out['headline_mode']]['value']
    # CPU baseline: the oracle's mala_step (float32, torch-CPU convs) on a bounded sample of the same chains
    if not args.no_cpu_baseline:
        ocfg = O.OracleConfig('B', dataset='celebA', z_dim=128)
        params = O.init_weights(ocfg, dtype=np.float32)
        nc = 4
        rs = np.random.RandomState(5)
        z = z0_host.numpy()[:nc, None, :].copy()
        with _CpuThreads(args.cpu_backend) as th:
            acc, times = [], []
            for i in range(3):
                eps = rs.normal(size=z.shape).astype(np.float32)
                u = rs.uniform(size=(nc, 1)).astype(np.float32)
                t0 = time.perf_counter()
                z, _, _, ratio = O.mala_step(ocfg, params, z, eps, u, 5e-5)
                if i:
                    times.append(time.perf_counter() - t0)
                    acc.append(float(np.mean(ratio)))
            cores = th.cores
        t = float(np.mean(times))
        out['cpu_baseline'] = {'value': nc / t, 'unit': 'chain-steps/s', 'cores': cores, 'kind': 'port', 'backend': args.cpu_backend,
                               'accept_mean': float(np.mean(acc)),
                               'sample': 'oracle mala_step (float32, 2 forward + 2 input-gradient passes of D o G per step as in the '
                                         'reference graph) on %d chains, %.2f s/step' % (nc, t)}
    return out


def aux_kernel_table(model, n, B, peaks, sampler):
    """Achieved HBM GB/s of the CUDA-core kernels of the estimator tail (and of the sampler when `sampler`): algorithmic
    bytes per launch (apvae_aux_time) / CUDA-event time per launch, against the measured copy peak."""
    from apvae import _lib
    from apvae._lib import lib, check
    rows = []
    name = C.create_string_buffer(64)
    by, ms = C.c_double(), C.c_float()
    ids = list(range(_lib.AUX_COUNT))
    for which in ids:
        is_sampler = which in (6, 7, 8, 9)
        if is_sampler and not sampler:
            continue
        nn = n if which != 5 else n * 8
        try:
            check(lib.apvae_aux_time(model._h, which, int(nn), int(B), 20, name, 64, C.byref(by), C.byref(ms), None), model._h)
        except Exception as e:                                   # a kernel this handle cannot run (e.g. no fused heads)
            rows.append({'kernel': 'aux %d' % which, 'error': str(e)[:80]})
            continue
        gbs = by.value / (ms.value * 1e-3) / 1e9 if ms.value > 0 else 0.0
        rows.append({'kernel': name.value.decode(), 'units': int(nn), 'ms': round(ms.value, 5),
                     'bytes_per_launch': int(by.value), 'hbm_gbs': round(gbs, 1), 'hbm_frac': round(gbs / peaks['hbm_gbs'], 4)})
    return rows


def _finite(v):
    """JSON has no NaN / Infinity: non-finite metric values (e.g. pd_energy's exp overflowing in float32 with random-init
    weights, exactly as the reference's graph would) are reported as strings."""
    if isinstance(v, (float, np.floating)):
        v = float(v)
        return v if np.isfinite(v) else repr(v)                  # 'nan' / 'inf' / '-inf'
    return v


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
    from apvae.evaluation import Evaluator, combine_accumulators, metrics_from_accumulator, shard_range

    kw = {'z_dim': args.z_dim}
    cfg = ExpConfig(args.arch, dataset=args.dataset or ('cifar10' if args.arch != 'M' else None),
                    max_batch=max(args.batch, 1), max_chunk=args.chunk, operand_dtype=args.operand_dtype,
                    tuning={k.strip(): int(v) for k, v in (kv.split('=') for kv in args.tuning.split(',') if kv.strip())}, **kw)
    model = Model(cfg, synthetic_weights(weight_schema(cfg)), device=local)
    # scalars of a trained-like model so that every printed metric is finite in float32 (pd_energy averages
    # D exp(-w xi D - log_Z + (d/2) log 2 pi): with the initial xi = 0.5 and log_Z = 1 the exponent overflows on random
    # weights, exactly as the reference's graph would): xi = sigmoid(-6), log_Z = (d/2) log 2 pi.  No effect on the timing.
    scalars = {'xi_raw': -6.0, 'beta': 0.0, 'log_Z': round(0.5 * cfg.z_dim * float(np.log(2 * np.pi)), 3)}
    model.set_scalars(scalars['xi_raw'], scalars['beta'], scalars['log_Z'])
    B, K, W, S = args.batch, args.n_z, args.warmup, args.steps
    prior = not args.no_prior_branch
    Kp = cfg.test_n_pz if prior else 0
    ev = Evaluator(model, n_z=K, batch_size=B, seed=1 + 1000 * rank, prior_branch=prior)
    strong = args.images > 0

    # this rank's images, resident in HBM before the timed region
    if strong:
        lo, hi = shard_range(args.images, rank, world)
        n_warm = min(W * B, hi - lo)
        # synthetic_images is seeded: every rank generates the same set and keeps its contiguous shard (ceil(N / R) rule)
        x_all = synthetic_images(args.images, cfg.x_shape, seed=0)
        x_host = torch.from_numpy(x_all[lo:hi]).pin_memory()
        n_img = hi - lo
    else:
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
        xs = x_dev[:n_warm][(i * B) % max(1, n_warm):][:B] if strong else x_dev[i * B:(i + 1) * B]
        wa = ev.run_shard(xs, accum=accum.clone())
        combine_accumulators(wa)                                 # warm-up of the collective path on a FILLED accumulator
                                                                 # (NCCL lazy init + first use of the torch kernels it runs)
    barrier()
    launches0 = model.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_region0 = time.perf_counter()
    e0.record()
    if strong:
        # one Evaluator pass over the shard, batch by batch (the last batch ragged), L2 flushed between batches
        for b0 in range(0, n_img, B):
            flush.fill_(b0 & 0xFF)
            ev.run_shard(x_dev[b0:b0 + B], accum=accum, first_index=lo + b0)
        steps_done = -(-n_img // B)
    else:
        for i in range(W, W + S):
            flush.fill_(i & 0xFF)                                # evict L2 between timed steps
            ev.run_shard(x_dev[i * B:(i + 1) * B], accum=accum, first_index=i * B)
        steps_done = S
    total = combine_accumulators(accum)                          # the one collective: all-gather of the metric accumulators
    e1.record()
    barrier()
    t_region1 = time.perf_counter()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=model.device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    clocks = sampler.stop(t_region0, t_region1) if rank == 0 else None
    launches = model.launch_count() - launches0
    metrics = metrics_from_accumulator(total)
    images_done = args.images if strong else S * B * world
    value = images_done / (ms / 1e3)
    shard_sizes = [shard_range(args.images, r, world)[1] - shard_range(args.images, r, world)[0] for r in range(world)] if strong else None

    # ---------------- end to end through the public API with HOST buffers ("e2e")
    if strong:
        # Evaluator.run on the pinned HOST array: every batch is copied H2D inside the loop (like feed_dict), the
        # metric dict comes back to the host at the end
        barrier()
        t0 = time.perf_counter()
        ev.run(torch.from_numpy(x_all).pin_memory())
        barrier()
        t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=model.device)
        if world > 1:
            dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
        e2e_value = args.images / float(t_e2e)
        h2d = B * int(np.prod(cfg.x_shape)) * 4
        d2h = _lib.ACCUM_SLOTS * 8
        e2e_steps = steps_done
    else:
        out_host = torch.empty((_lib.PER_IMAGE_ROWS, B), dtype=torch.float32).pin_memory()
        p4_host = torch.empty(4, dtype=torch.float32).pin_memory()
        for i in range(min(W, 2)):
            r = model.is_nll(x_host[i * B:(i + 1) * B].to(model.device, non_blocking=True), K, seed=5 + i)
            out_host.copy_(r['per_image'])
        barrier()
        t0 = time.perf_counter()
        for i in range(W, W + S):
            xb = x_host[i * B:(i + 1) * B].to(model.device, non_blocking=True)      # H2D of the step's inputs
            r = model.is_nll(xb, K, seed=77 + i)
            if prior:
                p4_host.copy_(model.prior_branch(Kp, seed=77 + i))
            out_host.copy_(r['per_image'])                                           # D2H of the step's result
            torch.cuda.synchronize()
        barrier()
        t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=model.device)
        if world > 1:
            dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
        e2e_value = S * B * world / float(t_e2e)
        h2d = B * int(np.prod(cfg.x_shape)) * 4
        d2h = _lib.PER_IMAGE_ROWS * B * 4 + (16 if prior else 0)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel + per-step tensor fraction
    peaks = load_peaks()
    mq, mg, md = model.macs()
    # algorithmic FLOPs of one step on one rank: q_net and D(x) per image, G and D per (image, sample) and per prior draw
    flops_per_step = 2.0 * (mq * B + (mg + md) * (K * B + Kp) + md * B)
    flops_total_rank = flops_per_step * steps_done if not strong else 2.0 * (mq * n_img + (mg + md) * (K * n_img + Kp * steps_done) + md * n_img)
    roofline, op_table, aux = None, None, None
    if not args.no_op_times:
        n_units = min(args.chunk, K * B)
        rows = model.op_times(n_units, reps=5)
        tot = sum(r[3] for r in rows if r[0] != 'q_net')
        op_table = []
        H, Wd, Cx = cfg.x_shape
        for net, name, macs, t_ms, nn in rows:
            tf = 2.0 * macs * nn / (t_ms * 1e-3) / 1e12 if t_ms > 0 else 0.0
            row = {'op': name, 'units': nn, 'ms': round(t_ms, 4), 'tflops': round(tf, 1),
                   'share': round(t_ms / tot, 4) if net != 'q_net' else None}
            # the HBM-bound CUDA-core ops: algorithmic bytes per unit -> achieved GB/s against the copy peak
            by = None
            if '[heads' in name and args.arch == 'A':
                by = H * Wd * (96 * 2 + 16 * 2 + Cx * 4)                 # trunk row in, D_psi's raw input row out, x in (likelihood fused)
            elif '[fused block' in name:
                by = H * Wd * (16 * 2) * 4                                # x in, r, raw and act out: 16-channel 16-bit rows
            if by and t_ms > 0:
                row['hbm_gbs'] = round(by * nn / (t_ms * 1e-3) / 1e9, 1)
                row['hbm_frac'] = round(row['hbm_gbs'] / peaks['hbm_gbs'], 4)
            op_table.append(row)
        dom = max((r for r in rows if r[0] != 'q_net'), key=lambda r: r[3])
        ach = 2.0 * dom[2] * dom[4] / (dom[3] * 1e-3) / 1e12
        all_flops = sum(2.0 * r[2] * r[4] for r in rows if r[0] != 'q_net')
        traffic, traffic_src = None, None
        for tj_name in ('r2_traffic.json', 'r1_traffic.json'):           # DRAM bytes per launch from the committed ncu capture
            try:
                with open(os.path.join(ROOT, 'profiles', tj_name)) as f:
                    tj = json.load(f)
                if args.chunk == tj.get('chunk') and args.arch == 'A' and args.z_dim == 128:
                    traffic = tj['bytes_per_launch'].get(dom[1].split(' [')[0])
                    traffic_src = tj['source']
                    if traffic is not None:
                        break
            except (OSError, ValueError, KeyError):
                pass
        roofline = {'bound': 'tensor', 'kernel': 'tcgen05/TMEM/TMA implicit GEMM, op ' + dom[1],
                    'traffic_unit': 'bytes of DRAM per launch', 'traffic_source': traffic_src,
                    'achieved': round(ach, 1), 'peak': peaks['bf16_tflops'], 'unit': 'TFLOP/s',
                    'frac': round(ach / peaks['bf16_tflops'], 4), 'traffic': traffic,
                    'peak_source': peaks['_source'] + ' burst 16-bit dense peak (kernel timed alone, CUDA events)',
                    'share_of_chunk_time': round(dom[3] / tot, 4),
                    'all_ops_tflops': round(all_flops / (tot * 1e-3) / 1e12, 1),
                    'all_ops_frac': round(all_flops / (tot * 1e-3) / 1e12 / peaks['bf16_tflops'], 4)}
        aux = aux_kernel_table(model, n_units, B, peaks, sampler=False)
    step_tflops = flops_total_rank / (ms / 1e3) / 1e12
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_throughput(args, steps=max(1, args.cpu_steps), warmup=max(0, args.cpu_warmup))
        cpu.pop('ms_per_step')
    mala = None
    model.close()
    del model
    torch.cuda.empty_cache()
    if world == 1 and not args.no_mala and not strong:
        try:
            mala = mala_bench(args, local, peaks)
        except Exception as e:                                   # the headline line must survive a sampler problem
            mala = {'error': repr(e)[:300]}

    line = {
        'metric': metric_name(args), 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': steps_done, 'warmup': W,
        'ms_per_step': ms / steps_done, 'higher_is_better': True, 'scaling': 'strong' if strong else 'weak', 'vs_baseline': None,
        'dtype': args.operand_dtype, 'data': 'synthetic', 'config': workload_config(args, world),
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h},
        'gpu_launches': int(launches),
        'roofline': roofline,
        'step_tensor': {'algorithmic_tflops_per_gpu': round(step_tflops, 1),      # flops are per rank
                        'frac_of_sustained_peak': round(step_tflops / peaks['bf16_tflops_sustained'], 4),
                        'gflop_per_image': round(2.0 * (mq + (mg + md) * K + md) / 1e9, 1),
                        'prior_branch_gflop_per_step': round(2.0 * (mg + md) * Kp / 1e9, 1)},
        'cpu_baseline': cpu,
        'metrics': {k: _finite(metrics.get(k)) for k in ('test_nll', 'adv_test_nll', 'test_lb', 'adv_test_lb', 'test_recon', 'reconstruct_energy',
                                                'real_energy', 'pn_energy', 'pd_energy', 'kl_adv_and_gaussian', 'test_mse')},
        'mala': mala,
        'aux_kernels': aux,
        'ops': op_table,
    }
    if strong:
        line['images'] = args.images
        line['shard_sizes'] = shard_sizes
        line['batches_per_rank'] = [-(-s // B) for s in shard_sizes]
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
