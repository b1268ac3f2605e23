"""BASELINE configs[3]: CelebA-shape (64x64x3, arch B = dmm_for_sampling.py) latent MALA sampling of the energy prior,
100 steps x 1024 chains, then x = G_theta(z) -> uint8.  Prints one JSON line (chain-steps/s, algorithmic TFLOP/s).
  python tools/bench_mala.py [--precise] [--chains 1024] [--steps 100] [--dataset celebA] [--arch B] [--chunk 1024] [--op-times]
Random-init weights, z0 ~ N(0, I) (stand-in for the G_omega latent GAN), in-kernel Philox noise.
Under torchrun every rank runs `--chains` independent chains on its own GPU (the chains shard with no collective on the
data path); rank 0 prints the aggregate over the slowest rank's device time."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from bench import synthetic_weights  # noqa: E402
from apvae.model import ExpConfig, Model, weight_schema  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--arch', default='B')
    ap.add_argument('--dataset', default='celebA')
    ap.add_argument('--z-dim', type=int, default=128)
    ap.add_argument('--chains', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--chunk', type=int, default=1024)
    ap.add_argument('--precise', type=int, nargs='?', const=1, default=0,
                    help='split-precision forward: 1 = activations and kernels hi + lo (3 terms), 2 = activations only (2 terms)')
    ap.add_argument('--operand-dtype', default='fp16', choices=('fp16', 'bf16'))
    ap.add_argument('--op-times', action='store_true', help='also print per-op timings of the forward and adjoint programs')
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', '0'), ('WORLD_SIZE', '1'), ('LOCAL_RANK', '0')))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    cfg = ExpConfig(a.arch, dataset=a.dataset if a.arch == 'B' else None, z_dim=a.z_dim, max_batch=8, max_chunk=a.chunk,
                    enable_grad=True, precise=a.precise, operand_dtype=a.operand_dtype)
    m = Model(cfg, synthetic_weights(weight_schema(cfg)), device=local)
    z0 = torch.from_numpy(np.random.RandomState(2 + rank).normal(size=(a.chains, cfg.z_dim)).astype(np.float32)).cuda()
    m.mala(z0, 2, seed=1, history=False)                      # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = m.mala(z0, a.steps, seed=7, history=True)
    img = m.sample_images(out['z'])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
        dist.barrier()
        if rank != 0:
            dist.destroy_process_group()
            return
    _, mg, md = m.macs()
    # algorithmic: one forward + one input-gradient backward of D o G per chain-step (backward ~ forward), SURVEY 8(d)
    flops = 4.0 * (mg + md) * a.chains * a.steps * world
    acc = float((out['history_ratio'] > 0.5).float().mean())
    if a.op_times and rank == 0:
        rows = m.op_times(min(a.chunk, a.chains), reps=5, grad=True)
        tot = sum(r[3] for r in rows if r[0] != 'q_net')
        for net, name, macs, t_ms, nn in rows:
            if net == 'q_net':
                continue
            tf = 2.0 * macs * nn / (t_ms * 1e-3) / 1e12 if t_ms > 0 else 0.0
            print('%-9s %-64s %8.4f ms %8.1f TFLOP/s %5.1f%%' % (net, name, t_ms, tf, 100 * t_ms / tot), file=sys.stderr)
        print('sum of ops per chunk of %d chains: %.3f ms (forward + adjoint programs; a MALA step runs them once per chunk)'
              % (min(a.chunk, a.chains), tot), file=sys.stderr)
    print(json.dumps({
        'metric': 'MALA chain-steps/sec (%s %s, d=%d, %d chains x %d steps)' % (a.arch, a.dataset, cfg.z_dim, a.chains, a.steps),
        'value': world * a.chains * a.steps / (ms / 1e3), 'unit': 'chain-steps/s', 'n_gpus': world, 'ms_total': ms,
        'algorithmic_tflops': flops / (ms / 1e3) / 1e12, 'precise': int(a.precise),
        'mean_accept_ratio': float(out['history_ratio'].mean()), 'frac_ratio_gt_half': acc,
        'images': list(img.shape), 'dtype': a.operand_dtype + ('' if not a.precise else ' x3 (split precision)' if a.precise == 1 else ' x2 (activation split)')}))


if __name__ == '__main__':
    main()
