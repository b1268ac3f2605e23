#!/usr/bin/env python
"""A/B of the G_theta heads kernel (heads.cu) on one B200: fused likelihood on / off, fp32 heads written or not.
Prints ms per launch over `--chunk` samples and the achieved HBM GB/s on the kernel's algorithmic bytes."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--chunk', type=int, default=7400)
    args = ap.parse_args()
    import torch
    from bench import synthetic_weights, synthetic_images
    from apvae.model import ExpConfig, Model, weight_schema
    from apvae._lib import lib, check
    cfg = ExpConfig('A', z_dim=128, max_batch=8, max_chunk=args.chunk)
    m = Model(cfg, synthetic_weights(weight_schema(cfg)))
    x = synthetic_images(8, cfg.x_shape)
    name = C.create_string_buffer(256)
    macs, ms, cnt = C.c_double(), C.c_float(), C.c_int()
    check(lib.apvae_op_count(m._h, 1, C.byref(cnt)))
    heads = cnt.value - 1

    def t(label, bytes_per_pixel):
        best = 1e9
        for _ in range(3):
            check(lib.apvae_op_time(m._h, 1, heads, args.chunk, 10, name, 256, C.byref(macs), C.byref(ms), None), m._h)
            best = min(best, ms.value)
        gb = bytes_per_pixel * 1024 * args.chunk / (best * 1e-3) / 1e9
        print('%-52s %.4f ms  %.0f GB/s (%d B/pixel)' % (label, best, gb, bytes_per_pixel))

    m.is_nll(x, 1000, seed=1)                    # leaves the heads op in its evaluation mode: likelihood on, no fp32 heads
    torch.cuda.synchronize()
    t('fused likelihood, raw out only (is_nll mode)', 192 + 32)
    m.decode(np.zeros((4, 128), np.float32))     # decode mode: fp32 heads written, likelihood off
    torch.cuda.synchronize()
    t('no likelihood, fp32 heads + raw out (decode mode)', 192 + 24 + 32)
    m.energy_of_latents(np.zeros((4, 128), np.float32))
    torch.cuda.synchronize()
    t('no likelihood, raw out only (energy mode)', 192 + 32)


if __name__ == '__main__':
    main()
