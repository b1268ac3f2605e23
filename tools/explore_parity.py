#!/usr/bin/env python
"""Which roundings carry the energy / likelihood error of the 16-bit operand path?  (CPU only, test tooling.)

Emulates the library's numerics in numpy on top of the oracle's layer functions: every conv / deconv / dense
input and kernel is rounded to the operand format (bf16 or fp16), accumulation is exact (float64; the tensor
cores accumulate in fp32), every stored block output is rounded (it is both the next block's shortcut and,
activated, the next conv's input).  Prints max / mean |error| of x_mean, D(G(z)), D(x) and the relative error
of log p(x|z) against the float64 oracle for a list of (G format, D format) schemes.

  python tools/explore_parity.py [--arch A] [--n 16]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracle as O                                     # noqa: E402
from oracle import layers as L, nets as N              # noqa: E402
from helpers import bf16_round                         # noqa: E402


def rounder(fmt):
    if fmt == 'bf16':
        return lambda a: bf16_round(np.asarray(a, np.float32)).astype(np.float64)
    if fmt == 'fp16':
        return lambda a: np.asarray(a, np.float64).astype(np.float16).astype(np.float64)
    if fmt == 'bf16x2':                                # hi + lo bf16 pair: 16 mantissa bits
        def r(a):
            a32 = np.asarray(a, np.float32)
            hi = bf16_round(a32)
            return (hi.astype(np.float64) + bf16_round(a32 - hi).astype(np.float64))
        return r
    return lambda a: np.asarray(a, np.float64)


def block(x, params, scope, cout, k, s, conv_fn, resize_at_exit, ra, rw):
    """oracle.layers._block with the library's rounding points: x arrives already rounded (stored tensor)."""
    cin = x.shape[-1]
    use_sc = (s != 1) or (cin != cout)
    p = lambda name: params[scope + '/' + name]
    sc = conv_fn(x, rw(p('shortcut/kernel')), p('shortcut/bias'), s) if use_sc else x
    s0, s1 = (1, s) if resize_at_exit else (s, 1)
    r0 = conv_fn(ra(L.leaky_relu(x)), rw(p('residual/conv_0/kernel')), p('residual/conv_0/bias'), s0)
    r1 = conv_fn(ra(L.leaky_relu(r0)), rw(p('residual/conv_1/kernel')), p('residual/conv_1/bias'), s1)
    return ra(sc + r1)


def g_theta(cfg, params, z, ra, rw, head_exact=True):
    gh, gw = N._g_base_hw(cfg)
    if cfg.arch == 'A':
        h = ra(z.reshape(z.shape[:-1] + (gh, gw, N._zc(cfg))))
    else:
        h = ra(L.dense(ra(z), rw(params['G_theta/level_0/kernel']), params['G_theta/level_0/bias']))
        h = h.reshape(h.shape[:-1] + (gh, gw, 64))
    for scope, cout, k, s in N._g_blocks(cfg):
        h = block(h, params, 'G_theta/' + scope, cout, k, s, L.deconv2d, False, ra, rw)
    name = 'G_theta/feature_map_mean_to_pixel' if cfg.arch == 'M' else 'G_theta/x_mean'
    xm = L.conv2d(h, params[name + '/kernel'], params[name + '/bias'], 1, np.tanh)      # heads.cu: fp32 weights
    xs = None
    if cfg.arch != 'M' and not cfg.uniform_scale:
        xs = L.conv2d(h, params['G_theta/x_logstd/kernel'], params['G_theta/x_logstd/bias'], 1)
    return xm, xs


def d_psi(cfg, params, x, ra, rw, first_exact=True):
    h = ra(x)
    for scope, cout, k, s in N._d_blocks(cfg):
        h = block(h, params, 'D_psi/' + scope, cout, k, s, L.conv2d, True, ra, rw)
    hf = h.reshape(h.shape[:-3] + (-1,))
    hf = L.dense(hf, rw(params['D_psi/level_-2/kernel']), params['D_psi/level_-2/bias'])
    hf = L.dense(hf, params['D_psi/level_-1/kernel'], params['D_psi/level_-1/bias'])
    return hf[..., 0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--arch', default='A')
    ap.add_argument('--dataset', default=None)
    ap.add_argument('--z-dim', type=int, default=128)
    ap.add_argument('--n', type=int, default=16)
    args = ap.parse_args()
    kw = {'z_dim': args.z_dim} if args.arch != 'M' else {}
    if args.arch == 'B':
        kw['dataset'] = args.dataset or 'cifar10'
    cfg = O.OracleConfig(args.arch, **kw)
    params = O.init_weights(cfg)
    rs = np.random.RandomState(11)
    z = rs.normal(size=(args.n, cfg.z_dim))
    u8 = np.random.RandomState(0).randint(0, 256, size=(args.n,) + tuple(cfg.x_shape))
    x = (u8 - 127.5) / 256.0 * 2
    xm_o = O.G_theta(cfg, params, z)
    dg_o = O.D_psi(cfg, params, xm_o)
    dx_o = O.D_psi(cfg, params, x)
    print('scale: |D(G(z))| max %.3f, |D(x)| max %.3f' % (np.abs(dg_o).max(), np.abs(dx_o).max()))
    schemes = [('bf16', 'bf16', 'bf16', 'bf16'), ('fp16', 'fp16', 'fp16', 'fp16'),
               ('bf16', 'bf16', 'fp16', 'fp16'), ('fp16', 'fp16', 'bf16', 'bf16'),
               ('bf16', 'none', 'bf16', 'none'), ('none', 'bf16', 'none', 'bf16'),
               ('bf16', 'bf16', 'bf16x2', 'bf16x2'), ('bf16', 'bf16', 'bf16', 'bf16x2'), ('bf16', 'bf16', 'bf16x2', 'bf16')]
    print('%-34s %10s %10s | %10s %10s | %10s %10s' % ('G act/w, D act/w', 'xm max', 'xm mean', 'D(G) max', 'D(G) mean', 'D(x) max', 'D(x) mean'))
    for ga, gw_, da, dw in schemes:
        xm, _ = g_theta(cfg, params, z, rounder(ga), rounder(gw_))
        dg = d_psi(cfg, params, xm, rounder(da), rounder(dw))
        dx = d_psi(cfg, params, x, rounder(da), rounder(dw))
        # D's own error on the oracle's x_mean (isolates D)
        dg_iso = d_psi(cfg, params, xm_o, rounder(da), rounder(dw))
        e_xm, e_dg, e_dx = np.abs(xm - xm_o), np.abs(dg - dg_o), np.abs(dx - dx_o)
        print('G %-6s/%-6s D %-6s/%-6s     %10.2e %10.2e | %10.2e %10.2e | %10.2e %10.2e | D alone on exact x_mean: max %.2e'
              % (ga, gw_, da, dw, e_xm.max(), e_xm.mean(), e_dg.max(), e_dg.mean(), e_dx.max(), e_dx.mean(),
                 np.abs(dg_iso - dg_o).max()))


if __name__ == '__main__':
    main()
