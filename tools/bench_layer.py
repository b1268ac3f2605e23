"""Times ONE conv / deconv layer on the GPU through the C ABI (apvae_plan_bench): kernel tuning and
ncu profiling of a single op in isolation.
  python tools/bench_layer.py [--transpose] --k 3 --stride 1 --h 16 --w 16 --cin 192 --cout 192 --n 256
--tuning patch=-1,patch_macc=2,... (apvae_tuning fields; the old APVAE_* variable names are still read, by the TOOLS only) selects the kernel variant."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))
import torch  # noqa: E402

from apvae.plan import Plan, pad16  # noqa: E402
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import _tuning  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--transpose', action='store_true')
    ap.add_argument('--k', type=int, default=3)
    ap.add_argument('--stride', type=int, default=1)
    ap.add_argument('--h', type=int, default=16)
    ap.add_argument('--w', type=int, default=16)
    ap.add_argument('--cin', type=int, default=192)
    ap.add_argument('--cout', type=int, default=192)
    ap.add_argument('--n', type=int, default=256)
    ap.add_argument('--reps', type=int, default=10)
    ap.add_argument('--outs', default='raw,act', help='comma list of raw, act, f32')
    ap.add_argument('--res', default='none', help='none | same | even (residual added in the epilogue)')
    ap.add_argument('--tuning', default='', help='apvae_tuning overrides, e.g. patch_2cta=-1,patch_macc=2')
    a = ap.parse_args(argv)
    rs = np.random.RandomState(0)
    shp = (a.k, a.k, a.cout, a.cin) if a.transpose else (a.k, a.k, a.cin, a.cout)
    kern = (rs.normal(size=shp) / np.sqrt(a.k * a.k * a.cin)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=a.transpose, stride=a.stride, in_h=a.h, in_w=a.w, kernel=kern,
                      bias=np.zeros(a.cout, np.float32))], a.n, dtype='bf16', tuning=_tuning.parse(a.tuning))
    i = plan.info
    src = torch.randn((a.n, a.h, a.w, pad16(a.cin)), device='cuda').to(torch.bfloat16)
    kw = {}
    if 'raw' in a.outs:
        kw['out_raw'] = torch.empty((a.n, i.ho, i.wo, i.npad), dtype=torch.bfloat16, device='cuda')
    if 'act' in a.outs:
        kw['out_act'] = torch.empty((a.n, i.ho, i.wo, i.npad), dtype=torch.bfloat16, device='cuda')
    if 'f32' in a.outs:
        kw['out_f32'] = torch.empty((a.n, i.ho, i.wo, a.cout), dtype=torch.float32, device='cuda')
        kw['f32_cols'] = a.cout
    if a.res != 'none':
        from apvae import _lib
        d = 1 if a.res == 'same' else 2
        kw['res'] = torch.randn((a.n, i.ho // d, i.wo // d, i.npad), device='cuda').to(torch.bfloat16)
        kw['res_mode'] = _lib.RES_SAME if a.res == 'same' else _lib.RES_EVEN
    ms = plan.run([src], a.n, bench_reps=a.reps, **kw)
    tf = 2.0 * i.macs_per_image * a.n / (ms * 1e-3) / 1e12
    print('%s k=%d s=%d %dx%d %d->%d n=%d res=%s | patch=%d macc=%d cluster=%d a_st=%d b_st=%d kc=%d | %.4f ms  %.1f TFLOP/s'
          % ('deconv' if a.transpose else 'conv', a.k, a.stride, a.h, a.w, a.cin, a.cout, a.n, a.res, i.patch, i.patch_macc,
             i.patch_cluster, i.patch_a_stages, i.patch_b_stages, i.kc, ms, tf))


if __name__ == '__main__':
    main()
