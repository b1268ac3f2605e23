"""GPU diagnostic for the tap-GEMM kernel: runs the simplest GEMM shapes with
structured inputs and prints where results differ, so one gpurun call is enough
to tell a descriptor / swizzle / TMEM-lane bug from a planner bug."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402

import apvae  # noqa: E402
from helpers import bf16_round  # noqa: E402


def run(n, fin, fout, mode):
    rs = np.random.RandomState(0)
    if mode == 'rand':
        x = bf16_round(rs.normal(size=(n, fin)))
        k = bf16_round(rs.normal(size=(fin, fout)) / np.sqrt(fin))
    elif mode == 'eye':      # out[r, c] = x[r, c]
        x = bf16_round(rs.randint(-8, 8, size=(n, fin)).astype(np.float32))
        k = np.zeros((fin, fout), np.float32)
        for i in range(min(fin, fout)):
            k[i, i] = 1
    else:                    # rowid: out[r, c] = r (checks row/lane mapping)
        x = np.zeros((n, fin), np.float32)
        x[:, 0] = np.arange(n) % 128
        k = np.zeros((fin, fout), np.float32)
        k[0, :] = 1
    got = apvae.dense(torch.tensor(x, device='cuda'), k, None).cpu().numpy()
    want = x.astype(np.float64) @ k.astype(np.float64)
    err = np.abs(got - want)
    ok = err.max() <= 2e-4 * max(1, np.abs(want).max())
    print('dense n=%d fin=%d fout=%d mode=%s  max_err=%.4g  %s' % (n, fin, fout, mode, err.max(), 'OK' if ok else 'FAIL'))
    if not ok:
        bad = np.argwhere(err > 1e-3)
        print('  bad count %d of %d; first bad (row, col):' % (len(bad), err.size), bad[:8].tolist())
        print('  bad rows:', sorted(set(bad[:, 0].tolist()))[:40])
        print('  bad cols:', sorted(set(bad[:, 1].tolist()))[:40])
        r, c = bad[0]
        print('  got row %d :' % r, np.round(got[r, :16], 3).tolist())
        print('  want row %d:' % r, np.round(want[r, :16], 3).tolist())
    return ok


if __name__ == '__main__':
    print(torch.cuda.get_device_name(0))
    allok = True
    for fin in (16, 32, 64, 128):
        for fout in (16, 64, 256):
            for mode in ('rowid', 'eye', 'rand'):
                try:
                    allok &= run(128, fin, fout, mode)
                except Exception as e:     # noqa: BLE001
                    print('EXC', fin, fout, mode, repr(e))
                    allok = False
                    if 'CUDA' in repr(e) or 'cuda' in repr(e):
                        sys.exit(2)
    print('ALL OK' if allok else 'SOME FAILED')
