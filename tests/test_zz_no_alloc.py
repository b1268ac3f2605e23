"""apvae_is_ll_noalloc — the estimator tail that allocates nothing (fixed device scratch inside the library) — against the
existing entry points it restates: apvae_discretized_logistic_log_prob into a caller-owned buffer followed by
apvae_is_reduce.  Through the C ABI, two calls in a row (different sizes, so the second call runs over the first one's
left-over scratch contents), plus the size limit.  Both sides run the same device functions in the same order, so the gate
is the tightest one the estimator tests use (rtol 2e-6, tests/test_model_gpu.py::test_estimator_tail_exact_given_same_inputs).

No counterpart in the reference.  This file is named so
that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 2e-6


def _p(t):
    return C.c_void_p(t.data_ptr())


def _case(K, B, group, seed):
    """Script-shaped inputs: pixels on the 8-bit grid in [-1, 1), tanh-range means, log-scales around -2."""
    rs = np.random.RandomState(seed)
    x = ((rs.randint(0, 256, size=(1, B, group)) - 127.5) / 256.0 * 2) * np.ones((K, 1, 1))
    mean = np.tanh(rs.normal(size=(K, B, group)))
    log_scale = rs.normal(size=(K, B, group)) * 0.5 - 2.0
    d = 128
    z = rs.normal(size=(K, B, d))
    log_pn = -0.5 * (z ** 2).sum(-1) - 0.5 * d * np.log(2 * np.pi)
    log_q = log_pn + rs.normal(size=(K, B)) * 3 + 100.0
    d_g = rs.normal(size=(K, B)) * 0.2
    return x, mean, log_scale, log_q, log_pn, d_g, d


def _both(K, B, group, seed):
    import torch
    from apvae import _lib
    from apvae._lib import lib, check
    x, mean, ls, lq, lpn, dg, d = _case(K, B, group, seed)
    dev = lambda a: torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device='cuda')
    x, mean, ls, lq, lpn, dg = (dev(a) for a in (x, mean, ls, lq, lpn, dg))
    bin_size, lo, hi, eps = 2.0 / 256.0, -1.0 + 1.0 / 256.0, 1.0 - 1.0 / 256.0, 1e-10
    w_xi, log_Z = 2.5, 1.0
    # existing entry points: caller-owned intermediate
    log_px = torch.empty(K * B, dtype=torch.float32, device='cuda')
    want = torch.zeros(_lib.PER_IMAGE_ROWS, B, dtype=torch.float32, device='cuda')
    check(lib.apvae_discretized_logistic_log_prob(_p(x), _p(mean), _p(ls), K * B, group, bin_size, 1, lo, 1, hi, 1, eps,
                                                  _p(log_px), None))
    check(lib.apvae_is_reduce(_p(log_px), _p(lq), _p(lpn), _p(dg), None, K, B, d, w_xi, log_Z, _p(want), None))
    # the no-allocation entry point; outputs allocated BEFORE the call, allocator statistics bracket it
    got = torch.zeros(_lib.PER_IMAGE_ROWS, B, dtype=torch.float32, device='cuda')
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()['allocation.all.allocated']
    check(lib.apvae_is_ll_noalloc(_p(x), _p(mean), _p(ls), group, bin_size, 1, lo, 1, hi, 1, eps, _p(lq), _p(lpn), _p(dg),
                                  K, B, d, w_xi, log_Z, _p(got), None))
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()['allocation.all.allocated'] == before      # nothing went through torch's allocator
    return got.cpu().numpy().astype(np.float64), want.cpu().numpy().astype(np.float64)


def test_no_alloc_estimator_tail_matches_the_existing_entry_points_twice_in_a_row():
    from apvae import _lib
    rows = [_lib.ROW_LL, _lib.ROW_LB, _lib.ROW_ADV_LL, _lib.ROW_ADV_LB, _lib.ROW_RECON, _lib.ROW_DG, _lib.ROW_LZ_MAX,
            _lib.ROW_LZ_SUM, _lib.ROW_MSE]
    # first call fills 64 * 37 scratch elements; the second, smaller and with other inputs, runs over what it left behind;
    # the third grows again (up to the documented maximum of 65536 samples)
    for K, B, group, seed in ((64, 37, 3072, 11), (10, 5, 784, 12), (1024, 64, 48, 13)):
        got, want = _both(K, B, group, seed)
        assert np.isfinite(want[rows]).all()
        np.testing.assert_allclose(got[rows], want[rows], rtol=RTOL, atol=0)
        assert np.all(got[_lib.ROW_REAL_E] == 0)                                    # untouched, like apvae_is_reduce


def test_no_alloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from apvae import _lib
    from apvae._lib import lib, last_error
    K, B, group = 1025, 64, 1                                                       # 65600 > 65536
    assert K * B > _lib.NOALLOC_MAX_SAMPLES
    t = torch.zeros(K * B, dtype=torch.float32, device='cuda')
    out = torch.zeros(_lib.PER_IMAGE_ROWS, B, dtype=torch.float32, device='cuda')
    st = lib.apvae_is_ll_noalloc(_p(t), _p(t), _p(t), group, 2.0 / 256.0, 1, -1.0, 1, 1.0, 1, 1e-10, _p(t), _p(t), _p(t),
                                 K, B, 128, 1.0, 0.0, _p(out), None)
    assert st == _lib.ERR_INVALID
    assert 'APVAE_NOALLOC_MAX_SAMPLES' in last_error(None)
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0                                            # nothing was launched
