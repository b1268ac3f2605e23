"""The CUDA estimator kernels on the reference's own known-answer vectors, through the C ABI
(apvae_discretized_logistic_log_prob / apvae_bernoulli_log_prob / apvae_is_reduce).

Fixtures: tests/golden/reference_kats.npz, written by tests/golden/make_golden.py which EXECUTES the reference's numpy
test helpers (tfsnippet/tests/distributions/test_discretized.py:15-42,232-363; tests/variational/test_evaluation.py:9-47;
tests/variational/test_objectives.py).  The same device functions (csrc/likelihood.cuh) score the pixels inside the fused
heads kernel of the hot path, so this pins the likelihood arithmetic of apvae_is_nll to the reference's vectors.

The reference runs these tests in float64; the path (like the reference's models) computes in float32, where
log(sigmoid(hi) - sigmoid(lo)) loses absolute accuracy ~2^-23 / (sigmoid(hi) - sigmoid(lo)) to cancellation.  The
tolerance below is written per element from that conditioning: 1e-4 + 2e-6 / p, p = exp(golden log-prob)
(the kernel evaluates the two sigmoids with the hardware ex2 / rcp: a few ulp of 1.0 in absolute terms).
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'reference_kats.npz'))


def _dev(a):
    import torch
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device='cuda')


def _dl(x, mean, log_scale, bin_size, min_val, max_val, biased_edges=True, epsilon=1e-7, group=1):
    """DiscretizedLogistic(mean, log_scale, ...).log_prob(x, group_ndims) on the GPU; arrays broadcast on the host."""
    import torch
    from apvae._lib import lib, check
    shape = np.broadcast_shapes(np.shape(x), np.shape(mean), np.shape(log_scale))
    xb, mb, sb = (_dev(np.broadcast_to(a, shape)) for a in (x, mean, log_scale))
    n = int(np.prod(shape))
    rows = n // group
    out = torch.empty(rows, dtype=torch.float32, device='cuda')
    p = lambda t: C.c_void_p(t.data_ptr())
    check(lib.apvae_discretized_logistic_log_prob(
        p(xb), p(mb), p(sb), rows, group, float(bin_size), int(min_val is not None), float(min_val or 0.0),
        int(max_val is not None), float(max_val or 0.0), int(biased_edges), float(epsilon), p(out), None))
    torch.cuda.synchronize()
    return out.cpu().numpy().astype(np.float64), shape


def _close(got, want):
    tol = 1e-4 + 2e-6 / np.exp(want)          # fp32 conditioning of log(cdf_high - cdf_low), see the module docstring
    err = np.abs(got - want)
    assert np.all(err <= tol), (float(err.max()), float((err / tol).max()))


def test_cuda_discretized_logistic_on_the_reference_payload():
    """test_discretized.py:232-310: every constructor variant of the reference's test_log_prob."""
    x, mean, ls = GOLD['dl_x'], GOLD['dl_mean'], GOLD['dl_log_scale']
    bin_size, lo, hi = GOLD['dl_params']
    for name, kw in (('dl_unbiased', dict(min_val=lo, max_val=hi, biased_edges=False)),
                     ('dl_nomin', dict(min_val=None, max_val=hi)),
                     ('dl_nomax', dict(min_val=lo, max_val=None)),
                     ('dl_biased', dict(min_val=lo, max_val=hi))):
        got, shape = _dl(x, mean, ls, bin_size, **kw)
        _close(got.reshape(shape), GOLD[name])
    # group_ndims = 2: sums over the last two axes (5 x 4 = 20 consecutive elements)
    got, shape = _dl(x, mean, ls, bin_size, lo, hi, group=20)
    want = GOLD['dl_biased_g2']
    tol = (1e-4 + 2e-6 / np.exp(GOLD['dl_biased'])).sum(axis=(-1, -2))
    assert np.all(np.abs(got.reshape(want.shape) - want) <= tol)


def test_cuda_discretized_logistic_probabilities_sum_to_one():
    """test_discretized.py:312-334: the eleven bins of [0, 1] at bin_size 1/10 carry total probability 1."""
    got, shape = _dl(GOLD['ps_x'], GOLD['ps_mean'], GOLD['ps_log_scale'], 1 / 10., 0., 1.)
    prob = np.exp(got.reshape(shape)).sum(axis=0)
    assert prob.shape == (3, 2, 5, 4)
    np.testing.assert_allclose(prob, np.ones_like(prob), rtol=1e-5, atol=2e-6)


def test_cuda_discretized_logistic_extreme_fallback_branch():
    """test_discretized.py:336-363: cdf_delta < epsilon -> x_mid + log_delta - 2 softplus(x_mid)."""
    ls = GOLD['ex_log_scale']
    got, _ = _dl(np.zeros_like(ls), np.zeros_like(ls), ls, 1 / 256., None, None, epsilon=1e-7)
    assert np.all(np.isfinite(got))
    np.testing.assert_allclose(got, np.log(1 / 256.) - ls - 2 * np.log(2.0), rtol=2e-6)
    naive = GOLD['ex_naive']
    ok = np.isfinite(naive)
    np.testing.assert_allclose(got[ok], naive[ok], rtol=1e-5, atol=1e-1)       # the reference test's own bound


def test_cuda_likelihood_with_the_scripts_parameters():
    """The p_net x node exactly as dmm_cifar10.py:612-619 builds it (bin 2/256, edges at -1 + 1/256 and 1 - 1/256,
    epsilon 1e-10) on the 256-level pixel grid, against the float64 oracle: edge bins, middle bins and the fallback
    branch (tiny scales) all occur."""
    rs = np.random.RandomState(0)
    u8 = np.concatenate([np.arange(256), rs.randint(0, 256, size=4096 - 256)])
    x = (u8 - 127.5) / 256.0 * 2
    mean = np.tanh(rs.normal(size=x.shape) * 1.5)
    ls = rs.uniform(-9.0, 1.0, size=x.shape)
    want = O.discretized_logistic_log_prob(x, mean, ls, 2.0 / 256.0, -1.0 + 1.0 / 256.0, 1.0 - 1.0 / 256.0, epsilon=1e-10)
    got, _ = _dl(x, mean, ls, 2.0 / 256.0, -1.0 + 1.0 / 256.0, 1.0 - 1.0 / 256.0, epsilon=1e-10)
    # outside the fp32-cancellation regime (p > 1e-3) the kernel is exact to fp32 rounding of its inputs
    good = want > np.log(1e-3)
    assert good.sum() > 1000
    np.testing.assert_allclose(got[good], want[good], rtol=2e-4, atol=2e-4)
    # everywhere: finite, and within the conditioning bound (fallback branch elements: relative)
    assert np.all(np.isfinite(got))
    tol = np.maximum(1e-4 + 2e-6 / np.exp(np.maximum(want, -60)), 2e-4 * np.abs(want))
    assert np.all(np.abs(got - want) <= tol)
    assert (want < -30).any()                                  # the fallback branch was exercised


def test_cuda_bernoulli_clip_branch():
    """Arch M's likelihood (dmm_bernouli_fashion.py:619-626): G_theta ends in tanh, so about half the means are
    negative and sit on clip(., 1e-7, .), and saturated ones sit on 1 - 1e-7.  x_mean fed in directly, so the kernel is
    compared exactly: against the oracle in float32 (the reference's dtype; 1 - 1e-7 rounds to 1 - 2^-23 there) and
    against float64 away from the upper clip."""
    import torch
    from apvae._lib import lib, check
    rs = np.random.RandomState(1)
    n, group = 64, 784
    mean = np.tanh(rs.normal(size=(n, group)) * 2.0).astype(np.float32)
    mean[0, :8] = [1.0, 0.99999999, 1.0 - 1e-7, 0.0, -0.0, 1e-7, 5e-8, -1.0]
    x = (rs.uniform(size=(n, group)) < 0.5).astype(np.float32)
    assert (mean < 1e-7).mean() > 0.4 and (mean >= np.float32(1.0 - 1e-7)).any()
    out = torch.empty(n, dtype=torch.float32, device='cuda')
    p = lambda t: C.c_void_p(t.data_ptr())
    xd, md = _dev(x), _dev(mean)                                   # (keep the device tensors alive across the call)
    check(lib.apvae_bernoulli_log_prob(p(xd), p(md), n, group, p(out), None))
    torch.cuda.synchronize()
    got = out.cpu().numpy().astype(np.float64)
    want32 = O.bernoulli_log_prob_from_mean(x[:, None].reshape(n, 28, 28, 1), mean.reshape(n, 28, 28, 1), 3).astype(np.float64)
    np.testing.assert_allclose(got, want32, rtol=2e-6)
    # float64 oracle: identical except where the mean reaches the upper clip (fp32 cannot hold 1 - 1e-7)
    rows = ~(mean >= np.float32(1.0 - 2e-7)).any(axis=1)
    want64 = O.bernoulli_log_prob_from_mean(x.astype(np.float64).reshape(n, 28, 28, 1),
                                            mean.astype(np.float64).reshape(n, 28, 28, 1), 3)
    np.testing.assert_allclose(got[rows], want64[rows], rtol=2e-6)


def test_cuda_is_reduce_on_the_reference_payload():
    """tests/variational/test_evaluation.py:9-47 (importance_sampling_log_likelihood, the reference's atol 1e-4) and
    test_objectives.py (elbo): log_p [13] broadcast over K = 7 samples, log_q [7, 13]."""
    import torch
    from apvae import _lib
    from apvae._lib import lib, check
    log_p, log_q = GOLD['vi_log_p'], GOLD['vi_log_q']
    K, B = log_q.shape
    lp = _dev(np.broadcast_to(log_p, (K, B)))
    lq = _dev(log_q)
    zero = torch.zeros(K * B, dtype=torch.float32, device='cuda')
    out = torch.full((_lib.PER_IMAGE_ROWS, B), 7.0, dtype=torch.float32, device='cuda')
    p = lambda t: C.c_void_p(t.data_ptr())
    check(lib.apvae_is_reduce(p(lp), p(lq), p(zero), p(zero), None, K, B, 0, 0.0, 0.0, p(out), None))
    torch.cuda.synchronize()
    o = out.cpu().numpy().astype(np.float64)
    np.testing.assert_allclose(o[_lib.ROW_LL], GOLD['vi_is_ll'], atol=1e-4, rtol=0)
    np.testing.assert_allclose(o[_lib.ROW_LB], GOLD['vi_elbo'], atol=1e-4, rtol=0)
    # with log N(z) = D(G(z)) = 0, w_xi = 0, log_Z = 0 and d = 0 the energy rows repeat the plain ones
    np.testing.assert_allclose(o[_lib.ROW_ADV_LL], GOLD['vi_is_ll'], atol=1e-4, rtol=0)
    assert np.all(o[_lib.ROW_REAL_E] == 7.0)                      # not this kernel's row


def test_cuda_log_mean_exp_on_the_reference_payload():
    """tests/ops/test_misc.py:39-87: log_mean_exp over the last axis of linspace(0, 10, 1000).reshape(5, 10, 20), through
    the same per-image reduction (K = 20 'samples', B = 50 'images')."""
    import torch
    from apvae import _lib
    from apvae._lib import lib, check
    x = GOLD['lse_x']                                               # [5, 10, 20]
    K, B = 20, 50
    lp = _dev(np.moveaxis(x, -1, 0).reshape(K, B))
    zero = torch.zeros(K * B, dtype=torch.float32, device='cuda')
    out = torch.empty((_lib.PER_IMAGE_ROWS, B), dtype=torch.float32, device='cuda')
    p = lambda t: C.c_void_p(t.data_ptr())
    check(lib.apvae_is_reduce(p(lp), p(zero), p(zero), p(zero), None, K, B, 0, 0.0, 0.0, p(out), None))
    torch.cuda.synchronize()
    got = out.cpu().numpy().astype(np.float64)[_lib.ROW_LL].reshape(5, 10)
    np.testing.assert_allclose(got, GOLD['lme_last'], rtol=2e-6)
