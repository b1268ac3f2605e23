"""Pins the CPU oracle: (1) against the reference's own known-answer tests (fixtures in
tests/golden/reference_kats.npz, produced by tests/golden/make_golden.py which EXECUTES the
reference's numpy test helpers), (2) for the arithmetic the reference holds no independent
values for (conv / deconv, gradients, MALA) against independent cross-checks: torch-CPU
conv ops, adjoint identities and finite differences.  No GPU needed."""
import os

import numpy as np
import pytest

import oracle as O

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'reference_kats.npz'))


# ---------------------------------------------------------------- reference KATs
def test_discretized_logistic_matches_reference_naive_pdf():
    """tfsnippet tests/distributions/test_discretized.py:232-310 (rtol 1e-5, atol 1e-6)."""
    x, mean, ls = GOLD['dl_x'], GOLD['dl_mean'], GOLD['dl_log_scale']
    bin_size, lo, hi = GOLD['dl_params']
    kw = dict(rtol=1e-5, atol=1e-6)
    f = O.discretized_logistic_log_prob
    np.testing.assert_allclose(f(x, mean, ls, bin_size, lo, hi, biased_edges=False), GOLD['dl_unbiased'], **kw)
    np.testing.assert_allclose(f(x, mean, ls, bin_size, None, hi), GOLD['dl_nomin'], **kw)
    np.testing.assert_allclose(f(x, mean, ls, bin_size, lo, None), GOLD['dl_nomax'], **kw)
    np.testing.assert_allclose(f(x, mean, ls, bin_size, lo, hi), GOLD['dl_biased'], **kw)
    np.testing.assert_allclose(f(x, mean, ls, bin_size, lo, hi, group_ndims=2), GOLD['dl_biased_g2'], **kw)


def test_discretized_logistic_probabilities_sum_to_one():
    """test_discretized.py:312-334."""
    lp = O.discretized_logistic_log_prob(GOLD['ps_x'], GOLD['ps_mean'], GOLD['ps_log_scale'], 1 / 10., 0., 1.)
    prob = np.exp(lp).sum(axis=0)
    assert prob.shape == (3, 2, 5, 4)
    np.testing.assert_allclose(prob, np.ones_like(prob), rtol=1e-5, atol=1e-6)


def test_discretized_logistic_extreme_fallback_branch():
    """test_discretized.py:336-363: cdf_delta < epsilon -> the PixelCNN++ log-pdf approximation.
    The naive answer is -inf/garbage there; the reference test only requires closeness where the naive
    value is finite, so check finiteness + the analytic mid-point density."""
    ls = GOLD['ex_log_scale']
    lp = O.discretized_logistic_log_prob(np.asarray(0.), np.asarray(0.), ls, 1 / 256., epsilon=1e-7)
    assert np.all(np.isfinite(lp))
    np.testing.assert_allclose(lp, np.log(1 / 256.) - ls - 2 * np.log(2.0), rtol=1e-6)
    naive = GOLD['ex_naive']
    ok = np.isfinite(naive)
    np.testing.assert_allclose(lp[ok], naive[ok], rtol=1e-5, atol=1e-1)


def test_log_sum_and_mean_exp():
    """tests/ops/test_misc.py:39-87."""
    x = GOLD['lse_x']
    np.testing.assert_allclose(O.log_sum_exp(x, axis=-1), GOLD['lse_last'])
    np.testing.assert_allclose(O.log_sum_exp(x, axis=(0, 2)), GOLD['lse_02'])
    np.testing.assert_allclose(O.log_sum_exp(x), GOLD['lse_all'])
    np.testing.assert_allclose(O.log_sum_exp(x, axis=(0, 2), keepdims=True), GOLD['lse_02'].reshape(1, 10, 1))
    np.testing.assert_allclose(O.log_mean_exp(x, axis=-1), GOLD['lme_last'])
    np.testing.assert_allclose(O.log_mean_exp(x, axis=(0, 2)), GOLD['lme_02'])
    np.testing.assert_allclose(O.log_mean_exp(x), GOLD['lme_all'])


def test_maybe_clip_value():
    """tests/ops/test_misc.py:90-110."""
    x = np.linspace(0, 1, 101, dtype=np.float32)
    assert O.maybe_clip_value(x) is x
    np.testing.assert_allclose(O.maybe_clip_value(x, min_val=.1), np.maximum(x, .1))
    np.testing.assert_allclose(O.maybe_clip_value(x, max_val=.9), np.minimum(x, .9))
    np.testing.assert_allclose(O.maybe_clip_value(x, .1, .9), np.maximum(np.minimum(x, .9), .1))


def test_importance_sampling_and_elbo_payload():
    """tests/variational/test_evaluation.py:9-47, test_objectives.py (atol 1e-4)."""
    log_p, log_q = GOLD['vi_log_p'], GOLD['vi_log_q']
    np.testing.assert_allclose(O.importance_sampling_log_likelihood(log_p, log_q, 0), GOLD['vi_is_ll'], atol=1e-4)
    np.testing.assert_allclose(O.elbo_objective(log_p, log_q, 0), GOLD['vi_elbo'], atol=1e-4)
    np.testing.assert_allclose(O.elbo_objective(log_p, log_q), log_p - log_q, atol=1e-6)
    with pytest.raises(ValueError):
        O.importance_sampling_log_likelihood(log_p, log_q, None)


def test_evaluator_weighting():
    """tests/trainer/test_evaluator.py:53-103: arange(6), batch 4 -> 2.5 weighted, 3.0 unweighted."""
    assert abs(O.evaluator_average(GOLD['ev_batches'], GOLD['ev_weights']) - GOLD['ev_expected'][0]) < 1e-12
    assert abs(O.evaluator_average(GOLD['ev_batches'], None) - GOLD['ev_expected'][1]) < 1e-12


def test_conv_1x1_is_strided_matmul_and_dense_is_dot():
    """tests/layers/convolutional/test_conv2d.py:117-137, tests/layers/core/test_dense.py."""
    x, k, b = GOLD['c11_x'], GOLD['c11_k'], GOLD['c11_b']
    np.testing.assert_allclose(O.conv2d(x, k, b, 1), GOLD['c11_s1'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(O.conv2d(x, k, b, 2), GOLD['c11_s2'], rtol=1e-5, atol=1e-6)
    rs = np.random.RandomState(0)
    xx, kk, bb = rs.normal(size=(5, 3, 7)), rs.normal(size=(7, 4)), rs.normal(size=4)
    np.testing.assert_allclose(O.dense(xx, kk, bb), np.dot(xx, kk) + bb)


# ---------------------------------------------------------------- independent cross-checks (parity unpinned upstream)
KS = [(3, 1), (5, 1), (3, 2), (5, 2), (1, 1), (1, 2)]


def _torch_same_conv(x, kernel, s):
    import torch
    import torch.nn.functional as F
    kh = kernel.shape[0]
    n, h, w, c = x.shape
    _, lo_h, hi_h = O.same_padding(h, kh, s)
    _, lo_w, hi_w = O.same_padding(w, kh, s)
    xt = torch.tensor(x).permute(0, 3, 1, 2)
    xt = F.pad(xt, (lo_w, hi_w, lo_h, hi_h))
    wt = torch.tensor(kernel).permute(3, 2, 0, 1)
    return F.conv2d(xt, wt, stride=s).permute(0, 2, 3, 1).numpy()


@pytest.mark.parametrize('k,s', KS)
def test_conv2d_matches_torch_cpu(k, s):
    rs = np.random.RandomState(k * 10 + s)
    x = rs.normal(size=(2, 8, 10, 3))
    kern = rs.normal(size=(k, k, 3, 4))
    np.testing.assert_allclose(O.conv2d(x, kern, None, s), _torch_same_conv(x, kern, s), rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize('k,s', KS)
def test_deconv2d_is_the_adjoint_of_same_conv(k, s):
    """<conv(x; W), y> == <x, deconv(y; W)> with the deconv kernel laid out [kh,kw,Cout,Cin] = conv's
    [kh,kw,Cin_conv,Cout_conv] (conv2d_.py:339,464-471: conv2d_transpose is the gradient of conv2d)."""
    rs = np.random.RandomState(k * 10 + s + 1)
    h = 8
    x = rs.normal(size=(2, h, h, 3))                 # conv input (the deconv OUTPUT grid)
    kern = rs.normal(size=(k, k, 3, 5))              # conv: 3 -> 5 ; as deconv kernel: Cout=3, Cin=5
    y = rs.normal(size=(2, h // s, h // s, 5))
    lhs = np.sum(O.conv2d(x, kern, None, s) * y)
    rhs = np.sum(x * O.deconv2d(y, kern, None, s))
    assert abs(lhs - rhs) < 1e-9 * max(1.0, abs(lhs))


@pytest.mark.parametrize('arch,kw', [('A', dict(z_dim=64)), ('B', dict(dataset='cifar10', z_dim=24)), ('M', dict(z_dim=12))])
def test_energy_gradient_by_finite_differences(arch, kw):
    cfg = O.OracleConfig(arch, **kw)
    params = O.init_weights(cfg)
    rs = np.random.RandomState(0)
    z = rs.normal(size=(2, 1, cfg.z_dim))
    e, g, pure = O.energy_and_grad(cfg, params, z, xi_raw=0.2)
    assert e.shape == (2, 1) and g.shape == z.shape and pure.shape == (2, 1)
    for idx in [(0, 0, 0), (1, 0, cfg.z_dim - 1), (0, 0, cfg.z_dim // 2)]:
        h = 1e-5
        zp, zm = z.copy(), z.copy()
        zp[idx] += h
        zm[idx] -= h
        ep = O.energy_and_grad(cfg, params, zp, xi_raw=0.2)[0]
        em = O.energy_and_grad(cfg, params, zm, xi_raw=0.2)[0]
        fd = (ep[idx[:2]] - em[idx[:2]]) / (2 * h)
        assert abs(fd - g[idx]) < 1e-4 * max(1.0, abs(fd)), (idx, fd, g[idx])


def test_mala_step_properties():
    """alpha -> tiny: proposals are accepted (ratio ~ 1); rejected chains keep z exactly (dmm_cifar10.py:224-226)."""
    cfg = O.OracleConfig('B', dataset='cifar10', z_dim=16)
    params = O.init_weights(cfg)
    rs = np.random.RandomState(1)
    z = rs.normal(size=(3, 1, 16))
    eps = rs.normal(size=z.shape)
    z1, e, pure, ratio = O.mala_step(cfg, params, z, eps, np.zeros((3, 1)), 1e-9)
    assert np.all(ratio > 0.99) and np.all(z1 != z)
    z2, _, _, ratio2 = O.mala_step(cfg, params, z, eps, np.ones((3, 1)) * 2, 1e-9)     # u > 1: never accept
    assert np.array_equal(z2, z)
    out = O.mala_sample(cfg, params, z, rs.normal(size=(4, 3, 1, 16)), rs.uniform(size=(4, 3, 1)), alpha=1e-4)
    assert out['history_z'].shape == (4, 3, 1, 16) and np.array_equal(out['history_z'][0], z)


def test_eval_batch_self_consistency_and_weight_schema():
    from apvae.model import ExpConfig, weight_schema
    for arch, kw in (('A', dict(z_dim=128)), ('A', dict()), ('B', dict(dataset='cifar10')), ('B', dict(dataset='celebA')), ('M', dict())):
        ocfg = O.OracleConfig(arch, **kw)
        ours = dict(weight_schema(ExpConfig(arch, dataset=kw.get('dataset'), z_dim=ocfg.z_dim)))
        theirs = {k: tuple(v) for k, v in O.weight_shapes(ocfg).items()}
        assert ours == theirs                        # the C library's TF variable schema == the oracle's
    cfg = O.OracleConfig('M', z_dim=8)
    params = O.init_weights(cfg)
    rs = np.random.RandomState(0)
    x = (rs.uniform(size=(2, 28, 28, 1)) < 0.5).astype(np.float64)
    out = O.eval_batch(cfg, params, x, x, np.clip(rs.normal(size=(3, 2, 8)), -2, 2))
    assert np.all(out['ele_test_lb'] <= out['ele_test_ll'] + 1e-9)       # Jensen: ELBO <= IS-LL
    assert np.all(out['ele_adv_test_lb'] <= out['ele_adv_test_ll'] + 1e-9)


def test_planar_flow_follows_the_reference_test_and_its_jacobian():
    """tests/layers/flows/test_planar_nf.py:16-71 restated: u_hat against the closed form the reference test uses, the
    transform against the test's own numpy `transform` on x = arange(30).reshape(2, 3, 5); plus two checks the
    reference does not have: the invertibility condition w.u_hat >= -1 and log_det against the determinant of a
    finite-difference Jacobian."""
    rs = np.random.RandomState(1234)
    n = 5
    w, u, b = rs.normal(size=(1, n)), rs.normal(size=(1, n)), rs.normal(size=(1,))
    m = lambda a: -1 + np.log(1 + np.exp(a))
    wu = np.dot(w, u.T)
    u_hat = O.planar_flow_u_hat(w, u)
    np.testing.assert_allclose(u + w * (m(wu) - wu) / np.sum(w ** 2), u_hat)           # test_u_hat
    assert np.dot(w, u_hat.T).item() >= -1.0

    def transform(x, u, w, b):                                                          # test_transform's reference
        wxb = np.dot(x, w.T) + b
        y = x + u * np.tanh(wxb)
        tanh_wxb = np.tanh(wxb)
        phi = (1 - tanh_wxb ** 2) * w
        log_det = np.log(np.abs(1 + np.dot(phi, u.T)))
        return y, np.squeeze(log_det, -1)

    x = np.arange(30, dtype=np.float64).reshape([2, 3, n]) / 10.0
    y, log_det = transform(x, u_hat, w, b)
    y2, log_det2 = O.planar_flow_transform(x, w, b, u)
    np.testing.assert_allclose(y2, y, rtol=1e-12)
    np.testing.assert_allclose(log_det2, log_det, rtol=1e-12)
    # finite-difference Jacobian of one row
    x0 = x[1, 2]
    J = np.zeros((n, n))
    for j in range(n):
        e = np.zeros(n); e[j] = 1e-6
        J[:, j] = (O.planar_flow_transform(x0 + e, w, b, u)[0] - O.planar_flow_transform(x0 - e, w, b, u)[0]) / 2e-6
    assert abs(np.log(abs(np.linalg.det(J))) - log_det2[1, 2]) < 1e-6
    # sequential flow: log-dets add, variable scopes posterior_flow/_i/
    params = {}
    for i in range(3):
        params['posterior_flow/_%d/w' % i] = rs.normal(size=(1, n))
        params['posterior_flow/_%d/u' % i] = rs.normal(size=(1, n))
        params['posterior_flow/_%d/b' % i] = rs.normal(size=(1,))
    ys, lds = O.planar_flows_transform(x, params, 3)
    yy, ld = x, 0.0
    for i in range(3):
        yy, l1 = O.planar_flow_transform(yy, params['posterior_flow/_%d/w' % i], params['posterior_flow/_%d/b' % i],
                                         params['posterior_flow/_%d/u' % i])
        ld = ld + l1
    np.testing.assert_allclose(ys, yy)
    np.testing.assert_allclose(lds, ld)


def test_flow_posterior_in_the_evaluation_graph_and_schema():
    """use_flow = True (dmm_bernouli_fashion.py:468-474): q(z|x) = FlowDistribution(TruncatedNormal, planar flows);
    log q = log q0(z0) - sum log_det; the C library's schema lists posterior_flow/_i/{w,b,u}."""
    from apvae.model import ExpConfig, weight_schema
    cfg = O.OracleConfig('M', z_dim=8, use_flow=True, nf_layers=3)
    ours = dict(weight_schema(ExpConfig('M', z_dim=8, use_flow=True, nf_layers=3)))
    assert ours == {k: tuple(v) for k, v in O.weight_shapes(cfg).items()}
    assert ours['posterior_flow/_2/w'] == (1, 8) and ours['posterior_flow/_0/b'] == (1,)
    one = dict(weight_schema(ExpConfig('M', z_dim=8, use_flow=True, nf_layers=1)))
    assert 'posterior_flow/w' in one and 'posterior_flow/_0/w' not in one
    assert not any(k.startswith('posterior_flow') for k, _ in weight_schema(ExpConfig('M', z_dim=8)))
    params = O.init_weights(cfg)
    rs = np.random.RandomState(0)
    x = (rs.uniform(size=(2, 28, 28, 1)) < 0.5).astype(np.float64)
    eps = np.clip(rs.normal(size=(3, 2, 8)), -2, 2)
    out = O.eval_batch(cfg, params, x, x, eps)
    base = O.eval_batch(O.OracleConfig('M', z_dim=8), params, x, x, eps)
    z, ld = O.planar_flows_transform(base['z'], params, 3)
    np.testing.assert_allclose(out['z'], z)
    np.testing.assert_allclose(out['log_q'], base['log_q'] - ld)
    assert np.abs(out['z'] - base['z']).max() > 1e-2                                     # the synthetic flow is not the identity
    assert np.all(out['ele_test_lb'] <= out['ele_test_ll'] + 1e-9)


def test_latent_gan_schema_and_shapes():
    from apvae.model import ExpConfig, weight_schema
    for arch, kw in (('A', dict(z_dim=128)), ('A', dict()), ('B', dict(dataset='cifar10', z_dim=64)), ('B', dict(dataset='celebA'))):
        ocfg = O.OracleConfig(arch, latent_gan=True, **kw)
        ours = dict(weight_schema(ExpConfig(arch, dataset=kw.get('dataset'), z_dim=ocfg.z_dim, latent_gan=True)))
        assert ours == {k: tuple(v) for k, v in O.weight_shapes(ocfg).items()}
        assert ours['G_omega/level_0/kernel'] == (128, 256 * 16) and ours['G_omega/level_3/shortcut/kernel'] == (1, 1, 128, 256)
    ocfg = O.OracleConfig('B', dataset='cifar10', z_dim=64, latent_gan=True)
    z0 = O.G_omega(ocfg, O.init_weights(ocfg), np.random.RandomState(0).normal(size=(2, 3, 128)))
    assert z0.shape == (2, 3, 64) and np.isfinite(z0).all()
    # arch M: six weight-normalised dense layers (kernel, kernel-shaped scale, bias)
    ocfg = O.OracleConfig('M', latent_gan=True)
    ours = dict(weight_schema(ExpConfig('M', z_dim=ocfg.z_dim, latent_gan=True)))
    assert ours == {k: tuple(v) for k, v in O.weight_shapes(ocfg).items()}
    assert ours['G_omega/dense/weight_norm/scale'] == (40, 40) and ours['G_omega/dense_5/kernel'] == (40, ocfg.z_dim)


def test_weight_norm_matches_the_reference_tests_numpy_statement():
    """tests/layers/normalization/test_weight_norm.py:23-53 (payload arange(120) as a (2,3,4,5) kernel, axis=-1): the
    expected values there are plain numpy; arch M's G_omega (dmm_bernouli_fashion.py:526-543) uses exactly this on its
    dense kernels."""
    kernel = np.reshape(np.arange(120, dtype=np.float64), (2, 3, 4, 5))
    scale = np.random.RandomState(0).normal(size=kernel.shape)
    want = kernel / np.sqrt(np.sum(kernel ** 2, axis=(0, 1, 2), keepdims=True))
    np.testing.assert_allclose(O.weight_norm(kernel), want, rtol=1e-12)
    np.testing.assert_allclose(O.weight_norm(kernel, scale), scale * want, rtol=1e-12)
    # dense kernels [in, out]: reduce over the input axis; an all-zero column hits the epsilon floor instead of dividing by 0
    k2 = kernel.reshape(24, 5).copy()
    k2[:, 2] = 0.0
    w = O.weight_norm(k2, scale.reshape(24, 5))
    assert np.all(w[:, 2] == 0.0) and np.all(np.isfinite(w))
    np.testing.assert_allclose(np.sum(O.weight_norm(k2)[:, [0, 1, 3, 4]] ** 2, axis=0), 1.0, rtol=1e-12)


def test_arch_m_latent_gan_oracle_shapes():
    cfg = O.OracleConfig('M', latent_gan=True)
    params = O.init_weights(cfg)
    z = np.random.RandomState(1).normal(size=(5, cfg.z_dim))
    out = O.G_omega(cfg, params, z)
    assert out.shape == (5, cfg.z_dim) and np.all(np.isfinite(out))
    # every layer ends in leaky_relu (the last included): negative outputs are 0.2 x the pre-activation
    pre = 'G_omega/dense_5/'
    h = z
    for i in range(5):
        p = 'G_omega/dense%s/' % ('_%d' % i if i else '')
        h = O.leaky_relu(h @ O.weight_norm(params[p + 'kernel'], params[p + 'weight_norm/scale']) + params[p + 'bias'])
    last = h @ O.weight_norm(params[pre + 'kernel'], params[pre + 'weight_norm/scale']) + params[pre + 'bias']
    np.testing.assert_allclose(out, np.where(last > 0, last, 0.2 * last), rtol=1e-12)


@pytest.mark.parametrize('k,s', KS)
def test_torch_conv_backend_matches_the_numpy_one(k, s):
    """oracle.set_conv_backend('torch') (the CPU baseline's oneDNN path) computes the same SAME conv / transposed conv."""
    rs = np.random.RandomState(k * 10 + s + 2)
    x = rs.normal(size=(2, 8, 10, 3))
    kern = rs.normal(size=(k, k, 3, 4))
    kd = rs.normal(size=(k, k, 4, 3))
    bias = rs.normal(size=4)
    want_c, want_d = O.conv2d(x, kern, bias, s), O.deconv2d(x, kd, bias, s)
    old = O.set_conv_backend('torch')
    try:
        assert O.get_conv_backend() == 'torch'
        np.testing.assert_allclose(O.conv2d(x, kern, bias, s), want_c, rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(O.deconv2d(x, kd, bias, s), want_d, rtol=1e-10, atol=1e-10)
    finally:
        O.set_conv_backend(old)
    with pytest.raises(ValueError):
        O.set_conv_backend('cudnn')


def test_torch_conv_backend_whole_graph():
    """The whole test graph (arch B, CIFAR shape) through both conv backends, float32 as in the reference."""
    cfg = O.OracleConfig('B', dataset='cifar10', z_dim=16)
    params = O.init_weights(cfg, dtype=np.float32)
    rs = np.random.RandomState(0)
    x = ((rs.randint(0, 256, size=(2, 32, 32, 3)) - 127.5) / 128.0).astype(np.float32)
    eps = rs.normal(size=(3, 2, 16)).astype(np.float32)
    a = O.eval_batch(cfg, params, x, x, eps)
    old = O.set_conv_backend('torch')
    try:
        b = O.eval_batch(cfg, params, x, x, eps)
    finally:
        O.set_conv_backend(old)
    for key in ('ele_test_ll', 'ele_adv_test_ll', 'd_g', 'ele_test_mse'):
        np.testing.assert_allclose(np.asarray(a[key], np.float64), np.asarray(b[key], np.float64), rtol=2e-4, atol=2e-4)


def test_test_mse_definition():
    """test_mse (dmm_cifar10.py:889-895): squared difference of the ROUNDED grey levels, summed over pixels, min over the
    sample axis, mean over images; tf.round = round half to even."""
    cfg = O.OracleConfig('A', z_dim=64)
    x = np.array([[-1.0, 0.0], [0.5, 0.00390625]]).reshape(1, 2, 2, 1)              # grey levels -0.5, 127.5, 191.5, 128
    xm = np.stack([x, x + 2.0 / 128.0, x * 0])                                        # K = 3: exact, two levels off, zero
    se = O.rounded_pixel_sq_err(cfg, x, xm[:, None].reshape(3, 1, 2, 2, 1))
    lv = np.round(x * 128 + 127.5).ravel()
    assert lv.tolist() == [-0.0, 128.0, 192.0, 128.0]                                 # 127.5 -> 128, 191.5 -> 192 (half to even)
    assert se.shape == (3, 1) and se[0, 0] == 0 and se[1, 0] == 4 * 2 ** 2
    assert se[2, 0] == np.sum((np.round(127.5) - lv) ** 2)
    cfg_m = O.OracleConfig('M')
    xb = np.array([0.0, 1.0, 1.0, 0.0]).reshape(1, 2, 2, 1)
    pm = np.array([-0.3, 0.999, 0.5, 0.2]).reshape(1, 1, 2, 2, 1)                      # clip(-0.3) -> 1e-7 -> level 0
    se = O.rounded_pixel_sq_err(cfg_m, xb, pm)
    assert se[0, 0] == (0 - 0) ** 2 + (255 - 255) ** 2 + (np.round(127.5) - 255) ** 2 + (51 - 0) ** 2
