"""GPU tests of the evaluation driver and of BASELINE-size runs through size-independent properties.

* Evaluator.run (spt.Evaluator.run, tfsnippet/trainer/evaluator.py:193-259) on ragged mini-batches: the
  size-weighted mean of batch means == the mean over images of the per-image rows, and both agree with the
  float64 oracle within the north_star tolerance; the log-Z loop (dmm_cifar10.py:1215-1227) against the oracle's
  two-level logsumexp.
* K = 1000 (BASELINE configs[1]) on a few images, where the oracle would take minutes: chunking invariance
  (7400-sample chunks vs ragged 999-sample chunks, which also swaps the kernel that runs the 8x8 layers),
  Jensen's bound (IS log-likelihood >= ELBO per image), and duplicate images with duplicate noise giving
  identical rows whatever their position in the batch.
"""
import math

import numpy as np
import pytest

import oracle as O
from helpers import synthetic_images

pytestmark = pytest.mark.gpu

NLL_RTOL = 1e-3


def _model(max_batch, max_chunk, **kw):
    from apvae.model import ExpConfig, Model
    ocfg = O.OracleConfig('A', z_dim=128)
    params = O.init_weights(ocfg)
    cfg = ExpConfig('A', z_dim=128, max_batch=max_batch, max_chunk=max_chunk, **kw)
    return ocfg, params, Model(cfg, params)


def test_evaluator_ragged_batches_match_per_image_means_and_oracle():
    from apvae.evaluation import Evaluator
    ocfg, params, m = _model(max_batch=4, max_chunk=8)
    N, K, bs = 10, 4, 4                                   # batches of 4, 4, 2
    x = synthetic_images((N,) + tuple(ocfg.x_shape), seed=0).astype(np.float64)
    eps = np.random.RandomState(3).normal(size=(K, N, ocfg.z_dim))
    xi_raw, beta, log_Z = 0.3, -0.5, 2.5
    m.set_scalars(xi_raw, beta, log_Z)
    ev = Evaluator(m, n_z=K, batch_size=bs)
    got = ev.run(x, x_origin=x, eps=eps)
    # (1) against the same library's per-image rows: sums of per-image values / N
    rows = {k: [] for k in ('ele_test_ll', 'ele_test_lb', 'ele_adv_test_ll', 'ele_adv_test_lb', 'ele_test_recon',
                            'ele_reconstruct_energy', 'ele_real_energy')}
    for b0 in range(0, N, bs):
        o = m.is_nll(x[b0:b0 + bs], K, x_origin=x[b0:b0 + bs], eps=eps[:, b0:b0 + bs])
        for k in rows:
            rows[k].append(o[k].cpu().numpy().astype(np.float64))
    mean = {k: np.concatenate(v).mean() for k, v in rows.items()}
    pairs = (('test_nll', -mean['ele_test_ll']), ('test_lb', mean['ele_test_lb']), ('adv_test_nll', -mean['ele_adv_test_ll']),
             ('adv_test_lb', mean['ele_adv_test_lb']), ('test_recon', mean['ele_test_recon']),
             ('reconstruct_energy', mean['ele_reconstruct_energy']), ('real_energy', mean['ele_real_energy']))
    for name, want in pairs:
        assert abs(got[name] - want) <= 1e-5 * max(1.0, abs(want)), (name, got[name], want)
    # (2) against the float64 oracle (np.average of batch means weighted by batch size == mean over images)
    want = O.eval_batch(ocfg, params, x, x, eps, xi_raw=xi_raw, beta=beta, log_Z=log_Z)
    assert abs(got['test_nll'] + want['ele_test_ll'].mean()) < NLL_RTOL * abs(want['ele_test_ll'].mean())
    assert abs(got['adv_test_nll'] + want['ele_adv_test_ll'].mean()) < NLL_RTOL * abs(want['ele_adv_test_ll'].mean())
    assert abs(got['test_lb'] - want['ele_test_lb'].mean()) < NLL_RTOL * abs(want['ele_test_lb'].mean())
    assert abs(got['real_energy'] - want['ele_real_energy'].mean()) < 1e-2
    # (3) the log-Z loop: logsumexp over batches of the per-batch log-mean-exp terms - log(#batches) + log(len_train)
    lz = ev.estimate_log_Z(x, x_origin=x, eps=eps, set_model=False)
    w_xi = ocfg.pull_back_energy_weight / (1 + math.exp(-xi_raw))
    per_batch = []
    for b0 in range(0, N, bs):
        sl = slice(b0, b0 + bs)
        t = -want['energy'][:, sl] - want['log_q'][:, sl]                     # another_log_Z_compute_op, dmm_cifar10.py:922-924
        per_batch.append(O.log_mean_exp(t.reshape(-1)))
    lz_o = O.log_sum_exp(np.asarray(per_batch)) - math.log(len(per_batch)) + math.log(m.config.len_train)
    assert abs(lz - lz_o) < 1e-3 * abs(lz_o) + w_xi * 1e-2, (lz, lz_o)        # energies within 1e-2, times the weight
    m.close()


def test_full_size_k1000_properties():
    B, K = 4, 1000
    ocfg, params, big = _model(max_batch=B, max_chunk=7400)
    x = synthetic_images((B,) + tuple(ocfg.x_shape), seed=0).astype(np.float32)
    x[3] = x[0]                                                    # duplicate image ...
    eps = np.random.RandomState(3).normal(size=(K, B, ocfg.z_dim)).astype(np.float32)
    eps[:, 3] = eps[:, 0]                                          # ... with duplicate noise
    keys = ('ele_test_ll', 'ele_test_lb', 'ele_adv_test_ll', 'ele_adv_test_lb', 'ele_test_recon', 'ele_reconstruct_energy')
    big.set_scalars(0.3, -0.5, 2.5)
    a = {k: v.cpu().numpy().astype(np.float64) for k, v in big.is_nll(x, K, x_origin=x, eps=eps).items() if k in keys}
    big.close()
    # Jensen: log mean exp >= mean, for the Gaussian-prior and the energy-prior estimators
    assert np.all(a['ele_test_ll'] >= a['ele_test_lb'] - 1e-3)
    assert np.all(a['ele_adv_test_ll'] >= a['ele_adv_test_lb'] - 1e-3)
    assert np.all(np.isfinite(np.stack(list(a.values()))))
    # duplicate (image, noise) -> the same rows whatever the position in the batch (every sample's arithmetic is
    # independent of its neighbours)
    for k in keys:
        assert abs(a[k][3] - a[k][0]) <= 1e-6 * abs(a[k][0]), (k, a[k][0], a[k][3])
    # chunking invariance: 999-sample chunks (ragged: 4000 = 4 * 999 + 4; odd capacity -> the 8x8 layers run on the
    # tap kernel instead of the patch kernel's image pairs) against one 4000-sample launch per op
    ocfg, params, small = _model(max_batch=B, max_chunk=999)
    small.set_scalars(0.3, -0.5, 2.5)
    b = {k: v.cpu().numpy().astype(np.float64) for k, v in small.is_nll(x, K, x_origin=x, eps=eps).items() if k in keys}
    small.close()
    for k in ('ele_test_ll', 'ele_test_lb', 'ele_adv_test_ll', 'ele_adv_test_lb', 'ele_test_recon'):
        rel = np.abs(a[k] - b[k]) / np.abs(a[k])
        assert rel.max() < 2e-4, (k, a[k], b[k])                   # same 16-bit rounding points, different fp32 summation order
    assert np.abs(a['ele_reconstruct_energy'] - b['ele_reconstruct_energy']).max() < 1e-2


def test_benchmarked_shape_against_the_oracle():
    """The configuration bench.py times (BASELINE configs[1]: arch A, z_dim 128, K = 1000, 37 images per step, chunks of
    7400 samples) against the oracle: the library evaluates the whole 37-image batch with fed-in noise; the oracle
    (float32 "as-reference" mode, ~1-2 minutes of CPU) evaluates two of its images — images are independent, so those
    per-image rows must agree.  north_star tolerances: per-image NLL 1e-3 relative, energies 1e-2 absolute."""
    from apvae.model import ExpConfig, Model
    B, K = 37, 1000
    ocfg = O.OracleConfig('A', z_dim=128)
    params = O.init_weights(ocfg)
    m = Model(ExpConfig('A', z_dim=128, max_batch=B, max_chunk=7400), params)
    x = synthetic_images((B,) + tuple(ocfg.x_shape), seed=0).astype(np.float32)
    eps = np.random.RandomState(3).normal(size=(K, B, ocfg.z_dim)).astype(np.float32)
    xi_raw, beta, log_Z = 0.0, 0.0, 1.0                             # the bench's scalars
    m.set_scalars(xi_raw, beta, log_Z)
    got = m.is_nll(x, K, x_origin=x, eps=eps, per_sample=True)
    g = {k: v.cpu().numpy().astype(np.float64) for k, v in got.items()}
    m.close()
    p32 = {k: v.astype(np.float32) for k, v in params.items()}
    old_backend = O.set_conv_backend('torch')       # torch-CPU convs (cross-checked against the numpy ones on the CPU): 4x faster
    try:
        _check_images_against_oracle(ocfg, p32, x, eps, g, K, xi_raw, beta, log_Z)
    finally:
        O.set_conv_backend(old_backend)


def _check_images_against_oracle(ocfg, p32, x, eps, g, K, xi_raw, beta, log_Z):
    for b in (0, 36):
        # the oracle in slices of 100 samples (the nets treat samples independently; bounds its memory), then the
        # reference's reductions over all K (variational/evaluation.py:9-39, objectives.py:9-42)
        parts = [O.eval_batch(ocfg, p32, x[b:b + 1], x[b:b + 1], eps[k0:k0 + 100, b:b + 1], xi_raw=xi_raw, beta=beta, log_Z=log_Z)
                 for k0 in range(0, K, 100)]
        cat = lambda key: np.concatenate([p[key] for p in parts], axis=0).astype(np.float64)
        log_px, log_q, log_pn, d_g, energy, sq = (cat(k) for k in ('log_px', 'log_q', 'log_pn', 'd_g', 'energy', 'sq_err'))
        lja = log_px - energy - max(log_Z, -20.0)
        want = {'ele_test_ll': O.importance_sampling_log_likelihood(log_px + log_pn, log_q, 0),
                'ele_test_lb': O.elbo_objective(log_px + log_pn, log_q, 0),
                'ele_adv_test_ll': O.importance_sampling_log_likelihood(lja, log_q, 0),
                'ele_adv_test_lb': O.elbo_objective(lja, log_q, 0)}
        for name in want:
            rel = abs(g[name][b] - float(want[name][0])) / abs(float(want[name][0]))
            assert rel < NLL_RTOL, (b, name, g[name][b], want[name][0])
        d_err = np.abs(g['d_g'][:, b] - d_g[:, 0])
        assert d_err.max() < 1e-2, (b, d_err.max(), d_err.mean())           # all 1000 samples of the image
        assert abs(g['ele_real_energy'][b] - float(parts[0]['ele_real_energy'][0])) < 1e-2
        lp_err = np.abs(g['log_px'][:, b] - log_px[:, 0]) / np.abs(log_px[:, 0])
        assert lp_err.max() < NLL_RTOL, (b, lp_err.max())
        assert abs(g['ele_test_mse'][b] - sq[:, 0].min()) < 2e-2 * sq[:, 0].min()


def test_k1_and_bad_shapes_are_rejected():
    """_require_multi_samples (variational/utils.py:1-5 via evaluation.py:25) and InputSpec shape errors."""
    from apvae._lib import ApvaeError
    ocfg, params, m = _model(max_batch=2, max_chunk=8)
    x = synthetic_images((2,) + tuple(ocfg.x_shape), seed=0).astype(np.float32)
    with pytest.raises((ValueError, ApvaeError)):
        m.is_nll(x, 0)
    with pytest.raises((ValueError, ApvaeError)):
        m.is_nll(x[:, :16], 3)                                      # wrong image height
    with pytest.raises((ValueError, ApvaeError)):
        m.is_nll(np.concatenate([x, x, x]), 3)                      # batch beyond max_batch
    m.close()
