"""Test-graph metrics, log-Z, MALA sampler and Evaluator averaging (test oracle).

Reference: experiment/dmm/dmm_cifar10.py:97-252 (EnergyDistribution, MALA),
:880-927 (test graph), :1215-1227 (log Z), :976-990 (image quantisation);
tfsnippet/trainer/evaluator.py:17-31,193-259.
"""
import numpy as np

from .dists import (normal_log_prob, truncated_normal_log_prob,
                    std_normal_log_prob, discretized_logistic_log_prob,
                    bernoulli_log_prob_from_mean, log_mean_exp, log_sum_exp,
                    importance_sampling_log_likelihood, elbo_objective,
                    maybe_clip_value, sigmoid, planar_flows_transform)
from .nets import q_net, G_theta, D_psi, G_theta_input_grad, D_psi_input_grad

__all__ = ['log_px_given_z', 'rounded_pixel_sq_err', 'energy_terms', 'eval_batch', 'prior_branch',
           'log_Z_batch', 'combine_log_Z', 'energy_and_grad', 'mala_step',
           'mala_sample', 'sample_images', 'evaluator_average', 'METRIC_NAMES']

METRIC_NAMES = ['test_nll', 'test_lb', 'adv_test_nll', 'adv_test_lb',
                'reconstruct_energy', 'real_energy', 'pd_energy', 'pn_energy',
                'test_recon', 'kl_adv_and_gaussian', 'test_mse']     # dmm_cifar10.py:1131-1136


def log_px_given_z(cfg, params, x, z, beta=0.0):
    """p_net's x node (dmm_cifar10.py:611-619; arch M: dmm_bernouli_fashion.py:619-626).
    x [B,H,W,C] broadcasts over the leading sample axis of z [K,B,d].
    Returns (log p(x|z) [K,B], x_mean [K,B,H,W,C])."""
    if cfg.arch == 'M':
        x_mean = G_theta(cfg, params, z)
        return bernoulli_log_prob_from_mean(x, x_mean, 3), x_mean
    x_mean, x_logstd = G_theta(cfg, params, z, return_std=True)
    log_scale = maybe_clip_value(
        np.asarray(beta, dtype=x_mean.dtype) if cfg.uniform_scale else x_logstd,
        min_val=cfg.epsilon)
    lp = discretized_logistic_log_prob(
        x, x_mean, log_scale, bin_size=2.0 / 256.0, min_val=-1.0 + 1.0 / 256.0,
        max_val=1.0 - 1.0 / 256.0, epsilon=1e-10, group_ndims=3)
    return lp, x_mean


def rounded_pixel_sq_err(cfg, x, x_mean):
    """The summand of test_mse (dmm_cifar10.py:889-891): sum over H, W, C of
    (round(x_mean * 128 + 127.5) - round(x * 128 + 127.5))^2 with tf.round = round half to even (np.round);
    arch M (dmm_bernouli_fashion.py:893-896): round(mean * 255) - round(x * 255), where distribution.mean of the
    Bernoulli is sigmoid(logits) = the clipped G_theta output (:619-626)."""
    if cfg.arch == 'M':
        p = np.clip(x_mean, 1e-7, 1.0 - 1e-7)
        d = np.round(p * 255.0) - np.round(x * 255.0)
    else:
        d = np.round(x_mean * 128.0 + 127.5) - np.round(x * 128.0 + 127.5)
    return np.sum(np.square(d), axis=(-1, -2, -3))


def energy_terms(cfg, z, d_of_g, xi_raw, log_Z):
    """EnergyDistribution.log_prob attributes (dmm_cifar10.py:164-177):
    energy = w * D(G(z)) * xi + 0.5 sum z^2 ; log_energy_prob = -energy -
    max(log_Z, -20) (:135) ; xi = sigmoid(xi_raw) (:607)."""
    xi = sigmoid(np.asarray(xi_raw, dtype=z.dtype))
    energy = cfg.pull_back_energy_weight * d_of_g * xi + 0.5 * np.sum(np.square(z), axis=-1)
    log_energy_prob = -energy - np.maximum(log_Z, -20.0)
    return energy, log_energy_prob


def eval_batch(cfg, params, x, x_origin, eps, xi_raw=0.0, beta=0.0, log_Z=1.0):
    """One Evaluator mini-batch of the test graph (dmm_cifar10.py:880-927),
    with the duplicate G_theta pass removed (the graph evaluates G(z) at :611
    and again inside :170 — same values).

    x        [B,H,W,C]  likelihood target (input_x)
    x_origin [B,H,W,C]  q_net / real_energy input for arch M
                         (dmm_bernouli_fashion.py:888-891); A/B feed x twice.
    eps      [K,B,d]    N(0,1) (resampled into [-2,2] for TruncatedNormal)
    Returns dict with per-sample [K,B], per-image [B] and batch scalars."""
    q_in = x_origin if cfg.arch == 'M' else x
    mu, logstd = q_net(cfg, params, q_in)
    z = mu + np.exp(logstd) * eps                                   # zd.Normal sample
    if cfg.use_truncated:
        log_q = truncated_normal_log_prob(z, mu, logstd, 1)
    else:
        log_q = normal_log_prob(z, mu, logstd, 1)
    if cfg.use_flow:
        # FlowDistribution(z_distribution, posterior_flow) (dmm_bernouli_fashion.py:468-474):
        # y = flow(x), log q(y) = log q0(x) - log_det (distributions/flow.py:118-150)
        z, log_det = planar_flows_transform(z, params, cfg.nf_layers)
        log_q = log_q - log_det
    log_px, x_mean = log_px_given_z(cfg, params, x, z, beta)
    log_pn = std_normal_log_prob(z)
    d_g = D_psi(cfg, params, x_mean)                                # pure_energy, :175
    energy, log_energy_prob = energy_terms(cfg, z, d_g, xi_raw, log_Z)
    out = dict(mu=mu, logstd=logstd, z=z, log_q=log_q, log_px=log_px,
               log_pn=log_pn, d_g=d_g, energy=energy, x_mean=x_mean)
    # :896-899 plain prior
    lj = log_px + log_pn
    out['ele_test_ll'] = importance_sampling_log_likelihood(lj, log_q, 0)
    out['ele_test_lb'] = elbo_objective(lj, log_q, 0)
    # :901-911 adversarial prior
    lja = log_px + log_energy_prob
    out['ele_adv_test_ll'] = importance_sampling_log_likelihood(lja, log_q, 0)
    out['ele_adv_test_lb'] = elbo_objective(lja, log_q, 0)
    out['test_nll'] = -np.mean(out['ele_test_ll'])
    out['test_lb'] = np.mean(out['ele_test_lb'])
    out['adv_test_nll'] = -np.mean(out['ele_adv_test_ll'])
    out['adv_test_lb'] = np.mean(out['ele_adv_test_lb'])
    out['test_recon'] = np.mean(log_px)                             # :886-888
    out['sq_err'] = rounded_pixel_sq_err(cfg, x, x_mean)            # [K,B]
    out['ele_test_mse'] = np.min(out['sq_err'], axis=0)             # :892 reduce_min over the sample axis
    out['test_mse'] = np.mean(out['ele_test_mse'])                  # :893-895 (test_x_samples = 1, :76)
    out['ele_real_energy'] = D_psi(cfg, params, x_origin)           # :913
    out['real_energy'] = np.mean(out['ele_real_energy'])
    out['reconstruct_energy'] = np.mean(d_g)                        # :914
    # :922-924 another_log_Z_compute_op (over ALL [K,B] elements)
    out['another_log_Z'] = log_mean_exp(-energy - log_q + np.log(cfg.len_train))
    return out


def prior_branch(cfg, params, eps_p, xi_raw=0.0, log_Z=1.0):
    """test_pn_net = p_net(n_z=test_n_pz) with prior samples z ~ N(0,I)
    (dmm_cifar10.py:883,915-927).  eps_p [K_p, 1, d]."""
    z = eps_p
    x_mean = G_theta(cfg, params, z)
    d_g = D_psi(cfg, params, x_mean)
    log_pn = std_normal_log_prob(z)
    energy, log_energy_prob = energy_terms(cfg, z, d_g, xi_raw, log_Z)
    return dict(
        pn_energy=np.mean(d_g),                                     # :918
        pd_energy=np.mean(d_g * np.exp(log_energy_prob - log_pn)),  # :915-917
        kl_adv_and_gaussian=np.mean(log_pn - log_energy_prob),      # :925-927
        log_Z_compute_op=log_mean_exp(-energy - log_pn),            # :919-920
        d_g=d_g)


def log_Z_batch(cfg, energy, log_q):
    """another_log_Z_compute_op for one train batch (dmm_cifar10.py:922-924)."""
    return log_mean_exp(-energy - log_q + np.log(cfg.len_train))


def combine_log_Z(log_Z_list):
    """Host combine (dmm_cifar10.py:1221-1222): logsumexp(list) - log(len)."""
    v = np.asarray(log_Z_list, dtype=np.float64)
    return log_sum_exp(v) - np.log(len(v))


def energy_and_grad(cfg, params, z, xi_raw=0.0):
    """get_sgld_proposal's energy, pure energy and tf.gradients(energy,[z])
    (dmm_cifar10.py:241-247): E = w*xi*D(G(z)) + 0.5|z|^2."""
    xi = sigmoid(np.asarray(xi_raw, dtype=z.dtype))
    x_mean = G_theta(cfg, params, z)
    d, g_x = D_psi_input_grad(cfg, params, x_mean)
    _, g_z = G_theta_input_grad(cfg, params, z, g_x)
    w = cfg.pull_back_energy_weight * xi
    energy = w * d + 0.5 * np.sum(np.square(z), axis=-1)
    grad = w * g_z + z
    return energy, grad, d


def mala_step(cfg, params, z, eps, u, alpha, xi_raw=0.0):
    """One iteration of EnergyDistribution.sample's loop
    (dmm_cifar10.py:197-226) with the noise fed in: eps ~ N(0,1) [like z],
    u ~ U(0,1) [z.shape[:-1]].  Returns (z_next, e_z, pure_e_z, ratio)."""
    e_z, g_z, pure = energy_and_grad(cfg, params, z, xi_raw)
    z_prime = z - alpha * g_z + eps * np.sqrt(alpha * 2)            # :248-251
    e_zp, g_zp, _ = energy_and_grad(cfg, params, z_prime, xi_raw)   # :199
    log_q_zprime_z = np.sum(np.square(z_prime - z + alpha * g_z), axis=-1) * (-1.0 / (4 * alpha))
    log_q_z_zprime = np.sum(np.square(z - z_prime + alpha * g_zp), axis=-1) * (-1.0 / (4 * alpha))
    log_ratio_1 = -e_zp + e_z                                       # :213
    log_ratio_2 = log_q_z_zprime - log_q_zprime_z                   # :214
    with np.errstate(over='ignore'):
        ratio = np.clip(np.exp(log_ratio_1 + log_ratio_2), 0.0, 1.0)   # :216-218
    mask = (u < ratio).astype(z.dtype)[..., None]                   # :224-225
    z_next = z_prime * mask + z * (1 - mask)                        # :226
    return z_next, e_z, pure, ratio


def mala_sample(cfg, params, z0, eps, u, alpha=None, xi_raw=0.0):
    """`steps` MALA iterations; eps [steps, n, 1, d], u [steps, n, 1].
    Returns final z and the histories the reference records (:235-238);
    history_z[i] is z BEFORE step i."""
    alpha = cfg.smallest_step if alpha is None else alpha
    z = z0
    hz, he, hp, hr = [], [], [], []
    for i in range(eps.shape[0]):
        hz.append(z)
        z, e_z, pure, ratio = mala_step(cfg, params, z, eps[i], u[i], alpha, xi_raw)
        he.append(e_z), hp.append(pure), hr.append(ratio)
    return dict(z=z, history_z=np.stack(hz), history_e_z=np.stack(he),
                history_pure_e_z=np.stack(hp), history_ratio=np.stack(hr))


def sample_images(cfg, params, z):
    """x_plots (dmm_cifar10.py:976-977,990) + the host np.round (:1067,1240):
    uint8 = round(clip(256*G(z)/2 + 127.5, 0, 255))."""
    x_mean = G_theta(cfg, params, z)
    x = np.clip(256.0 * x_mean / 2 + 127.5, 0, 255)
    return np.round(x.reshape((-1,) + tuple(cfg.x_shape))).astype(np.uint8)


def evaluator_average(batch_values, batch_weights=None):
    """Evaluator.run's merge (trainer/evaluator.py:249-254): np.average of the
    stacked per-batch metric rows with weights = batch_arrays[0].size
    (auto_batch_weight, :17-31) or 1 when batch_weight_func is None."""
    vals = np.stack([np.asarray(v, dtype=np.float64) for v in batch_values], axis=0)
    w = None if batch_weights is None else np.asarray(batch_weights, dtype=np.float64)
    return np.average(vals, axis=0, weights=w)
