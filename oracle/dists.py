"""Distribution / reduction arithmetic of the hot path (test oracle, numpy).

Reference: tfsnippet/distributions/{discretized,univariate,utils,batch_to_value}.py,
tfsnippet/ops/misc.py, tfsnippet/variational/{evaluation,objectives}.py,
experiment/truncated_normal.py; ZhuSuan Normal/Bernoulli formulas (third party,
unpinned: git+https://github.com/thu-ml/zhusuan.git, requirements.txt:15).
"""
import numpy as np

__all__ = [
    'softplus', 'sigmoid', 'log_sum_exp', 'log_mean_exp', 'maybe_clip_value',
    'normal_log_prob', 'truncated_normal_log_prob', 'std_normal_log_prob',
    'discretized_logistic_log_prob', 'bernoulli_log_prob_from_mean',
    'importance_sampling_log_likelihood', 'elbo_objective',
    'planar_flow_u_hat', 'planar_flow_transform', 'planar_flows_transform',
    'TRUNCATED_AREA',
]

_LOG_2PI = np.log(2.0 * np.pi)
TRUNCATED_AREA = 0.95449974      # truncated_normal.py:146


def softplus(x):
    """tf.nn.softplus = log(1 + exp(x)), evaluated stably."""
    return np.logaddexp(0.0, x)


def sigmoid(x):
    """tf.sigmoid, evaluated stably (same split as the reference tests'
    safe_sigmoid, tests/distributions/test_discretized.py:11-12)."""
    x = np.asarray(x)
    e = np.exp(-np.abs(x))
    return np.where(x >= 0, 1.0 / (1.0 + e), e / (1.0 + e))


def log_sum_exp(x, axis=None, keepdims=False):
    """ops/misc.py:32-68: x_max + log(sum(exp(x - x_max)))."""
    x = np.asarray(x)
    m = np.max(x, axis=axis, keepdims=True)
    s = np.sum(np.exp(x - m), axis=axis, keepdims=keepdims)
    m_out = m if keepdims else np.squeeze(m, axis=axis)
    return m_out + np.log(s)


def log_mean_exp(x, axis=None, keepdims=False):
    """ops/misc.py:71-107: x_max + log(mean(exp(x - x_max)))."""
    x = np.asarray(x)
    m = np.max(x, axis=axis, keepdims=True)
    s = np.mean(np.exp(x - m), axis=axis, keepdims=keepdims)
    m_out = m if keepdims else np.squeeze(m, axis=axis)
    return m_out + np.log(s)


def maybe_clip_value(x, min_val=None, max_val=None):
    """ops/misc.py:110-131."""
    if min_val is not None and max_val is not None:
        return np.clip(x, min_val, max_val)
    if min_val is not None:
        return np.maximum(x, min_val)
    if max_val is not None:
        return np.minimum(x, max_val)
    return x


def _reduce_group(lp, group_ndims):
    """distributions/utils.py:8-51 reduce_group_ndims(tf.reduce_sum, ...)."""
    if group_ndims > 0:
        lp = np.sum(lp, axis=tuple(range(-group_ndims, 0)))
    return lp


def normal_log_prob(x, mean, logstd, group_ndims=0):
    """zd.Normal._log_prob through ZhuSuanDistribution.log_prob
    (distributions/univariate.py:41-47, wrapper.py:127-129):
    -0.5 log 2pi - logstd - 0.5 exp(-2 logstd) (x - mean)^2, then sum over the
    last `group_ndims` axes."""
    prec = np.exp(-2.0 * logstd)
    lp = -0.5 * _LOG_2PI - logstd - 0.5 * prec * np.square(x - mean)
    return _reduce_group(lp, group_ndims)


def truncated_normal_log_prob(x, mean, logstd, group_ndims=0):
    """experiment/truncated_normal.py:142-149: Normal log-prob minus
    log(0.95449974) PER ELEMENT, then the group reduction."""
    lp = normal_log_prob(x, mean, logstd, 0) - np.log(TRUNCATED_AREA)
    return _reduce_group(lp, group_ndims)


def std_normal_log_prob(z):
    """Normal(0, I).batch_ndims_to_value(1) (dmm_cifar10.py:601-603,
    distributions/batch_to_value.py:111-117): sum_d(-0.5 log 2pi - 0.5 z^2)."""
    return np.sum(-0.5 * _LOG_2PI - 0.5 * np.square(z), axis=-1)


def discretized_logistic_log_prob(x, mean, log_scale, bin_size, min_val=None,
                                  max_val=None, biased_edges=True,
                                  epsilon=1e-7, group_ndims=0):
    """DiscretizedLogistic.log_prob, line by line (distributions/discretized.py:198-259)."""
    x = np.asarray(x)
    inv_scale = np.exp(-log_scale)                              # :203-204
    half_bin = bin_size * 0.5                                   # :206
    half_delta = half_bin * inv_scale                           # :208
    log_delta = np.log(bin_size) - log_scale                    # :210
    x_mid = (x - mean) * inv_scale                              # :212
    x_low = x_mid - half_delta                                  # :213
    x_high = x_mid + half_delta                                 # :214
    cdf_low = sigmoid(x_low)                                    # :216
    cdf_high = sigmoid(x_high)                                  # :217
    cdf_delta = cdf_high - cdf_low                              # :223
    middle = np.where(                                          # :224-232
        cdf_delta > epsilon,
        np.log(np.maximum(cdf_delta, 1e-12)),
        x_mid + log_delta - 2.0 * softplus(x_mid))
    log_prob = middle
    given = np.broadcast_to(x, x_mid.shape)                     # :236
    if biased_edges and min_val is not None:                    # :240-245
        left_edge = min_val + half_bin
        log_prob = np.where(given < left_edge, -softplus(-x_high), log_prob)
    if biased_edges and max_val is not None:                    # :249-254
        right_edge = max_val - half_bin
        log_prob = np.where(given >= right_edge, -softplus(x_low), log_prob)
    return _reduce_group(log_prob, group_ndims)                 # :257


def bernoulli_log_prob_from_mean(x, x_mean, group_ndims=0):
    """Arch-M likelihood (dmm_bernouli_fashion.py:619-626): the tanh head
    output is clipped to [1e-7, 1-1e-7], logits = log(p) - log1p(-p), and
    zd.Bernoulli._log_prob = -sigmoid_cross_entropy_with_logits(labels=x,
    logits) = x*logits - softplus(logits)."""
    p = np.clip(x_mean, 1e-7, 1.0 - 1e-7)
    logits = np.log(p) - np.log1p(-p)
    lp = x * logits - softplus(logits)
    return _reduce_group(lp, group_ndims)


def planar_flow_u_hat(w, u):
    """u_hat = u + (m(w.u) - w.u) w / |w|^2, m(a) = -1 + log(1 + e^a)
    (layers/flows/planar_nf.py:104-106; w, u: [1, d])."""
    w, u = np.asarray(w, dtype=np.float64), np.asarray(u, dtype=np.float64)
    wu = np.dot(w, u.T)                                   # [1, 1]
    return u + (-1.0 + softplus(wu) - wu) * w / np.sum(np.square(w))


def planar_flow_transform(x, w, b, u):
    """One PlanarNormalizingFlow layer (planar_nf.py:112-135): x [..., d] ->
    (y [..., d], log_det [...]):  y = x + u_hat tanh(w.x + b),
    log_det = log|1 + ((1 - tanh^2) w) . u_hat|."""
    u_hat = planar_flow_u_hat(w, u)
    w = np.asarray(w, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    xf = x.reshape(-1, x.shape[-1])
    wxb = np.dot(xf, w.T) + np.asarray(b, dtype=np.float64)   # [?, 1]
    t = np.tanh(wxb)
    y = xf + u_hat * t
    phi = (1.0 - np.square(t)) * w                        # [?, d]
    log_det = np.log(np.abs(1.0 + np.dot(phi, u_hat.T)))  # [?, 1]
    return y.reshape(x.shape), log_det[:, 0].reshape(x.shape[:-1])


def planar_flows_transform(x, params, n_layers, scope='posterior_flow'):
    """planar_normalizing_flows(n_layers, name=scope) as a SequentialFlow
    (planar_nf.py:208-221; sequential.py sums the layers' log-dets).
    Variables scope/_i/{w,b,u}, or scope/{w,b,u} for a single layer."""
    y = np.asarray(x, dtype=np.float64)
    log_det = np.zeros(y.shape[:-1])
    for i in range(n_layers):
        pre = scope + '/' if n_layers == 1 else '%s/_%d/' % (scope, i)
        y, ld = planar_flow_transform(y, params[pre + 'w'], params[pre + 'b'], params[pre + 'u'])
        log_det = log_det + ld
    return y, log_det


def importance_sampling_log_likelihood(log_joint, latent_log_prob, axis):
    """variational/evaluation.py:9-39."""
    if axis is None:
        raise ValueError('`importance_sampling_log_likelihood` requires '
                         '`axis` to be specified.')         # variational/utils.py:1-5
    return log_mean_exp(log_joint - latent_log_prob, axis=axis)


def elbo_objective(log_joint, latent_log_prob, axis=None):
    """variational/objectives.py:9-42: mean over `axis` of
    (log_joint - latent_log_prob)."""
    obj = log_joint - latent_log_prob
    if axis is not None:
        obj = np.mean(obj, axis=axis)
    return obj
