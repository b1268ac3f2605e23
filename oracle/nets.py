"""q_net / G_theta / D_psi builders of the three script families (test oracle).

arch 'A' : experiment/dmm/dmm_cifar10.py            (conv latent, 96/192 wide)
arch 'B' : experiment/dmm/dmm_for_sampling.py       (dense latent, cifar/celebA)
arch 'M' : experiment/dmm/dmm_bernouli_fashion.py   (28x28x1 Bernoulli)

Weights live in a dict keyed by the TF variable names the reference creates
under `global_reuse` scopes, e.g.
'G_theta/level_3/residual/conv_0/kernel'.
"""
import numpy as np

from .layers import (conv2d, deconv2d, dense, resnet_conv2d_block,
                     resnet_deconv2d_block, block_input_grad,
                     conv2d_input_grad, leaky_relu)
from .dists import maybe_clip_value

__all__ = ['OracleConfig', 'weight_shapes', 'init_weights', 'q_net',
           'G_theta', 'D_psi', 'G_omega', 'weight_norm', 'G_theta_input_grad', 'D_psi_input_grad']


class OracleConfig(object):
    """The ExpConfig fields the path reads (dmm_cifar10.py:27-91,
    dmm_for_sampling.py:28-92, dmm_bernouli_fashion.py ExpConfig)."""

    def __init__(self, arch='A', z_dim=None, dataset=None, **kw):
        self.arch = arch
        if arch == 'A':
            self.x_shape = (32, 32, 3)
            self.z_dim = 1024 if z_dim is None else z_dim     # dmm_cifar10.py:29
            self.pull_back_energy_weight = 1000.0             # :46
            self.uniform_scale = False                        # :47
            self.smallest_step = 5e-5                         # :52
            self.min_logstd_of_q = -3.0                       # :85
            self.use_truncated = False
            self.len_train = 50000                            # :80
            self.dataset = 'cifar10'
        elif arch == 'B':
            self.dataset = dataset or 'celebA'                # dmm_for_sampling.py:92
            self.x_shape = (64, 64, 3) if self.dataset == 'celebA' else (32, 32, 3)
            self.z_dim = 128 if z_dim is None else z_dim      # :30
            self.pull_back_energy_weight = 1000.0
            self.uniform_scale = True                         # :48
            self.smallest_step = 5e-5
            self.min_logstd_of_q = -3.0
            self.use_truncated = False
            self.len_train = 162770 if self.dataset == 'celebA' else 50000
        elif arch == 'M':
            self.dataset = dataset or 'fashion'
            self.x_shape = (28, 28, 1)
            self.z_dim = 40 if z_dim is None else z_dim
            self.pull_back_energy_weight = 20.0
            self.uniform_scale = False
            self.smallest_step = 5e-6
            self.min_logstd_of_q = -5.0
            self.use_truncated = True
            self.len_train = 50000
        else:
            raise ValueError('unknown arch %r' % (arch,))
        self.kernel_size = 3
        self.shortcut_kernel_size = 1
        self.epsilon = -20.0
        self.initial_xi = 0.0
        self.test_n_pz = 1000
        self.latent_gan = False                               # build G_omega too (dmm_cifar10.py:503-532)
        self.use_flow = False                                 # dmm_bernouli_fashion.py:69
        self.nf_layers = 20                                   # :48
        for k, v in kw.items():
            if not hasattr(self, k):
                raise AttributeError(k)
            setattr(self, k, v)

    @property
    def likelihood(self):
        return 'bernoulli' if self.arch == 'M' else 'discretized_logistic'


# (scope, out_channels, kernel_size, strides) per network -------------------

def _q_blocks(cfg):
    if cfg.arch == 'A':         # dmm_cifar10.py:432-438
        return [('level_0', 96, 5, 1), ('level_1', 96, 3, 1), ('level_2', 192, 3, 2),
                ('level_4', 192, 3, 1), ('level_5', 192, 3, 2), ('level_7', 192, 3, 1),
                ('level_9', 192, 3, 1)]
    if cfg.arch == 'B':         # dmm_for_sampling.py:433-438
        s0 = 2 if cfg.dataset == 'celebA' else 1
        return [('level_0', 32, 5, s0), ('level_1', 32, 3, 1), ('level_2', 64, 3, 2),
                ('level_4', 64, 3, 1), ('level_5', 128, 3, 2)]
    return [('level_0', 16, 3, 1), ('level_2', 32, 3, 1), ('level_3', 32, 3, 1),   # M :444-448
            ('level_4', 32, 3, 2), ('level_6', 64, 3, 2)]


def _g_blocks(cfg):
    if cfg.arch == 'A':         # dmm_cifar10.py:481-488
        return [('level_0', 192, 3, 1), ('level_1', 192, 3, 1), ('level_2', 192, 3, 1),
                ('level_3', 192, 3, 2), ('level_5', 192, 3, 1), ('level_6', 192, 3, 1),
                ('level_8', 96, 3, 2), ('level_9', 96, 5, 1)]
    if cfg.arch == 'B':         # dmm_for_sampling.py:497-502
        s9 = 2 if cfg.dataset == 'celebA' else 1
        return [('level_3', 64, 3, 2), ('level_5', 64, 3, 1), ('level_6', 32, 3, 2),
                ('level_8', 32, 3, 1), ('level_9', 32, 5, s9)]
    return [('level_3', 64, 3, 2), ('level_5', 32, 3, 2), ('level_6', 32, 3, 1),   # M :510-514
            ('level_7', 32, 3, 1), ('level_8', 16, 3, 1)]


def _omega_blocks(cfg):          # dmm_cifar10.py:523-528 == dmm_for_sampling.py:536-541
    return [('level_1', 256, 3, 1), ('level_2', 256, 3, 1), ('level_3', 128, 3, 2),
            ('level_4', 64, 3, 1), ('level_5', 32, 3, 1), ('level_6', 16, 3, 1)]


def _omega_base_hw(cfg):
    if cfg.arch == 'A':          # dmm_cifar10.py:516-521
        return cfg.x_shape[0] // 8, cfg.x_shape[1] // 8
    return 4, 4                  # dmm_for_sampling.py:529-535


def _d_blocks(cfg):
    if cfg.arch == 'A':         # dmm_cifar10.py:551-556
        return [('level_0', 16, 3, 1), ('level_1', 32, 3, 1), ('level_2', 64, 3, 1),
                ('level_3', 128, 3, 2), ('level_4', 256, 3, 2), ('level_5', 256, 3, 2)]
    if cfg.arch == 'B':         # dmm_for_sampling.py:565-570
        s0 = 2 if cfg.dataset == 'celebA' else 1
        return [('level_0', 32, 3, s0), ('level_1', 32, 3, 1), ('level_3', 64, 3, 2),
                ('level_4', 64, 3, 1), ('level_5', 64, 3, 2)]
    return [('level_0', 16, 3, 1), ('level_2', 32, 3, 1), ('level_3', 32, 3, 1),   # M :561-565
            ('level_4', 32, 3, 2), ('level_6', 64, 3, 2)]


def _zc(cfg):
    """z_dim_channel = 16*z_dim // H // W (dmm_cifar10.py:440,478)."""
    return 16 * cfg.z_dim // cfg.x_shape[0] // cfg.x_shape[1]


def _g_base_hw(cfg):
    if cfg.arch == 'A':
        return cfg.x_shape[0] // 4, cfg.x_shape[1] // 4        # dmm_cifar10.py:480
    if cfg.arch == 'B':
        return 8, 8                                            # dmm_for_sampling.py:488-492
    return cfg.x_shape[0] // 4, cfg.x_shape[1] // 4            # M :500-506


def _block_shapes(shapes, root, blocks, cin, deconv):
    for scope, cout, k, s in blocks:
        pre = '%s/%s/' % (root, scope)
        if deconv:      # resize_at_exit=False; deconv kernels are [kh,kw,Cout,Cin]
            shapes[pre + 'residual/conv_0/kernel'] = (k, k, cout, cin)
            shapes[pre + 'residual/conv_0/bias'] = (cout,)
            shapes[pre + 'residual/conv_1/kernel'] = (k, k, cout, cout)
            shapes[pre + 'residual/conv_1/bias'] = (cout,)
            if s != 1 or cin != cout:
                shapes[pre + 'shortcut/kernel'] = (1, 1, cout, cin)
                shapes[pre + 'shortcut/bias'] = (cout,)
        else:           # resize_at_exit=True; conv kernels are [kh,kw,Cin,Cout]
            shapes[pre + 'residual/conv_0/kernel'] = (k, k, cin, cin)
            shapes[pre + 'residual/conv_0/bias'] = (cin,)
            shapes[pre + 'residual/conv_1/kernel'] = (k, k, cin, cout)
            shapes[pre + 'residual/conv_1/bias'] = (cout,)
            if s != 1 or cin != cout:
                shapes[pre + 'shortcut/kernel'] = (1, 1, cin, cout)
                shapes[pre + 'shortcut/bias'] = (cout,)
        cin = cout
    return cin


def weight_shapes(cfg):
    """Ordered {TF variable name: shape} for q_net, G_theta, D_psi
    (names per conv2d_.py:147-173, dense_.py:112-137, resnet.py:215-246)."""
    shapes = {}
    H, W, C = cfg.x_shape
    # q_net
    cq = _block_shapes(shapes, 'q_net', _q_blocks(cfg), C, False)
    if cfg.arch == 'A':
        zc = _zc(cfg)
        for head in ('z_mean', 'z_logstd'):        # dmm_cifar10.py:441-444
            _block_shapes(shapes, 'q_net', [(head, zc, 3, 1)], cq, False)
    else:
        hq, wq = H, W
        for _, _, _, s in _q_blocks(cfg):
            hq, wq = -(-hq // s), -(-wq // s)
        for head in ('z_mean', 'z_logstd'):        # dmm_for_sampling.py:445-446
            shapes['q_net/%s/kernel' % head] = (hq * wq * cq, cfg.z_dim)
            shapes['q_net/%s/bias' % head] = (cfg.z_dim,)
    # G_theta
    gh, gw = _g_base_hw(cfg)
    if cfg.arch == 'A':
        cg = _zc(cfg)
    else:
        cg = 64
        shapes['G_theta/level_0/kernel'] = (cfg.z_dim, 64 * gh * gw)
        shapes['G_theta/level_0/bias'] = (64 * gh * gw,)
    cg = _block_shapes(shapes, 'G_theta', _g_blocks(cfg), cg, True)
    if cfg.arch == 'M':                            # dmm_bernouli_fashion.py:517-521
        shapes['G_theta/feature_map_mean_to_pixel/kernel'] = (1, 1, cg, C)
        shapes['G_theta/feature_map_mean_to_pixel/bias'] = (C,)
    else:                                          # dmm_cifar10.py:489-496
        for head in ('x_mean', 'x_logstd'):
            shapes['G_theta/%s/kernel' % head] = (1, 1, cg, C)
            shapes['G_theta/%s/bias' % head] = (C,)
    # D_psi
    cd = _block_shapes(shapes, 'D_psi', _d_blocks(cfg), C, False)
    hd, wd = H, W
    for _, _, _, s in _d_blocks(cfg):
        hd, wd = -(-hd // s), -(-wd // s)
    shapes['D_psi/level_-2/kernel'] = (hd * wd * cd, 64)     # dmm_cifar10.py:558-559
    shapes['D_psi/level_-2/bias'] = (64,)
    shapes['D_psi/level_-1/kernel'] = (64, 1)                # :561
    shapes['D_psi/level_-1/bias'] = (1,)
    if cfg.latent_gan and cfg.arch == 'M':       # dmm_bernouli_fashion.py:526-543: six weight-normalised dense layers
        for i in range(6):
            pre = 'G_omega/dense%s/' % ('_%d' % i if i else '')
            fin, fout = (cfg.z_dim if i == 0 else 40), (cfg.z_dim if i == 5 else 40)
            shapes[pre + 'kernel'] = (fin, fout)
            shapes[pre + 'weight_norm/scale'] = (fin, fout)      # kernel-shaped, weight_norm_.py:84-92
            shapes[pre + 'bias'] = (fout,)
    elif cfg.latent_gan:
        oh, ow = _omega_base_hw(cfg)
        shapes['G_omega/level_0/kernel'] = (128, 256 * oh * ow)
        shapes['G_omega/level_0/bias'] = (256 * oh * ow,)
        co = _block_shapes(shapes, 'G_omega', _omega_blocks(cfg), 256, True)
        zch = cfg.z_dim // (4 * oh * ow)
        shapes['G_omega/feature_map_mean_to_pixel/kernel'] = (1, 1, co, zch)
        shapes['G_omega/feature_map_mean_to_pixel/bias'] = (zch,)
    if cfg.use_flow:                                         # planar_nf.py:79-103,208-221
        for i in range(cfg.nf_layers):
            pre = 'posterior_flow/' if cfg.nf_layers == 1 else 'posterior_flow/_%d/' % i
            shapes[pre + 'w'] = (1, cfg.z_dim)
            shapes[pre + 'b'] = (1,)
            shapes[pre + 'u'] = (1, cfg.z_dim)
    return shapes


def init_weights(cfg, seed=1234, dtype=np.float64, bias_std=0.1, head_gain=1.0):
    """Non-degenerate synthetic weights: EVERY kernel is
    glorot-normal (std = sqrt(2/(fan_in+fan_out)), layers/initialization.py:8-23)
    — including the heads the reference zero-initialises — and every bias is
    N(0, bias_std^2).  Identical arrays are fed to the oracle and the CUDA
    path; values are rounded through float32 so both sides hold the same
    numbers."""
    rng = np.random.RandomState(seed)
    params = {}
    for name, shp in weight_shapes(cfg).items():
        if name.endswith('/kernel'):
            if len(shp) == 4:
                rf = shp[0] * shp[1]
                fan_in, fan_out = shp[2] * rf, shp[3] * rf
            else:
                fan_in, fan_out = shp
            std = np.sqrt(2.0 / (fan_in + fan_out))
            w = rng.normal(0.0, std, size=shp)
        elif name.endswith('/weight_norm/scale'):
            w = 1.0 + rng.normal(0.0, 0.25, size=shp)      # the reference initialises ones (weight_norm_.py:64-65)
        elif name.startswith('posterior_flow/'):
            # the reference draws w, u ~ N(0, 0.01^2) (planar_nf.py:37-41), which leaves the flow at the
            # identity to 1e-4; the synthetic test weights are larger so that the layers do something
            w = rng.normal(0.0, 0.1 if name.endswith('/b') else 1.0 / np.sqrt(shp[-1]) * 2.0, size=shp)
        else:
            w = rng.normal(0.0, bias_std, size=shp)
        params[name] = w.astype(np.float32).astype(dtype)
    return params


def q_net(cfg, params, x, tape=None):
    """q_net hidden features + heads -> (z_mean, clipped z_logstd), [B, z_dim]
    (A: dmm_cifar10.py:418-452; B: dmm_for_sampling.py:421-460;
    M: dmm_bernouli_fashion.py:432-478)."""
    h = x
    for scope, cout, k, s in _q_blocks(cfg):
        h = resnet_conv2d_block(h, params, 'q_net/' + scope, cout, k, s, tape)
    if cfg.arch == 'A':
        zc = _zc(cfg)
        z_mean = resnet_conv2d_block(h, params, 'q_net/z_mean', zc, 3, 1)
        z_logstd = resnet_conv2d_block(h, params, 'q_net/z_logstd', zc, 3, 1)
        z_mean = z_mean.reshape(z_mean.shape[:-3] + (-1,))      # reshape_tail(...,3,[-1])
        z_logstd = z_logstd.reshape(z_logstd.shape[:-3] + (-1,))
    else:
        hf = h.reshape(h.shape[:-3] + (-1,))
        z_mean = dense(hf, params['q_net/z_mean/kernel'], params['q_net/z_mean/bias'])
        z_logstd = dense(hf, params['q_net/z_logstd/kernel'], params['q_net/z_logstd/bias'])
    z_logstd = maybe_clip_value(z_logstd, min_val=cfg.min_logstd_of_q)
    return z_mean, z_logstd


def _g_trunk(cfg, params, z, tape=None):
    gh, gw = _g_base_hw(cfg)
    if cfg.arch == 'A':
        h = z.reshape(z.shape[:-1] + (gh, gw, _zc(cfg)))        # dmm_cifar10.py:480
    else:
        h = dense(z, params['G_theta/level_0/kernel'], params['G_theta/level_0/bias'])
        h = h.reshape(h.shape[:-1] + (gh, gw, 64))
    for scope, cout, k, s in _g_blocks(cfg):
        h = resnet_deconv2d_block(h, params, 'G_theta/' + scope, cout, k, s, tape)
    return h


def weight_norm(kernel, scale=None, epsilon=1e-12):
    """tfsnippet.layers.weight_norm(kernel, axis=-1) (layers/normalization/weight_norm_.py:76-97): tf.nn.l2_normalize over
    every axis but the last (x / sqrt(max(sum x^2, epsilon))), times the kernel-shaped `scale`.  The reference's own test
    states it in numpy: scale * kernel / sqrt(sum(kernel ** 2, axis=all but last, keepdims=True))
    (tests/layers/normalization/test_weight_norm.py:23-53)."""
    kernel = np.asarray(kernel)
    red = tuple(range(kernel.ndim - 1))
    w = kernel / np.sqrt(np.maximum(np.sum(kernel ** 2, axis=red, keepdims=True), epsilon))
    return w if scale is None else w * np.asarray(scale)


def G_omega(cfg, params, z_omega):
    """gan_z = flatten(G_omega(z_omega, z_channel)) (dmm_cifar10.py:503-532,630-632; dmm_for_sampling.py:519-546,
    646-648): z_omega [..., 128] -> [..., z_dim].  The 1x1 output conv has no activation.
    Arch M (dmm_bernouli_fashion.py:526-543,636-640): z_omega [..., z_dim]; six dense layers with weight_norm=True and
    leaky_relu on every one of them, the last included."""
    if cfg.arch == 'M':
        h = np.asarray(z_omega)
        for i in range(6):
            pre = 'G_omega/dense%s/' % ('_%d' % i if i else '')
            h = leaky_relu(dense(h, weight_norm(params[pre + 'kernel'], params[pre + 'weight_norm/scale']), params[pre + 'bias']))
        return h
    oh, ow = _omega_base_hw(cfg)
    h = dense(z_omega, params['G_omega/level_0/kernel'], params['G_omega/level_0/bias'])
    h = h.reshape(h.shape[:-1] + (oh, ow, 256))
    for scope, cout, k, s in _omega_blocks(cfg):
        h = resnet_deconv2d_block(h, params, 'G_omega/' + scope, cout, k, s)
    x = conv2d(h, params['G_omega/feature_map_mean_to_pixel/kernel'], params['G_omega/feature_map_mean_to_pixel/bias'])
    return x.reshape(x.shape[:-3] + (-1,))


def G_theta(cfg, params, z, return_std=False, tape=None):
    """G_theta(z[, return_std]) (A: dmm_cifar10.py:466-500; B:
    dmm_for_sampling.py:476-514; M: dmm_bernouli_fashion.py:492-521).  The
    heads read the trunk output WITHOUT a preceding activation; x_mean has a
    tanh (conv2d activation_fn, applied after the bias)."""
    h = _g_trunk(cfg, params, z, tape)
    if tape is not None:
        tape.append(dict(head_in=h))
    if cfg.arch == 'M':
        x_mean = conv2d(h, params['G_theta/feature_map_mean_to_pixel/kernel'],
                        params['G_theta/feature_map_mean_to_pixel/bias'], 1, np.tanh)
        if return_std:
            raise TypeError('arch M G_theta has no return_std')
        return x_mean
    x_mean = conv2d(h, params['G_theta/x_mean/kernel'], params['G_theta/x_mean/bias'], 1, np.tanh)
    if not return_std:
        return x_mean
    x_logstd = conv2d(h, params['G_theta/x_logstd/kernel'], params['G_theta/x_logstd/bias'], 1)
    return x_mean, x_logstd


def D_psi(cfg, params, x, tape=None):
    """D_psi(x) -> [...] (A: dmm_cifar10.py:535-562).  No activation before the
    flatten, none between the two dense layers (the arg_scope only covers
    resnet_conv2d_block)."""
    h = x
    for scope, cout, k, s in _d_blocks(cfg):
        h = resnet_conv2d_block(h, params, 'D_psi/' + scope, cout, k, s, tape)
    if tape is not None:
        tape.append(dict(flat_shape=h.shape))
    hf = h.reshape(h.shape[:-3] + (-1,))
    hf = dense(hf, params['D_psi/level_-2/kernel'], params['D_psi/level_-2/bias'])
    hf = dense(hf, params['D_psi/level_-1/kernel'], params['D_psi/level_-1/bias'])
    return hf[..., 0]


def D_psi_input_grad(cfg, params, x):
    """(D(x), dD/dx) by manual reverse mode (for ele_grad_norm,
    evaluation/dmm_cifar.py:955-970, and for grad_z E)."""
    tape = []
    d = D_psi(cfg, params, x, tape)
    flat_shape = tape.pop()['flat_shape']
    g = np.ones(d.shape + (1,), dtype=x.dtype) @ params['D_psi/level_-1/kernel'].T
    g = g @ params['D_psi/level_-2/kernel'].T
    g = g.reshape(flat_shape)
    for rec in reversed(tape):
        g = block_input_grad(g, params, rec)
    return d, g


def G_theta_input_grad(cfg, params, z, g_x_mean):
    """Vector-Jacobian product of x_mean = G_theta(z) w.r.t. z."""
    tape = []
    x_mean = G_theta(cfg, params, z, tape=tape)
    h = tape.pop()['head_in']
    name = 'G_theta/feature_map_mean_to_pixel' if cfg.arch == 'M' else 'G_theta/x_mean'
    g = g_x_mean * (1.0 - np.square(x_mean))                     # tanh'
    g = conv2d_input_grad(g, params[name + '/kernel'], 1, h.shape[-3:-1])
    for rec in reversed(tape):
        g = block_input_grad(g, params, rec)
    if cfg.arch == 'A':
        g = g.reshape(z.shape)
    else:
        g = g.reshape(g.shape[:-3] + (-1,)) @ params['G_theta/level_0/kernel'].T
    return x_mean, g
