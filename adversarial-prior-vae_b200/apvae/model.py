"""Device handle of the evaluation / sampling path: ExpConfig + weights -> libapvae_b200.so.

`ExpConfig` carries the fields of the experiment scripts' config classes that the
path reads (experiment/dmm/dmm_cifar10.py:27-91, dmm_for_sampling.py:28-92,
dmm_bernouli_fashion.py ExpConfig); `Model` owns one `apvae_handle` (one per
GPU / rank) and exposes the fused entry points of include/apvae.h on torch CUDA
tensors.  torch is plumbing only (device buffers, streams): every number comes
out of the library's kernels, and there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import lib, check

_ARCH = {'A': _lib.ARCH_A, 'B': _lib.ARCH_B, 'M': _lib.ARCH_M}


class ExpConfig(object):
    """Same field names as the scripts' ExpConfig.  arch: 'A' = dmm_cifar10.py,
    'B' = dmm_for_sampling.py (dataset 'celebA' | 'cifar10'), 'M' = dmm_bernouli_fashion.py."""

    def __init__(self, arch='A', dataset=None, **kw):
        if arch not in _ARCH:
            raise ValueError('arch must be one of A, B, M')
        self.arch = arch
        self.dataset = dataset or {'A': 'cifar10', 'B': 'celebA', 'M': 'fashion'}[arch]
        c = _lib.Config()
        check(lib.apvae_default_config(_ARCH[arch], int(self.dataset == 'celebA'), C.byref(c)))
        self.x_shape = (c.x_h, c.x_w, c.x_c)
        self.z_dim = c.z_dim
        self.pull_back_energy_weight = c.pull_back_energy_weight
        self.min_logstd_of_q = c.min_logstd_of_q
        self.epsilon = c.epsilon
        self.uniform_scale = bool(c.uniform_scale)
        self.use_truncated = bool(c.use_truncated)
        self.smallest_step = c.smallest_step
        self.len_train = c.len_train
        self.initial_xi = 0.0
        self.test_n_qz = {'A': 10, 'B': 10, 'M': 1000}[arch]       # dmm_cifar10.py:70, dmm_bernouli_fashion.py:81
        self.test_n_pz = 1000                                       # :69
        self.test_batch_size = {'A': 64, 'B': 64, 'M': 1}[arch]     # :71, dmm_bernouli_fashion.py:82
        self.sample_n_z = 100
        self.use_flow = False                                       # dmm_bernouli_fashion.py:69 ("VAEPP+Flow")
        self.nf_layers = 20                                         # :48
        self.latent_gan = False    # also build G_omega (gan_z = the MALA chains' initial_z, dmm_cifar10.py:503-532,623-639)
        # capacities of this build (not reference fields)
        self.max_batch = 64
        self.max_chunk = 512
        self.enable_grad = False
        self.precise = False       # split-precision forward of G_theta / D_psi for the sampler: True / 1 = activations AND kernels
                                   # as hi + lo 16-bit pairs (3 tensor-core terms, fp32-class energies); 2 = activations only
                                   # (2 terms: the kernels keep their 16-bit rounding, a fixed perturbation that cancels in
                                   # E(z) - E(z'))
        self.tuning = {}           # apvae_tuning overrides (kernel-variant selection for A-B runs and tests; {} = defaults)
        self.operand_dtype = 'fp16'  # 16-bit operand format of the tensor-core path: 'fp16' (11 significant bits; default)
                                     # or 'bf16' (8 bits, fp32's exponent range); fp32 accumulation either way
        for k, v in kw.items():
            if not hasattr(self, k):
                raise AttributeError('ExpConfig has no field %r' % (k,))
            setattr(self, k, v)

    def to_c(self):
        c = _lib.Config()
        c.arch = _ARCH[self.arch]
        c.x_h, c.x_w, c.x_c = [int(v) for v in self.x_shape]
        c.z_dim = int(self.z_dim)
        c.celeba = int(self.dataset == 'celebA')
        c.pull_back_energy_weight = float(self.pull_back_energy_weight)
        c.min_logstd_of_q = float(self.min_logstd_of_q)
        c.epsilon = float(self.epsilon)
        c.uniform_scale = int(self.uniform_scale)
        c.use_truncated = int(self.use_truncated)
        c.bernoulli = int(self.arch == 'M')
        c.max_batch, c.max_chunk = int(self.max_batch), int(self.max_chunk)
        c.enable_grad = int(self.enable_grad)
        c.smallest_step = float(self.smallest_step)
        c.len_train = int(self.len_train)
        c.precise = int(self.precise)
        c.nf_layers = int(self.nf_layers) if self.use_flow else 0
        c.latent_gan = int(self.latent_gan)
        if self.operand_dtype not in ('fp16', 'bf16'):
            raise ValueError("operand_dtype must be 'fp16' or 'bf16'")
        c.operand_bf16 = int(self.operand_dtype == 'bf16')
        c.tuning = _lib.Tuning.make(self.tuning)
        return c


def weight_schema(config):
    """[(tf_variable_name, shape)] the path needs (works without a GPU)."""
    c = config.to_c()
    n = C.c_int()
    check(lib.apvae_weight_count(C.byref(c), C.byref(n)))
    out = []
    name = C.create_string_buffer(256)
    shape = (C.c_int64 * 4)()
    nd = C.c_int()
    for i in range(n.value):
        check(lib.apvae_weight_info(C.byref(c), i, name, 256, shape, C.byref(nd)))
        out.append((name.value.decode(), tuple(int(shape[j]) for j in range(nd.value))))
    return out


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Model(object):
    """One `apvae_handle`.  Inputs / outputs are float32 CUDA tensors in the reference layouts
    (NHWC images, [K,B,...] sample-major)."""

    def __init__(self, config, params, device=0, xi_raw=None, beta=0.0, log_Z=1.0):
        import torch
        self.config = config
        self.device = torch.device('cuda', int(device))
        self._h = C.c_void_p()
        c = config.to_c()
        check(lib.apvae_create(C.byref(c), int(device), C.byref(self._h)))
        names, arrs = [], []
        for name, shape in weight_schema(config):
            if name not in params:
                raise KeyError('missing variable %r' % (name,))
            a = np.ascontiguousarray(params[name], dtype=np.float32)
            if tuple(a.shape) != tuple(shape):
                raise ValueError('variable %r: shape %r, expected %r' % (name, a.shape, shape))
            names.append(name.encode())
            arrs.append(a)
        n = len(names)
        c_names = (C.c_char_p * n)(*names)
        c_data = (C.c_void_p * n)(*[a.ctypes.data for a in arrs])
        shapes = [(C.c_int64 * a.ndim)(*a.shape) for a in arrs]
        c_shapes = (C.c_void_p * n)(*[C.addressof(s) for s in shapes])
        c_nd = (C.c_int * n)(*[a.ndim for a in arrs])
        check(lib.apvae_load_weights(self._h, n, c_names, c_data, c_shapes, c_nd), self._h)
        self.set_scalars(config.initial_xi if xi_raw is None else xi_raw, beta, log_Z)

    def close(self):
        if getattr(self, '_h', None):
            lib.apvae_destroy(self._h)
            self._h = None

    __del__ = close

    @classmethod
    def from_checkpoint(cls, config, save_path, device=0):
        """The model restored from a reference checkpoint (`tf.train.Saver` V2 files written by tfsnippet's
        CheckpointSaver, scaffold/checkpoint.py:230-253): `save_path` is a checkpoint prefix or a directory holding a
        `checkpoint` state file (tf.train.latest_checkpoint).  xi / beta / log_Z come from the checkpoint when present."""
        import os
        from . import checkpoint as ck
        prefix = save_path
        if os.path.isdir(save_path):
            prefix = ck.latest_checkpoint(save_path)
            if prefix is None:
                raise IOError('no checkpoint state in %r' % (save_path,))
        weights, scalars = ck.model_weights_from_checkpoint(prefix, config)
        return cls(config, weights, device=device, xi_raw=scalars.get('xi_raw'), beta=scalars.get('beta', 0.0),
                   log_Z=scalars.get('log_Z', 1.0))

    def set_scalars(self, xi_raw, beta, log_Z):
        """'p_net/xi' (pre-sigmoid), 'beta', 'log_Z' (dmm_cifar10.py:604-607,860,458-463)."""
        self.xi_raw, self.beta, self.log_Z = float(xi_raw), float(beta), float(log_Z)
        check(lib.apvae_set_scalars(self._h, self.xi_raw, self.beta, self.log_Z), self._h)

    # ------------------------------------------------------------------ helpers
    def _f32(self, t, shape=None):
        import torch
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(np.asarray(t, dtype=np.float32))
        t = t.to(device=self.device, dtype=torch.float32).contiguous()
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError('expected shape %r, got %r' % (tuple(shape), tuple(t.shape)))
        return t

    def _images(self, x):
        """float32 device tensor [B, H, W, C]; the trailing dims must be the model's x_shape (the reference's
        InputSpec raises ValueError at graph-build time for a wrong static shape)."""
        t = self._f32(x)
        want = tuple(int(v) for v in self.config.x_shape)
        if t.dim() != 4 or tuple(t.shape[1:]) != want:
            raise ValueError('expected images of shape [B, %d, %d, %d], got %r' % (want + (tuple(t.shape),)))
        return t

    def _empty(self, *shape, **kw):
        import torch
        return torch.empty(shape, dtype=kw.get('dtype', torch.float32), device=self.device)

    # ------------------------------------------------------------------ nets
    def encode(self, x):
        """q_net(x) -> (z_mean, max(z_logstd, min_logstd_of_q)), each [B, z_dim]."""
        x = self._images(x)
        B = x.shape[0]
        mu, ls = self._empty(B, self.config.z_dim), self._empty(B, self.config.z_dim)
        check(lib.apvae_encode(self._h, _ptr(x), B, _ptr(mu), _ptr(ls), _stream()), self._h)
        return mu, ls

    def decode(self, z, return_std=False):
        """G_theta(z, return_std): z [..., z_dim] -> x_mean [..., H, W, C] (and x_logstd)."""
        z = self._f32(z)
        lead = tuple(z.shape[:-1])
        z2 = z.reshape(-1, self.config.z_dim)
        n = z2.shape[0]
        H, W, Cx = self.config.x_shape
        xm = self._empty(n, H, W, Cx)
        xs = self._empty(n, H, W, Cx) if return_std else None
        check(lib.apvae_decode(self._h, _ptr(z2), n, _ptr(xm), _ptr(xs), _stream()), self._h)
        xm = xm.reshape(lead + (H, W, Cx))
        return (xm, xs.reshape(lead + (H, W, Cx))) if return_std else xm

    def energy_of_images(self, x):
        """D_psi(x): x [..., H, W, C] -> [...]."""
        x = self._images(x)
        lead = tuple(x.shape[:-3])
        x4 = x.reshape((-1,) + tuple(self.config.x_shape))
        out = self._empty(x4.shape[0])
        check(lib.apvae_energy(self._h, _ptr(x4), None, x4.shape[0], _ptr(out), _stream()), self._h)
        return out.reshape(lead)

    def energy_of_latents(self, z):
        """D_psi(G_theta(z)) ('pure_energy', dmm_cifar10.py:175): z [..., z_dim] -> [...]."""
        z = self._f32(z)
        lead = tuple(z.shape[:-1])
        z2 = z.reshape(-1, self.config.z_dim)
        out = self._empty(z2.shape[0])
        check(lib.apvae_energy(self._h, None, _ptr(z2), z2.shape[0], _ptr(out), _stream()), self._h)
        return out.reshape(lead)

    def sample_z(self, x_origin, K, eps=None, seed=0):
        """q_net(x, n_z=K)['z'] (dmm_cifar10.py:418-452): returns (z [K,B,d], log q(z|x) [K,B], log N(z) [K,B])."""
        x = self._images(x_origin)
        B = x.shape[0]
        e = self._f32(eps, (K, B, self.config.z_dim)) if eps is not None else None
        z, lq, lpn = self._empty(K, B, self.config.z_dim), self._empty(K, B), self._empty(K, B)
        check(lib.apvae_sample_z(self._h, _ptr(x), B, int(K), _ptr(e), int(seed), _ptr(z), _ptr(lq), _ptr(lpn),
                                 _stream()), self._h)
        return z, lq, lpn

    # ------------------------------------------------------------------ estimator
    def is_nll(self, x, K, x_origin=None, eps=None, seed=0, per_sample=False):
        """The test graph of one mini-batch (dmm_cifar10.py:880-927).  Returns a dict of
        per-image [B] tensors (and per-sample [K,B] if asked)."""
        x = self._images(x)
        B = x.shape[0]
        xo = self._f32(x_origin, x.shape) if x_origin is not None else None
        e = self._f32(eps, (K, B, self.config.z_dim)) if eps is not None else None
        pi = self._empty(_lib.PER_IMAGE_ROWS, B)
        ps = self._empty(_lib.PER_SAMPLE_ROWS, K * B) if per_sample else None
        check(lib.apvae_is_nll(self._h, _ptr(x), _ptr(xo), B, int(K), _ptr(e), int(seed), _ptr(pi),
                               _ptr(ps), _stream()), self._h)
        out = dict(ele_test_ll=pi[_lib.ROW_LL], ele_test_lb=pi[_lib.ROW_LB],
                   ele_adv_test_ll=pi[_lib.ROW_ADV_LL], ele_adv_test_lb=pi[_lib.ROW_ADV_LB],
                   ele_test_recon=pi[_lib.ROW_RECON], ele_reconstruct_energy=pi[_lib.ROW_DG],
                   ele_lz_max=pi[_lib.ROW_LZ_MAX], ele_lz_sum=pi[_lib.ROW_LZ_SUM],
                   ele_real_energy=pi[_lib.ROW_REAL_E], ele_test_mse=pi[_lib.ROW_MSE], per_image=pi)
        if per_sample:
            out.update(log_px=ps[_lib.SROW_LOG_PX].reshape(K, B), log_q=ps[_lib.SROW_LOG_Q].reshape(K, B),
                       log_pn=ps[_lib.SROW_LOG_PN].reshape(K, B), d_g=ps[_lib.SROW_DG].reshape(K, B),
                       sq_err=ps[_lib.SROW_SQ_ERR].reshape(K, B))
        return out

    def prior_stats(self, Kp, eps=None, seed=0):
        """test_pn_net branch (dmm_cifar10.py:883,915-927): D(G(z)) and log N(z) for Kp prior draws."""
        e = self._f32(eps).reshape(Kp, self.config.z_dim) if eps is not None else None
        dg, lpn = self._empty(Kp), self._empty(Kp)
        check(lib.apvae_prior_stats(self._h, int(Kp), _ptr(e), int(seed), _ptr(dg), _ptr(lpn), None, _stream()), self._h)
        return dg, lpn

    def prior_branch(self, Kp=None, eps=None, seed=0, out=None):
        """The batch scalars of the test_pn_net branch (dmm_cifar10.py:883,915-927), reduced on the device: a float32
        tensor [4] = (pn_energy, pd_energy, kl_adv_and_gaussian, log_Z_compute_op) from Kp = config.test_n_pz draws
        z ~ N(0, I) (fed as eps [Kp, z_dim] or drawn in-kernel)."""
        Kp = int(self.config.test_n_pz if Kp is None else Kp)
        e = self._f32(eps).reshape(Kp, self.config.z_dim) if eps is not None else None
        out = self._empty(4) if out is None else out
        check(lib.apvae_prior_stats(self._h, Kp, _ptr(e), int(seed), None, None, _ptr(out), _stream()), self._h)
        return out

    # ------------------------------------------------------------------ sampler
    def energy_grad(self, z):
        """get_sgld_proposal's (energy, grad, pure_energy) at z [n, z_dim] (dmm_cifar10.py:241-247)."""
        z = self._f32(z)
        lead = tuple(z.shape[:-1])
        z2 = z.reshape(-1, self.config.z_dim)
        n = z2.shape[0]
        e, g, p = self._empty(n), self._empty(n, self.config.z_dim), self._empty(n)
        check(lib.apvae_energy_grad(self._h, _ptr(z2), n, _ptr(e), _ptr(g), _ptr(p), _stream()), self._h)
        return e.reshape(lead), g.reshape(lead + (self.config.z_dim,)), p.reshape(lead)

    def image_energy_grad(self, x):
        """(D_psi(x), ele_grad = dD/dx, ele_grad_norm = sum(ele_grad^2)) of evaluation/dmm_cifar.py:958-959."""
        x = self._images(x)
        x4 = x.reshape((-1,) + tuple(self.config.x_shape))
        n = x4.shape[0]
        d, g, s = self._empty(n), self._empty(*x4.shape), self._empty(n)
        check(lib.apvae_image_energy_grad(self._h, _ptr(x4), n, _ptr(d), _ptr(g), _ptr(s), _stream()), self._h)
        return d, g.reshape(x.shape), s

    def mala(self, z0, steps, alpha=None, eps=None, u=None, seed=0, history=True, history_z=False):
        """EnergyDistribution.sample (dmm_cifar10.py:179-239): `steps` MALA iterations.
        z0 [n, z_dim] (or [n,1,z_dim]); eps [steps,n,z_dim], u [steps,n] optional fed-in noise.
        history_z=True also returns history_z [steps, ...z0.shape]: the state BEFORE each step (:200,235); the reference's
        plot_samples restarts its next run from history_z[-1] (:1061)."""
        self._check_sampler_precision()
        z = self._f32(z0).clone()
        shape = tuple(z.shape)
        z2 = z.reshape(-1, self.config.z_dim)
        n = z2.shape[0]
        alpha = self.config.smallest_step if alpha is None else alpha
        e = self._f32(eps).reshape(steps, n, self.config.z_dim) if eps is not None else None
        uu = self._f32(u).reshape(steps, n) if u is not None else None
        he = self._empty(steps, n) if history else None
        hp = self._empty(steps, n) if history else None
        hr = self._empty(steps, n) if history else None
        hz = self._empty(steps, n, self.config.z_dim) if history_z else None
        check(lib.apvae_mala(self._h, _ptr(z2), n, int(steps), float(alpha), _ptr(e), _ptr(uu), int(seed),
                             _ptr(he), _ptr(hp), _ptr(hr), _ptr(hz), _stream()), self._h)
        out = dict(z=z2.reshape(shape), history_e_z=he, history_pure_e_z=hp, history_ratio=hr)
        if history_z:
            out['history_z'] = hz.reshape((steps,) + shape)
        return out

    # absolute error of D(G(z)) a single-pass 16-bit forward leaves (measured maxima against the float64 oracle)
    SINGLE_PASS_ENERGY_ERROR = {'fp16': 2e-3, 'bf16': 2e-2}

    def _check_sampler_precision(self):
        """The accept / reject exponent of MALA carries pull_back_energy_weight * sigmoid(xi) * D(G(z)): at the reference's
        weight of 1000 a single-pass 16-bit forward puts ~1 nat (fp16) to ~10 nats (bf16) of rounding noise into it and
        the chain silently walks away from the reference's.  Say so, once per model, unless the caller has chosen a
        split-precision forward (config.precise = 1 or 2)."""
        if int(self.config.precise) != 0 or getattr(self, '_precision_warned', False):
            return
        import math
        import warnings
        xi = 1.0 / (1.0 + math.exp(-self.xi_raw))
        noise = self.config.pull_back_energy_weight * xi * self.SINGLE_PASS_ENERGY_ERROR[self.config.operand_dtype]
        if noise > 0.1:
            self._precision_warned = True
            warnings.warn('MALA with a single-pass %s forward: pull_back_energy_weight * xi = %.0f turns the ~%.0e error of '
                          'D(G(z)) into ~%.1f nats inside the acceptance ratio; build the model with ExpConfig(precise=2) '
                          '(or precise=True) to sample the reference\'s chain' %
                          (self.config.operand_dtype, self.config.pull_back_energy_weight * xi,
                           self.SINGLE_PASS_ENERGY_ERROR[self.config.operand_dtype], noise), RuntimeWarning, stacklevel=3)

    def fill_normal(self, shape, seed=0, offset=0, truncated=False):
        """N(0, 1) draws (Philox, in-kernel) as tf.random_normal inside zd.Normal.sample; `truncated`: resampled
        beyond 2 sigma (experiment/truncated_normal.py:96-107)."""
        out = self._empty(*tuple(int(v) for v in shape))
        check(lib.apvae_fill_normal(_ptr(out), out.numel(), int(seed), int(offset), int(bool(truncated)), _stream()))
        return out

    def latent_gan(self, n, z_omega=None, seed=0):
        """gan_z = p_omega_net(n_z=n)['f_z'].distribution.mean (dmm_cifar10.py:623-635,967-969): G_omega applied to
        z_omega ~ N(0, I_128) (fed [n,128] or drawn in-kernel) -> [n, z_dim], the `initial_z` of plot_samples' MALA loop."""
        din = self.config.z_dim if self.config.arch == 'M' else 128        # dmm_bernouli_fashion.py:636-639: z ~ N(0, I_z_dim)
        zo = self._f32(z_omega, (n, din)) if z_omega is not None else None
        out = self._empty(n, self.config.z_dim)
        check(lib.apvae_latent_gan(self._h, _ptr(zo), int(n), int(seed), _ptr(out), _stream()), self._h)
        return out

    def sample_images(self, z):
        """x_plots + np.round (dmm_cifar10.py:976-977,990,1067): uint8 [n,H,W,C]."""
        import torch
        z2 = self._f32(z).reshape(-1, self.config.z_dim)
        n = z2.shape[0]
        out = self._empty(n, *self.config.x_shape, dtype=torch.uint8)
        check(lib.apvae_sample_images(self._h, _ptr(z2), n, _ptr(out), _stream()), self._h)
        return out

    # ------------------------------------------------------------------ bookkeeping
    def launch_count(self):
        n = C.c_int64()
        check(lib.apvae_launch_count(self._h, C.byref(n)))
        return n.value

    def macs(self):
        """Algorithmic multiply-accumulates: (q per image, G per sample, D per sample)."""
        q, g, d = C.c_double(), C.c_double(), C.c_double()
        check(lib.apvae_macs(self._h, C.byref(q), C.byref(g), C.byref(d)))
        return q.value, g.value, d.value

    def op_times(self, n, reps=5, grad=False, best_of=3):
        """Per-op CUDA-event timings (bench.py roofline): [(net, name, macs_per_unit, ms_per_launch)]
        for every planned op of q_net (n capped at max_batch), G_theta and D_psi on `n` units; grad=True adds the
        adjoint programs of the sampler (model_grad.cu)."""
        out = []
        name = C.create_string_buffer(256)
        macs, ms, cnt = C.c_double(), C.c_float(), C.c_int()
        nets = [(0, 'q_net'), (1, 'G_theta'), (2, 'D_psi')]
        if grad and self.config.enable_grad:
            nets += [(3, 'D_psi^T'), (4, 'G_theta^T')]
        for net, label in nets:
            check(lib.apvae_op_count(self._h, net, C.byref(cnt)))
            nn = min(n, self.config.max_batch) if net == 0 else min(n, self.config.max_chunk)
            for i in range(cnt.value):
                best = None
                for _ in range(max(1, int(best_of))):      # a kernel timed alone: the best of a few event-timed groups
                    check(lib.apvae_op_time(self._h, net, i, int(nn), int(reps), name, 256, C.byref(macs),
                                            C.byref(ms), _stream()), self._h)
                    best = ms.value if best is None else min(best, ms.value)
                out.append((label, name.value.decode(), macs.value, best, nn))
        return out
