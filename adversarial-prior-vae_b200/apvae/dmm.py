"""The experiment scripts' call surface (experiment/dmm/dmm_cifar10.py, dmm_for_sampling.py,
dmm_bernouli_fashion.py) over one `Model`: same builder names, arguments and attribute paths, eager
instead of graph-building.  A script's test section

    test_q_net = q_net(input_origin_x, posterior_flow, n_z=config.test_n_qz)
    test_chain = test_q_net.chain(p_net, latent_axis=0, observed={'x': input_x}, n_z=..., beta=beta, log_Z=...)
    ele_test_ll = test_chain.vi.evaluation.is_loglikelihood()             # :896-898
    test_recon  = tf.reduce_mean(test_chain.model['x'].log_prob())        # :886-888
    adv_test_nll via test_chain.model['z'].log_prob().energy / .log_energy_prob    # :901-910

reads the same here with `s = Script(model)` and `s.q_net(...)`, `s.p_net`, `s.G_theta`, `s.D_psi`.
Everything is computed by the library (apvae_sample_z / apvae_is_nll / apvae_decode / apvae_energy /
apvae_mala); torch only holds the buffers.  Reductions over the sample axis that the reference writes as
tf.reduce_mean(..) on the returned tensors are the caller's, as in the reference."""
import math


class _Lazy(object):
    """Evaluates the mini-batch once (one apvae_is_nll call with per-sample rows) on first use."""

    def __init__(self, model, x, x_origin, K, eps, seed, beta, log_Z):
        self.model, self.x, self.x_origin, self.K, self.eps, self.seed = model, x, x_origin, K, eps, seed
        self.beta, self.log_Z = beta, log_Z
        self._out = None

    def get(self):
        if self._out is None:
            m = self.model
            if self.beta is not None or self.log_Z is not None:
                m.set_scalars(m.xi_raw, m.beta if self.beta is None else self.beta,
                              m.log_Z if self.log_Z is None else self.log_Z)
            self._out = m.is_nll(self.x, self.K, x_origin=self.x_origin, eps=self.eps, seed=self.seed, per_sample=True)
        return self._out


class _LogProb(object):
    """What EnergyDistribution.log_prob returns (dmm_cifar10.py:164-177): the Gaussian log p_N(z) as a tensor with
    .energy, .log_energy_prob and .pure_energy attached."""

    def __init__(self, log_pn, d_g, w_xi, log_Z, z_dim):
        self.tensor = log_pn
        self.pure_energy = d_g                                                        # :175
        half_sq = -log_pn - 0.5 * z_dim * math.log(2 * math.pi)                       # 0.5 |z|^2
        self.energy = w_xi * d_g + half_sq                                            # :169-171
        self.log_energy_prob = -self.energy - max(float(log_Z), -20.0)                # :135,173


class _StochasticTensor(object):
    def __init__(self, tensor=None, log_prob=None, distribution=None):
        self.tensor, self._log_prob, self.distribution = tensor, log_prob, distribution

    def log_prob(self, group_ndims=None):
        return self._log_prob() if callable(self._log_prob) else self._log_prob


class _Distribution(object):
    def __init__(self, mean=None, sample=None):
        self.mean = mean
        self._sample = sample

    def sample(self, n_samples=None):
        return self._sample(n_samples)


class _Evaluation(object):
    def __init__(self, lazy):
        self._lazy = lazy

    def is_loglikelihood(self):
        """variational/evaluation.py:9-39 on log_joint = log p(x|z) + log p_N(z), latent log q(z|x)."""
        return self._lazy.get()['ele_test_ll']

    importance_sampling_log_likelihood = is_loglikelihood


class _LowerBound(object):
    def __init__(self, lazy):
        self._lazy = lazy

    def elbo(self):
        """variational/objectives.py:9-42, averaged over the sample axis as the scripts do (:899)."""
        return self._lazy.get()['ele_test_lb']


class _VI(object):
    def __init__(self, lazy):
        self.evaluation = _Evaluation(lazy)
        self.lower_bound = _LowerBound(lazy)


class Chain(object):
    """q_net(...).chain(p_net, ...) (bayes.py:347-417, variational/chain.py:25-68)."""

    def __init__(self, script, lazy):
        m = script.model
        self.vi = _VI(lazy)
        g = lazy.get

        def z_log_prob():
            o = g()
            return _LogProb(o['log_pn'], o['d_g'], script.w_xi(), m.log_Z, m.config.z_dim)
        self.model = {
            'x': _StochasticTensor(log_prob=lambda: g()['log_px']),
            'z': _StochasticTensor(log_prob=z_log_prob),
        }
        self.variational = {'z': _StochasticTensor(log_prob=lambda: g()['log_q'])}
        self._lazy = lazy

    # the adversarial-prior estimators the scripts assemble by hand (:901-911)
    def adv_is_loglikelihood(self):
        return self._lazy.get()['ele_adv_test_ll']

    def adv_elbo(self):
        return self._lazy.get()['ele_adv_test_lb']


class QNet(object):
    """BayesianNet returned by q_net (dmm_cifar10.py:418-452)."""

    def __init__(self, script, x, n_z, eps, seed):
        if n_z is None or int(n_z) < 1:
            raise ValueError('`n_z` (the number of latent samples) must be a positive integer')
        self._script, self._x, self.n_z, self._eps, self._seed = script, x, int(n_z), eps, seed
        self._z = None

    def _draw(self):
        if self._z is None:
            self._z = self._script.model.sample_z(self._x, self.n_z, eps=self._eps, seed=self._seed)
        return self._z

    def __getitem__(self, name):
        if name != 'z':
            raise KeyError(name)
        return _StochasticTensor(tensor=self._draw()[0], log_prob=lambda: self._draw()[1])

    def chain(self, model_builder, latent_axis=0, observed=None, n_z=None, beta=None, log_Z=None, **kw):
        if latent_axis not in (0, None):
            raise ValueError('only latent_axis=0 (sample axis first) is used by the scripts')
        if not observed or 'x' not in observed:
            raise ValueError("chain() needs observed={'x': ...}")
        K = self.n_z if n_z is None else int(n_z)
        lazy = _Lazy(self._script.model, observed['x'], self._x, K, self._eps, self._seed, beta, log_Z)
        return Chain(self._script, lazy)


class PNet(object):
    """BayesianNet returned by p_net (dmm_cifar10.py:596-620).  With z observed it wraps the given latents; otherwise
    z is a sample of the energy prior: N(0, I) draws (n_z of them; pz.sample, :188-192) or `initial_z`, moved by
    `mcmc_iterator` MALA steps (:197-226).  `p_net(n_z=config.test_n_pz, mcmc_iterator=0)` is the test graph's
    test_pn_net (:883): its 'z'.log_prob() carries .energy / .log_energy_prob / .pure_energy at the prior draws."""

    def __init__(self, script, z, n_z, mcmc_iterator, mcmc_alpha, initial_z, seed, log_Z=None):
        self._script, self._z = script, z
        m = script.model
        self._log_Z = m.log_Z if log_Z is None else float(log_Z)
        self.history = None
        if z is None:
            if initial_z is None:
                if n_z is None:
                    raise ValueError('p_net needs observed z, an initial_z, or n_z prior samples')
                initial_z = m.fill_normal((int(n_z), 1, m.config.z_dim), seed=seed, offset=1 << 48)
            steps = int(mcmc_iterator)
            out = m.mala(initial_z, steps, alpha=mcmc_alpha, seed=seed, history_z=True) if steps > 0 else {'z': m._f32(initial_z)}
            self._z = out['z']
            self.history = out
        self._xm = None
        self._lp = None

    def _z_log_prob(self):
        if self._lp is None:
            m = self._script.model
            z = m._f32(self._z)
            d = m.config.z_dim
            d_g = m.energy_of_latents(z)
            log_pn = -0.5 * (z * z).sum(-1) - 0.5 * d * math.log(2 * math.pi)
            self._lp = _LogProb(log_pn, d_g, self._script.w_xi(), self._log_Z, d)
        return self._lp

    def __getitem__(self, name):
        m = self._script.model
        if name == 'z':
            t = _StochasticTensor(tensor=self._z, log_prob=self._z_log_prob)
            if self.history is not None:
                for k in ('history_z', 'history_e_z', 'history_pure_e_z', 'history_ratio'):
                    setattr(t, k, self.history.get(k))
            return t
        if name == 'x':
            if self._xm is None:
                self._xm = m.decode(self._z)
            return _StochasticTensor(tensor=None, distribution=_Distribution(mean=self._xm))
        raise KeyError(name)


class POmegaNet(object):
    """BayesianNet returned by p_omega_net (dmm_cifar10.py:623-639): 'f_z' has mean gan_z = G_omega(z_omega),
    'x' has mean G_theta(gan_z)."""

    def __init__(self, script, n_z, z_omega, seed):
        self._script = script
        self._fz = script.model.latent_gan(int(n_z), z_omega=z_omega, seed=seed)
        self._xm = None

    def __getitem__(self, name):
        if name == 'f_z':
            return _StochasticTensor(tensor=None, distribution=_Distribution(mean=self._fz))
        if name == 'x':
            if self._xm is None:
                self._xm = self._script.model.decode(self._fz)
            return _StochasticTensor(tensor=None, distribution=_Distribution(mean=self._xm))
        raise KeyError(name)


class _UnitNormal(object):
    """spt.Normal(mean=0, logstd=0) as the scripts build the base p(z) (dmm_cifar10.py:600-603)."""
    is_continuous = True
    is_reparameterized = True


class EnergyDistribution(object):
    """EnergyDistribution(pz, G, D, log_Z, xi, mcmc_iterator, mcmc_alpha, initial_z) (dmm_cifar10.py:97-252):
    p(z) proportional to exp(-w * xi * D(G(z)) - 0.5 |z|^2).  G and D must be the bound Script's G_theta / D_psi
    (the library evaluates D(G(z)) and its gradient itself); `xi` is the sigmoid-ed weight the scripts pass
    (`tf.nn.sigmoid(xi_node)`, :607) and must agree with the model's `xi_raw`.

    log_prob(given) returns the Gaussian log p_N(z) carrying .energy / .log_energy_prob / .pure_energy (:164-177);
    sample(n_samples) runs `mcmc_iterator` MALA steps from `initial_z` (or a N(0, I) draw) and returns a stochastic
    tensor carrying .history_z / .history_e_z / .history_pure_e_z / .history_ratio (:179-239)."""

    def __init__(self, pz, G, D, log_Z=0., xi=1.0, mcmc_iterator=0, mcmc_alpha=None, mcmc_algorithm='mala',
                 mcmc_space='z', initial_z=None, seed=0):
        if not getattr(pz, 'is_continuous', False):
            raise TypeError('`base_distribution` must be a continuous distribution.')         # :117-119
        script = getattr(G, '__self__', None)
        if not isinstance(script, Script) or getattr(D, '__self__', None) is not script:
            raise TypeError('G and D must be the G_theta / D_psi of one apvae.dmm.Script')
        if mcmc_algorithm != 'mala' or mcmc_space != 'z':
            raise ValueError("only mcmc_algorithm='mala' in mcmc_space='z' exists (as in the reference)")
        self._script, self._pz, self._G, self._D = script, pz, G, D
        m = script.model
        want = 1.0 / (1.0 + math.exp(-m.xi_raw))
        if abs(float(xi) - want) > 1e-6 * max(1.0, abs(want)):
            raise ValueError('xi = %g disagrees with sigmoid(xi_raw) = %g of the model' % (float(xi), want))
        self._xi = float(xi)
        self._log_Z = max(float(log_Z), -20.0)                                                # :135
        self._mcmc_iterator = int(mcmc_iterator)
        self._mcmc_alpha = m.config.smallest_step if mcmc_alpha is None else float(mcmc_alpha)
        self._initial_z = initial_z
        self._seed = seed

    pz = property(lambda self: self._pz)
    G = property(lambda self: self._G)
    D = property(lambda self: self._D)
    xi = property(lambda self: self._xi)
    log_Z = property(lambda self: self._log_Z)

    def set_log_Z(self, new_log_Z):
        self._log_Z = new_log_Z
        return new_log_Z

    def log_prob(self, given, group_ndims=0, name=None, y=None):
        m = self._script.model
        z = m._f32(given)
        d = m.config.z_dim
        d_g = m.energy_of_latents(z)
        log_pn = -0.5 * (z * z).sum(-1) - 0.5 * d * math.log(2 * math.pi)      # spt.Normal(0, 1).log_prob, group_ndims=1
        return _LogProb(log_pn, d_g, self._script.w_xi(), self._log_Z, d)

    def sample(self, n_samples=None, group_ndims=0, is_reparameterized=None, compute_density=None, name=None):
        m = self._script.model
        z0 = self._initial_z
        if z0 is None:
            if n_samples is None:
                raise ValueError('sample() needs n_samples or an initial_z')
            z0 = m.fill_normal((int(n_samples), m.config.z_dim), seed=self._seed)
        if self._mcmc_iterator > 0:
            out = m.mala(z0, self._mcmc_iterator, alpha=self._mcmc_alpha, seed=self._seed, history_z=True)
        else:
            out = {'z': m._f32(z0)}
        t = _StochasticTensor(tensor=out['z'], distribution=self)
        t.history_z = out.get('history_z')                       # state BEFORE each step (:200,235)
        t.history_e_z = out.get('history_e_z')
        t.history_pure_e_z = out.get('history_pure_e_z')
        t.history_ratio = out.get('history_ratio')
        return t


class PlanarFlows(object):
    def __init__(self, n_layers, name):
        if n_layers < 1:
            raise ValueError('n_layers must be positive')          # validate_positive_int_arg, planar_nf.py:200
        self.n_layers, self.name = n_layers, name


class Script(object):
    """Namespace of one experiment script bound to a Model: q_net / p_net / G_theta / D_psi."""

    def __init__(self, model):
        self.model = model
        self.config = model.config

    def w_xi(self):
        return self.config.pull_back_energy_weight / (1.0 + math.exp(-self.model.xi_raw))

    def planar_normalizing_flows(self, n_layers=1, name='posterior_flow'):
        """spt.layers.planar_normalizing_flows(config.nf_layers, name='posterior_flow') (dmm_cifar10.py:850-851): a
        descriptor only; the flow's variables (`posterior_flow/_i/{w,b,u}`) live in the model's weight dict."""
        return PlanarFlows(int(n_layers), name)

    def q_net(self, x, posterior_flow=None, observed=None, n_z=None, eps=None, seed=0):
        # like the reference (dmm_bernouli_fashion.py:468-474; dmm_cifar10.py's q_net ignores the argument), the flow is
        # applied iff config.use_flow; the handle was then created with nf_layers flow layers
        if self.config.use_flow and isinstance(posterior_flow, PlanarFlows) and posterior_flow.n_layers != self.config.nf_layers:
            raise ValueError('posterior_flow has %d layers, the model was built with nf_layers=%d'
                             % (posterior_flow.n_layers, self.config.nf_layers))
        return QNet(self, x, n_z, eps, seed)

    def p_net(self, observed=None, n_z=None, beta=1.0, mcmc_iterator=0, log_Z=0.0, initial_z=None, mcmc_alpha=None,
              seed=0):
        z = None if not observed else observed.get('z')
        return PNet(self, z, n_z, mcmc_iterator, mcmc_alpha, initial_z, seed, log_Z=log_Z)

    def p_omega_net(self, observed=None, n_z=None, beta=1.0, mcmc_iterator=0, log_Z=0.0, initial_z=None, z_omega=None,
                    seed=0):
        """gan_net = p_omega_net(n_z=sample_n_z, beta=beta); gan_z = gan_net['f_z'].distribution.mean
        (dmm_cifar10.py:965-969).  Needs ExpConfig(latent_gan=True)."""
        if n_z is None:
            raise ValueError('p_omega_net needs n_z')
        return POmegaNet(self, n_z, z_omega, seed)

    def Normal(self, mean=0.0, logstd=0.0):
        """The base p(z) of the scripts, spt.Normal(mean=zeros, logstd=zeros) (dmm_cifar10.py:600-603)."""
        if mean != 0.0 or logstd != 0.0:
            raise ValueError('the energy prior is built on N(0, I)')
        return _UnitNormal()

    def G_theta(self, z, return_std=False):
        return self.model.decode(z, return_std=return_std)

    def D_psi(self, x, y=None):
        return self.model.energy_of_images(x)
