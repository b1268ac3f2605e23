"""Diagnostic: one MALA step, oracle vs GPU (prints energies, ratios, z error)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests')); sys.path.insert(0, os.path.join(ROOT, 'adversarial-prior-vae_b200'))
import oracle as O
from apvae.model import ExpConfig, Model
arch, kw = sys.argv[1], ({'dataset': sys.argv[2]} if len(sys.argv) > 2 and sys.argv[1] == 'B' else ({'z_dim': 128} if sys.argv[1] == 'A' else {}))
alpha = float(os.environ.get('ALPHA', '2e-3'))
ocfg = O.OracleConfig(arch, **kw); params = O.init_weights(ocfg)
cfg = ExpConfig(arch, dataset=kw.get('dataset'), z_dim=ocfg.z_dim, max_batch=8, max_chunk=4, enable_grad=True)
m = Model(cfg, params); m.set_scalars(xi_raw=0.0, beta=0.0, log_Z=1.0)
n = 6; rs = np.random.RandomState(5)
z0 = rs.normal(size=(n, 1, ocfg.z_dim)); eps = rs.normal(size=(1, n, 1, ocfg.z_dim)); u = np.full((1, n, 1), 0.5)
print('w', ocfg.pull_back_energy_weight, 'alpha', alpha)
e_o, g_o, p_o = O.energy_and_grad(ocfg, params, z0, 0.0)
e, g, p = [t.cpu().numpy().reshape(x.shape) for t, x in zip(m.energy_grad(z0), (e_o, g_o, p_o))]
print('pure  o', p_o.ravel()); print('pure  g', p.ravel())
print('E err', (e - e_o).ravel())
zp_o = z0 - alpha * g_o + eps[0] * np.sqrt(2 * alpha)
zp = z0 - alpha * g + eps[0] * np.sqrt(2 * alpha)
print('zprime max abs diff', np.abs(zp - zp_o).max(), ' |g| max', np.abs(g_o).max(), 'g abs err max', np.abs(g - g_o).max())
e2_o, g2_o, p2_o = O.energy_and_grad(ocfg, params, zp_o, 0.0)
e2, g2, p2 = [t.cpu().numpy().reshape(x.shape) for t, x in zip(m.energy_grad(zp_o), (e2_o, g2_o, p2_o))]
print('dE oracle', (e_o - e2_o).ravel()); print('dE gpu   ', (e - e2).ravel())
zo, eo, po, ro = O.mala_step(ocfg, params, z0, eps[0], u[0], alpha, 0.0)
out = m.mala(z0, 1, alpha=alpha, eps=eps, u=u)
print('ratio o', ro.ravel()); print('ratio g', out['history_ratio'].cpu().numpy().ravel())
print('z err', np.abs(out['z'].cpu().numpy().reshape(zo.shape) - zo).max())
