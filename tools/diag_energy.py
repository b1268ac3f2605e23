"""Diagnostic: D(G(z)) error of the bf16 path vs the float64 oracle per GRAD case (run under the tuning toggles)."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, './adversarial-prior-vae_b200'); sys.path.insert(0, './tests')
import oracle as O
from test_model_gpu import GRAD_CASES, _make
for case in sorted(GRAD_CASES):
    arch, kw, n = GRAD_CASES[case]
    for mc in (3, 4):
        ocfg, params, cfg, m = _make(arch, kw, max_chunk=mc, enable_grad=True)
        m.set_scalars(xi_raw=0.4, beta=0.0, log_Z=1.0)
        z = np.random.RandomState(11).normal(size=(n, ocfg.z_dim))
        e_o, g_o, p_o = O.energy_and_grad(ocfg, params, z, xi_raw=0.4)
        e, g, p = [t.cpu().numpy().astype(np.float64) for t in m.energy_grad(z)]
        xm_o = O.G_theta(ocfg, params, z)
        xm_o = xm_o[0] if isinstance(xm_o, tuple) else xm_o
        xm = m.decode(z)
        xm = (xm[0] if isinstance(xm, tuple) else xm).cpu().numpy()
        print(case, 'max_chunk', mc, 'D(G(z)) err', np.round(p - p_o, 4), 'x_mean err max/mean', np.abs(xm - xm_o).max(), np.abs(xm - xm_o).mean())
        m.close()
