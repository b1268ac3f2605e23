import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'./adversarial-prior-vae_b200'); sys.path.insert(0,'./tests')
import oracle as O
from test_model_gpu import CASES, _make, _inputs
for case in sorted(CASES):
    arch, kw, B, K = CASES[case]
    ocfg, params, cfg, m = _make(arch, kw)
    x, xo, eps = _inputs(ocfg, B, K)
    want = O.eval_batch(ocfg, params, x, xo, eps, xi_raw=0.3, beta=-0.5, log_Z=2.5)
    m.set_scalars(0.3,-0.5,2.5)
    got = m.is_nll(x, K, x_origin=xo, eps=eps, per_sample=True)
    g = {k: v.cpu().numpy().astype(np.float64) for k, v in got.items()}
    mu, ls = m.encode(xo)
    print(case, 'mu err', np.abs(mu.cpu().numpy()-want['mu']).max(), 'scale', np.abs(want['mu']).max(), 'ls err', np.abs(ls.cpu().numpy()-want['logstd']).max())
    for k in ('log_q','log_pn','d_g','log_px'):
        print('  ',k,'abs err',np.abs(g[k]-want[k]).max(),'scale',np.abs(want[k]).max())
    for k in ('ele_test_ll','ele_adv_test_ll','ele_test_lb'):
        print('  ',k,'rel err',(np.abs(g[k]-want[k])/np.abs(want[k])).max(), want[k][:2])
    print('   real_e err', np.abs(g['ele_real_energy']-want['ele_real_energy']).max())
    m.close()
