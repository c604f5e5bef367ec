import sys, warnings
import numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import golden_names, load_golden
from oracle import gp_numpy as G
from polyaxon_b200 import _lib
h = _lib.Handle(0)
for name in [n for n in golden_names() if '_m19_' in n]:
    g = load_golden(name)
    h.set_data(g['X'], g['y'], 1, float(g['nu']))
    h.set_theta(float(g['theta']))
    L, alpha, _ = h.get_state()
    cond = float(g['cond'])
    print(name, 'cond %.1e' % cond)
    print('  Ldiag rel', np.abs(np.diag(L) / g['L_diag'] - 1).max(), ' alpha rel', np.abs(alpha - g['alpha']).max() / np.abs(g['alpha']).max())
    args = dict(acq=int(g['acq']), y_max=float(g['y_max']), kappa=float(np.nan_to_num(g['kappa'])), eps=float(np.nan_to_num(g['eps'])))
    _, _, _, mean, std, acq = h.sweep(g['x_tries_head'], dump=True, **args)
    print('  mean abs', np.abs(mean - g['mean']).max(), 'scale', max(np.abs(g['mean']).max(), np.abs(g['y']).max()), ' var abs', np.abs(std**2 - g['std']**2).max())
    lml, grad = h.lml(g['fixed_thetas'])
    for i, th in enumerate(g['fixed_thetas']):
        fd = (G.log_marginal_likelihood(g['X'], g['y'], float(th) + 1e-5, 1, float(g['nu']), eval_gradient=False)
              - G.log_marginal_likelihood(g['X'], g['y'], float(th) - 1e-5, 1, float(g['nu']), eval_gradient=False)) / 2e-5
        print('  theta %5.2f lml dev %.12g ref %.12g | grad dev %.8g fd %.8g sk %.8g' % (th, lml[i], g['lml_fixed'][i], grad[i], fd, g['grad_fixed'][i]))
    th, l, per = h.fit(np.log(float(g['length_scale'])), g['theta_inits'])
    print('  fit theta', th, 'golden', float(g['theta']), 'lml', l, 'golden', float(g['lml']), 'per_run', per.tolist(), 'ref', g['per_restart'].tolist())
    # K matrix entry check vs oracle
    K = G.kernel_matrix(g['X'], float(g['theta']), 1, float(g['nu']))
    Kd = (L @ L.T)
    print('  K recon rel', np.abs(Kd - K - 1e-10 * np.eye(len(K))).max())
