"""General-nu Matern in the sweeps: the Chebyshev table of the kernel value (option gen_table 1, default) against the
Bessel routine for every value (0): kernel time and posterior differences.  python tools/gen_nu_probe.py"""
import sys
sys.path.insert(0, '.')
import numpy as np
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem

h = _lib.Handle(0)
for n, d, N, nu, theta in [(2048, 32, 148 * 128 * 8, 1.9, 1.2), (300, 2, 20000, 1.9, -1.0), (40, 3, 20000, 0.7, 0.0),
                           (700, 8, 30000, 3.3, 0.5), (2048, 32, 148 * 128, 1.9, 4.0)]:
    X, y, _ = synthetic_problem(n, d)
    h.set_data(X, y, 1, nu)
    h.set_theta(theta)
    xc = np.random.RandomState(1).uniform(0, 1, (N, d))
    xc[:min(n, 50)] = X[:min(n, 50)]
    out = {}
    for tab in (0, 1):
        h.set_option('gen_table', tab)
        res = h.sweep(xc, _lib.ACQ_UCB, float(y.max()), kappa=1.0, dump=True)
        res = h.sweep(xc, _lib.ACQ_UCB, float(y.max()), kappa=1.0, dump=True)
        out[tab] = (res, h.last_sweep_stats()[0])
    (a, ta), (b, tb) = out[0], out[1]
    print('n %d d %d N %d nu %.1f theta %.1f: routine %.2f ms, table %.2f ms; arg-max %d / %d; max |dmean| %.2e (scale %.2e), max |dvar| %.2e' % (
        n, d, N, nu, theta, ta, tb, a[1], b[1], np.abs(a[3] - b[3]).max(), np.abs(a[3]).max(), np.abs(a[4] ** 2 - b[4] ** 2).max()), flush=True)
h.set_option('gen_table', 1)
h.close()
