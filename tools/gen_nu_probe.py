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

# fit and refinement with the tables (value + derivative) against the routine
import time
h = _lib.Handle(0)
for n, d, R in [(512, 16, 5), (2048, 32, 3)]:
    X, y, bounds = synthetic_problem(n, d)
    rs = np.random.RandomState(3)
    restarts = rs.uniform(np.log(1e-5), np.log(1e5), R - 1)
    res = {}
    for tab in (0, 1):
        h.set_option('gen_table', tab)
        h.set_data(X, y, 1, 1.9)
        h.fit(0.0, restarts)
        t0 = time.perf_counter()
        theta, lml, per_run = h.fit(0.0, restarts)
        ms_fit = (time.perf_counter() - t0) * 1e3
        seeds = rs.uniform(0, 1, (10, d))
        h.refine(seeds, bounds, _lib.ACQ_EI, float(y.max()), eps=0.0)
        t0 = time.perf_counter()
        xr, vr, ok = h.refine(seeds, bounds, _lib.ACQ_EI, float(y.max()), eps=0.0)
        ms_ref = (time.perf_counter() - t0) * 1e3
        res[tab] = (theta, lml, ms_fit, ms_ref, vr)
    print('n %d d %d R %d nu 1.9: fit %.1f -> %.1f ms, theta* %.12g / %.12g, lml %.12g / %.12g; refinement %.1f -> %.1f ms, max |dvalue| %.2e' % (
        n, d, R, res[0][2], res[1][2], res[0][0], res[1][0], res[0][1], res[1][1], res[0][3], res[1][3], np.abs(res[0][4] - res[1][4]).max()), flush=True)
h.set_option('gen_table', 1)
h.close()
