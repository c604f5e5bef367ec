"""A/B of the builders' distance forms in the int8 sweep (option i8_distance 0 direct / 1 FP64 tensor path)
against the numpy oracle: kernel time and posterior mean / std differences, over shapes that exercise the padding and the
cancellation fallback (candidates that ARE observations).  python tools/ab_distance.py"""
import sys
import numpy as np
sys.path.insert(0, '.')
from polyaxon_b200 import _lib
from oracle import gp_numpy
from oracle.reference_path import synthetic_problem

KINDS = {0: ('rbf', None), 1: ('matern', 0.5), 2: ('matern', 1.5), 3: ('matern', 2.5)}


def run(n, d, N, kind, impl=5, theta=1.2, check=True):
    X, y, _ = synthetic_problem(n, d)
    h = _lib.Handle(0)
    kname, nu = KINDS[kind]
    h.set_data(X, y, 0 if kname == 'rbf' else 1, nu if nu else 1.5)
    h.set_theta(theta)
    h.set_option('sweep_impl', impl)
    rs = np.random.RandomState(0)
    Xc = rs.uniform(0, 1, (N, d))
    Xc[:min(n, 64)] = X[:min(n, 64)]                     # exact zero distances
    Xc[64:96] = X[:32] + 1e-9                            # and near-zero ones
    outs = {}
    for dist in (0, 1):
        h.set_option('i8_distance', dist)
        ms = []
        for _ in range(3):
            out = h.sweep(Xc, _lib.ACQ_UCB, float(y.max()), kappa=2.0, dump=True)
            ms.append(h.last_sweep_stats()[0])
        outs[dist] = (out, min(ms))
    line = 'n %d d %d N %d kind %d impl %d:' % (n, d, N, kind, impl)
    if check:
        L, alpha, _ = h.get_state(want_L=True)
        m = min(N, 4096)
        mean_o, std_o = gp_numpy.predict(Xc[:m], X, L, alpha, theta, 0 if kname == 'rbf' else 1, nu if nu else 1.5)
    for dist in (0, 1):
        (v, i, x, mean, std, acq), ms = outs[dist]
        line += '\n   dist %d: %.3f ms  best %.12g @%d' % (dist, ms, v, i)
        if check:
            line += '  mean rel %.2e  std rel %.2e (floor 1e-6: %.2e)' % (
                np.max(np.abs(mean[:m] - mean_o) / np.maximum(np.abs(mean_o), 1e-12)),
                np.max(np.abs(std[:m] - std_o) / np.maximum(std_o, 1e-300)),
                np.max(np.abs(std[:m] - std_o) / np.maximum(std_o, 1e-6)))
        line += '  vs dist 0: dmean %.2e dstd %.2e' % (np.max(np.abs(mean - outs[0][0][3])), np.max(np.abs(std - outs[0][0][4])))
    print(line, flush=True)
    h.close()


if __name__ == '__main__':
    for kind in (3, 0, 1, 2):
        run(2048, 32, 148 * 128 * 8, kind)
    run(2048, 32, 148 * 128 * 8, 3, impl=4)
    run(2048, 32, 148 * 128 * 8, 3, impl=6)
    run(300, 7, 5000, 3, impl=4)
    run(1000, 2, 5000, 2, impl=4)
    run(1999, 32, 20000, 3)
    run(4096, 17, 148 * 128 * 2, 0)
    run(2048, 2, 148 * 128 * 2, 2, check=False)       # few dimensions: many pairs under the cancellation bound
    run(2048, 16, 148 * 128 * 2, 2, check=False)
