"""A/B of the sweep variants on one GP state: kernel time (CUDA events inside the library) and results."""
import sys
import numpy as np
sys.path.insert(0, '.')
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem

n, d = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2048, 32)
N = int(sys.argv[3]) if len(sys.argv) > 3 else 148 * 128 * 8
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, 1, 1.5)
h.set_theta(1.2)
rs = np.random.RandomState(0)
Xc = rs.uniform(0, 1, (N, d))
ref = None
for impl in (1, 2, 4, 5, 6):
    h.set_option('sweep_impl', impl)
    ms = []
    for _ in range(3):
        out = h.sweep(Xc, _lib.ACQ_POI, float(y.max()), dump=True)
        ms.append(h.last_sweep_stats()[0])
    v, i, x, mean, std, acq = out
    if ref is None:
        ref = out
    print('impl', impl, 'kernel ms', ['%.3f' % m for m in ms], 'evals/s %.4g' % (N / min(ms) * 1e3), 'best', v, i,
          'max|dmean| %.2e max|dvar| %.2e max|dacq| %.2e' % (np.max(np.abs(mean - ref[3])), np.max(np.abs(std ** 2 - ref[4] ** 2)),
                                                          np.max(np.abs(acq - ref[5]))))
