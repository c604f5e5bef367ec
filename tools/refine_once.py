"""One refinement of ten seeds at (n, d), three times, with wall time and evaluation counts (also the ncu driver for the
refine_* kernels): python tools/refine_once.py 512 16"""
import sys
sys.path.insert(0, '.')
import numpy as np
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
n, d = int(sys.argv[1]), int(sys.argv[2])
X, y, bounds = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, _lib.KERNEL_MATERN, 2.5)
h.set_theta(0.3)
rs = np.random.RandomState(3)
seeds = rs.uniform(0, 1, (10, d))
import time
for _ in range(3):
    t0 = time.perf_counter()
    out = h.refine(seeds, bounds, _lib.ACQ_EI, float(y.max()), eps=0.0, want_nfev=True)
    print('%.2f ms' % ((time.perf_counter() - t0) * 1e3), out[3].tolist())
h.close()
