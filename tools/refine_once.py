import sys
import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib
n, d = int(sys.argv[1]), int(sys.argv[2])
X, y, bounds = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, _lib.KERNEL_RBF, 1.5)
h.fit(0.0, [])
rs = np.random.RandomState(3)
seeds = rs.uniform(0, 1, (10, d))
out = h.refine(seeds, bounds, _lib.ACQ_EI, float(y.max()), eps=0.0, want_nfev=True)
print(out[3].tolist())
