"""Small end-to-end pass over every kernel for compute-sanitizer (memcheck)."""
import sys
import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib
h = _lib.Handle(0)
for n, d, kern, nu in [(37, 3, _lib.KERNEL_MATERN, 2.5), (200, 33, _lib.KERNEL_RBF, 1.5), (130, 5, _lib.KERNEL_MATERN, 1.9)]:
    X, y, b = synthetic_problem(n, d)
    h.set_data(X, y, kern, nu)
    th, lml, _ = h.fit(0.0, [1.0])
    xc = np.random.RandomState(0).uniform(0, 1, (300, d))
    for impl in (1, 0):
        h.set_option('sweep_impl', impl)
        print(n, d, impl, h.sweep(xc, _lib.ACQ_EI, float(y.max()), eps=0.01, dump=True)[:2])
    h.candidates_set(xc)
    print(h.sweep_resident(_lib.ACQ_UCB, 0.0, kappa=1.0)[:2])
    print(h.refine(xc[:5], b, _lib.ACQ_UCB, 0.0, kappa=1.0)[1])
    print(h.acq_value_grad(xc[:3], _lib.ACQ_POI, float(y.max()))[0])
print(h.kernel_values(1, 1.5, np.array([0.0, 1.0, 1e6]))[:2])
h.close()
print('done')
