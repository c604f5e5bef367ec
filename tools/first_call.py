"""Where the first sweep of a fresh handle spends its time (allocation, first launch of each kernel)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
n, d = int(sys.argv[1]), int(sys.argv[2])
X, y, _ = synthetic_problem(n, d)
xc = np.random.RandomState(0).uniform(0, 1, (200000, d))
for rep in range(2):
    h = _lib.Handle(0)
    t0 = time.perf_counter(); h.set_data(X, y, 1, 1.5); h.set_theta(0.3); t1 = time.perf_counter()
    h.candidates_set(xc); t2 = time.perf_counter()
    for k in range(3):
        s0 = time.perf_counter(); h.sweep_resident(_lib.ACQ_EI, float(y.max())); s1 = time.perf_counter()
        print('handle %d sweep %d: %.1f ms (kernel %.2f)' % (rep, k, (s1 - s0) * 1e3, h.last_sweep_stats()[0]))
    print('handle %d: state %.1f ms, candidates %.1f ms' % (rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3))
    h.close()
