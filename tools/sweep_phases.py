"""Profiling aid: time the sweep kernel whole, without phase A (K* built for the first tile only)
and without the DMMA work, to see how the time splits."""
import sys
import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib
import torch

n, d, N = 2048, 32, 148 * 128 * 80
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, _lib.KERNEL_MATERN, 1.5)
h.set_theta(1.2)
xc = torch.rand(N, d, dtype=torch.float64, device='cuda')
torch.cuda.synchronize()
for impl in [int(a) for a in sys.argv[1:]] or [1]:
    h.set_option('sweep_impl', impl)
    for dbg in (0, 1, 2, 3, 2 | 4, 2 | 8, 2 | 4 | 8):
        h.set_option('sweep_dbg', dbg)
        ms = []
        for _ in range(3):
            h.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
            ms.append(h.last_sweep_stats()[0])
        print('impl %d dbg %d  kernel %.2f ms  (%.2f M evals/s)' % (impl, dbg, min(ms), N / min(ms) / 1e3))
