"""Time split of the fused int8-sliced sweep: whole, tensor instructions skipped (builder-bound), builders skipped.
Needs the profiling build:  python -m polyaxon_b200.build --profile
                            PLX_BO_LIB=polyaxon_b200/libplxbo_profile.so python tools/i8_phases.py [impl] [i8_operand] [dbg,...] [i8_distance]"""
import sys
import numpy as np
sys.path.insert(0, '.')
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
import torch
n, d, N = 2048, 32, 148 * 128 * 8
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, 1, 1.5)
h.set_theta(1.2)
xc = torch.rand(N, d, dtype=torch.float64, device='cuda')
torch.cuda.synchronize()
h.set_option('sweep_impl', int(sys.argv[1]) if len(sys.argv) > 1 else 5)
h.set_option('i8_operand', int(sys.argv[2]) if len(sys.argv) > 2 else 0)
h.set_option('i8_distance', int(sys.argv[4]) if len(sys.argv) > 4 else 0)
for dbg in [int(a) for a in sys.argv[3].split(',')] if len(sys.argv) > 3 else (0, 64, 128, 64 | 128):
    h.set_option('sweep_dbg', dbg)
    ms = []
    for _ in range(3):
        h.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
        ms.append(h.last_sweep_stats()[0])
    print('impl %s operand %s dbg %d: %.2f ms' % (sys.argv[1:2], sys.argv[2:3], dbg, min(ms)))
