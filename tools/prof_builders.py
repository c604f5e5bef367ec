"""ncu driver for the builders of the cooperative int8 sweep: one warm launch, then one launch per debug mask
(profiling build only; 0 = whole kernel, 64 = tensor instructions skipped, i.e. the builders beside the operand stream).
PLX_BO_LIB=polyaxon_b200/libplxbo_profile.so python tools/prof_builders.py [impl=5] [dbg,dbg,...=0,64] [i8_distance=0]"""
import sys
sys.path.insert(0, '.')
import torch
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem

impl = int(sys.argv[1]) if len(sys.argv) > 1 else 5
masks = [int(a) for a in sys.argv[2].split(',')] if len(sys.argv) > 2 else [0, 64]
n, d, N = 2048, 32, 148 * 128 * 8
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, 1, 1.5)
h.set_theta(1.2)
h.set_option('sweep_impl', impl)
h.set_option('i8_distance', int(sys.argv[3]) if len(sys.argv) > 3 else 0)
xc = torch.rand(N, d, dtype=torch.float64, device='cuda', generator=torch.Generator(device='cuda').manual_seed(1))
torch.cuda.synchronize()
h.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
for m in masks:
    h.set_option('sweep_dbg', m)
    h.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
    print('impl %d dbg %d: %.3f ms' % (impl, m, h.last_sweep_stats()[0]))
h.close()
