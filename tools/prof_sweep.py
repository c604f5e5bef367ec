"""Short driver for ncu: a few launches of one sweep form at the C3 state on 151 552 device-resident candidates
(8 tiles per SM).  python tools/prof_sweep.py [sweep_impl=3] [i8_operand=0] [launches=3]"""
import sys
sys.path.insert(0, '.')
import torch
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem

impl = int(sys.argv[1]) if len(sys.argv) > 1 else 3
operand = int(sys.argv[2]) if len(sys.argv) > 2 else 0
launches = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n, d, N = 2048, 32, 148 * 128 * 8
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, 1, 1.5)
h.set_theta(1.2)
h.set_option('sweep_impl', impl)
h.set_option('i8_operand', operand)
xc = torch.rand(N, d, dtype=torch.float64, device='cuda', generator=torch.Generator(device='cuda').manual_seed(1))
torch.cuda.synchronize()
for _ in range(launches):
    v, i, _ = h.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
    print('impl %d operand %d: %.3f ms, best %r %d' % (impl, operand, h.last_sweep_stats()[0], v, i))
h.close()
