"""General-nu Matern: acquisition value / gradient and LML / gradient with the Chebyshev tables (gen_table 1) against the Bessel
routine (0) at a few states.  python tools/gen_nu_grad_check.py"""
import sys
sys.path.insert(0, '.')
import numpy as np
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
h = _lib.Handle(0)
for n, d, theta in [(512, 16, 1.1), (300, 4, -0.5), (300, 4, 2.5)]:
    X, y, bounds = synthetic_problem(n, d)
    rs = np.random.RandomState(5)
    pts = rs.uniform(0, 1, (64, d))
    out = {}
    for tab in (0, 1):
        h.set_option('gen_table', tab)
        h.set_data(X, y, 1, 1.9)
        h.set_theta(theta)
        out[tab] = h.acq_value_grad(pts, _lib.ACQ_EI, float(y.max()), eps=0.0)
        lml = h.lml(np.array([theta, theta + 0.3]))
        out[tab] = out[tab] + (lml,)
    v0, g0, l0 = out[0]; v1, g1, l1 = out[1]
    print(n, d, theta, 'max |dv| %.2e (scale %.2e)  max |dg| %.2e (scale %.2e)' % (np.abs(v0 - v1).max(), np.abs(v0).max(), np.abs(g0 - g1).max(), np.abs(g0).max()),
          'lml', l0[0], l1[0], 'grad', l0[1], l1[1])
h.close()
