"""One batched LML+gradient evaluation at (n, d, B) for an ncu launch list: python tools/lml_once.py [n=256] [d=8] [B=1]"""
import sys
sys.path.insert(0, '.')
import numpy as np
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
n, d, B = (int(a) for a in (sys.argv[1:4] + ['256', '8', '1'][len(sys.argv) - 1:]))
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_data(X, y, 1, 1.5)
for _ in range(3):
    print(h.lml(np.full(B, 0.5)))
h.close()
