import sys
import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib
n, R = int(sys.argv[1]), int(sys.argv[2])
X, y, _ = synthetic_problem(n, 32)
h = _lib.Handle(0)
h.set_data(X, y, _lib.KERNEL_MATERN, 1.5)
print(h.lml(np.linspace(0.2, 1.5, R))[0][:3])
