PLX_BO_LIB=polyaxon_b200/libplxbo_profile.so timeout 100 python - <<EOF 2>&1 | grep "fs_eval n=128" | tail -2
import numpy as np, sys
sys.path.insert(0, ".")
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
h = _lib.Handle(0)
X, y, _ = synthetic_problem(128, 16)
h.set_data(X, y, 1, 2.5)
h.fit(0.0, [])
EOF
for f in 1 0; do PLX_FS=$f timeout 100 python - <<EOF
import os, time, numpy as np, sys
sys.path.insert(0, ".")
from polyaxon_b200 import _lib
from oracle.reference_path import synthetic_problem
h = _lib.Handle(0)
h.set_option("fit_small", int(os.environ["PLX_FS"]))
for n in (16, 32, 64, 96, 128):
    X, y, _ = synthetic_problem(n, 16)
    h.set_data(X, y, 1, 2.5)
    rs = np.random.RandomState(1).uniform(-11, 11, 4)
    h.fit(0.0, rs)
    t0 = time.perf_counter()
    for _ in range(5): th, lml, pr = h.fit(0.0, rs)
    print("fit_small", os.environ["PLX_FS"], "n", n, "ms %.3f" % ((time.perf_counter() - t0) / 5 * 1e3), "evals", pr[:, 3].astype(int).tolist(), "theta %.9f lml %.9f" % (th, lml))
EOF
done
