"""Where get_suggestions time goes on the device path: fit / sweep / refine per BASELINE config."""
import sys
import time
import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib

h = _lib.Handle(0)


def t(fn, reps=5):
    fn()
    out = []
    for _ in range(reps):
        t0 = time.perf_counter()
        r = fn()
        out.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(out)), r


for label, n, d, kern, nu, R, N, S, acq in [
        ('C1', 15, 2, _lib.KERNEL_MATERN, 2.5, 1, 100000, 10, _lib.ACQ_UCB),
        ('C5m', 128, 16, _lib.KERNEL_MATERN, 2.5, 5, 100000, 10, _lib.ACQ_EI),
        ('C5', 512, 16, _lib.KERNEL_MATERN, 2.5, 5, 100000, 10, _lib.ACQ_EI),
        ('C2', 256, 8, _lib.KERNEL_RBF, 1.5, 1, 1000000, 10, _lib.ACQ_EI),
        ('C3fit', 2048, 32, _lib.KERNEL_MATERN, 1.5, 17, 0, 10, _lib.ACQ_POI),
        ('C4-512', 512, 32, _lib.KERNEL_MATERN, 1.5, 33, 0, 0, _lib.ACQ_POI),
        ('C4-4096', 4096, 32, _lib.KERNEL_MATERN, 1.5, 33, 0, 0, _lib.ACQ_POI)]:
    X, y, bounds = synthetic_problem(n, d)
    rs = np.random.RandomState(3)
    restarts = rs.uniform(np.log(1e-5), np.log(1e5), R - 1)
    ms_data, _ = t(lambda: h.set_data(X, y, kern, nu))
    reps = 3 if n >= 2048 else 5
    l0 = h.launch_count()
    ms_fit, (theta, lml, per_run) = t(lambda: h.fit(0.0, restarts), reps)
    launches = (h.launch_count() - l0) / (reps + 1)
    ms_lml, _ = t(lambda: h.lml(np.full(R, 0.5)), reps)
    line = '%-8s n=%-5d R=%-3d set_data %.2f ms | fit %.1f ms (evals/run %s, %d launches) | one batched LML+grad %.2f ms' % (
        label, n, R, ms_data, ms_fit, per_run[:, 3].astype(int).tolist()[:6], launches, ms_lml)
    if N:
        xc = rs.uniform(0, 1, (N, d))
        ms_rng, _ = t(lambda: rs.uniform(0, 1, (N, d)), 3)
        ms_sweep, _ = t(lambda: h.sweep(xc, acq, float(y.max()), kappa=2.576, eps=0.0), 5)
        kms = h.last_sweep_stats()[0]
        line += ' | rng %.1f ms | sweep(host) %.2f ms (kernel %.2f)' % (ms_rng, ms_sweep, kms)
    if S:
        seeds = rs.uniform(0, 1, (S, d))
        ms_ref, out = t(lambda: h.refine(seeds, bounds, acq, float(y.max()), kappa=2.576, eps=0.0, want_nfev=True), 5)
        line += ' | refine %.2f ms (nfev %s)' % (ms_ref, out[3].tolist())
    print(line)
    sys.stdout.flush()
