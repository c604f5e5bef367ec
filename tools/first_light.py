"""Bring-up check: each sweep variant in its own process so that a hang cannot take the rest down."""
import subprocess
import sys

CODE = r'''
import sys, time, numpy as np
sys.path.insert(0, '.')
from oracle import gp_numpy as G
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import _lib
impl, n, d = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
X, y, _ = synthetic_problem(n, d)
h = _lib.Handle(0)
h.set_option('sweep_impl', impl)
h.set_data(X, y, _lib.KERNEL_MATERN, 1.5)
t0 = time.time(); h.set_theta(0.8); t1 = time.time()
L, a = G.posterior_state(X, y, 0.8, G.KERNEL_MATERN, 1.5)
Ld, ad, Z = h.get_state(want_L=True, want_Linv=True)
print('impl', impl, 'n', n, 'd', d, 'set_theta %.3fs' % (t1 - t0))
print(' L err', np.abs(Ld - L).max(), ' alpha relerr', np.abs(ad - a).max() / np.abs(a).max(),
      ' |Z L - I|', np.abs(Z @ L - np.eye(n)).max())
xc = np.random.RandomState(0).uniform(0, 1, (20000, d))
t0 = time.time()
val, idx, xb, mean, std, acq = h.sweep(xc, _lib.ACQ_POI, float(y.max()), dump=True)
t1 = time.time()
m, s = G.predict(xc[:4096], X, L, a, 0.8, G.KERNEL_MATERN, 1.5)
print(' sweep %.3fs kernel_ms %.3f' % (t1 - t0, h.last_sweep_stats()[0]))
print(' mean err', np.abs(mean[:4096] - m).max(), ' var err', np.abs(std[:4096] ** 2 - s ** 2).max(),
      ' argmax', idx, int(np.argmax(acq)), val)
lml, grad = h.lml([0.3, 0.8])
print(' lml', lml, grad, [G.log_marginal_likelihood(X, y, t, G.KERNEL_MATERN, 1.5) for t in (0.3, 0.8)])
'''
for impl, n, d in [(0, 100, 5), (1, 100, 5), (1, 600, 32), (1, 2048, 32), (1, 300, 70)]:
    try:
        r = subprocess.run([sys.executable, '-c', CODE, str(impl), str(n), str(d)], timeout=120,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        print(r.stdout[-3000:])
        print('exit', r.returncode)
    except subprocess.TimeoutExpired as e:
        print('TIMEOUT impl', impl, n, d, (e.stdout or b'')[-2000:])
    sys.stdout.flush()
