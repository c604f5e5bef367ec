import sys, time
import numpy as np
sys.path.insert(0, '.')
from polyaxon_b200 import _lib
h = _lib.Handle(0)
for N, d in [(100000, 2), (1000000, 8), (1 << 24, 32)]:
    lo, hi = np.zeros(d), np.ones(d)
    r = np.random.RandomState(5)
    t0 = time.perf_counter(); assert h.candidates_draw(r, lo, hi, N); t1 = time.perf_counter()
    r2 = np.random.RandomState(5)
    t2 = time.perf_counter(); x = r2.uniform(lo, hi, size=(min(N, 1 << 21), d)); t3 = time.perf_counter()
    print('N=%d d=%d device %.1f ms | numpy %.1f ms for %d rows' % (N, d, (t1 - t0) * 1e3, (t3 - t2) * 1e3, x.shape[0]))
    assert np.array_equal(h.candidates_get(0, 1000, d), x[:1000])
