"""ctypes binding of libplxbo.so (the C ABI declared in include/plxbo.h).

There is no CPU path: if the shared library has not been built (``python __graft_entry__.py`` or
``python -m polyaxon_b200.build``) or no CUDA device is present, the calls below raise.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_longlong, c_void_p

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# PLX_BO_LIB: another build of the same library (tools/*_phases.py load the -DPLX_PROFILE build through it)
LIB_PATH = os.environ.get('PLX_BO_LIB') or os.path.join(HERE, 'libplxbo.so')

PLX_OK, PLX_ERR_INVALID, PLX_ERR_CUDA, PLX_ERR_NOT_PD, PLX_ERR_STATE, PLX_ERR_UNSUPPORTED = range(6)
KERNEL_RBF, KERNEL_MATERN = 0, 1
ACQ_UCB, ACQ_EI, ACQ_POI = 0, 1, 2

_dp = POINTER(c_double)
FG_CALLBACK = ctypes.CFUNCTYPE(None, c_double, _dp, _dp, c_void_p)

# every symbol include/plxbo.h declares: name -> (restype, argtypes)
SIGNATURES = {
    'plx_bo_create': (c_int, [c_int, POINTER(c_void_p)]),
    'plx_bo_destroy': (None, [c_void_p]),
    'plx_last_error': (c_char_p, [c_void_p]),
    'plx_version': (c_char_p, []),
    'plx_bo_set_stream': (c_int, [c_void_p, c_void_p]),
    'plx_bo_set_option': (c_int, [c_void_p, c_char_p, c_longlong]),
    'plx_bo_reserve': (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_longlong]),
    'plx_gp_set_data': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_double, c_double]),
    'plx_gp_lml': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    'plx_gp_fit': (c_int, [c_void_p, c_double, c_void_p, c_int, c_double, c_double, _dp, _dp, c_void_p]),
    'plx_gp_set_theta': (c_int, [c_void_p, c_double]),
    'plx_gp_get_state': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    'plx_gp_append': (c_int, [c_void_p, c_void_p, c_double]),
    'plx_acq_sweep': (c_int, [c_void_p, c_void_p, c_longlong, c_longlong, c_int, c_double, c_double, c_double,
                              _dp, POINTER(c_longlong), c_void_p, c_void_p, c_void_p, c_void_p]),
    'plx_acq_sweep_dev': (c_int, [c_void_p, c_void_p, c_longlong, c_longlong, c_int, c_double, c_double,
                                  c_double, _dp, POINTER(c_longlong), c_void_p]),
    'plx_candidates_set': (c_int, [c_void_p, c_void_p, c_longlong, c_int]),
    'plx_candidates_draw_mt19937': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_longlong, c_int,
                                            c_void_p, POINTER(c_int)]),
    'plx_candidates_draw_mt19937_range': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_longlong, c_int,
                                                  c_longlong, c_longlong, c_void_p, POINTER(c_int)]),
    'plx_candidates_get': (c_int, [c_void_p, c_void_p, c_longlong, c_longlong]),
    'plx_acq_sweep_resident': (c_int, [c_void_p, c_longlong, c_int, c_double, c_double, c_double, _dp,
                                       POINTER(c_longlong), c_void_p]),
    'plx_kernel_values': (c_int, [c_void_p, c_int, c_double, c_void_p, c_int, c_void_p]),
    'plx_acq_sweep_resident_range': (c_int, [c_void_p, c_longlong, c_longlong, c_int, c_double, c_double, c_double,
                                             _dp, POINTER(c_longlong), c_void_p]),
    'plx_last_sweep_stats': (c_int, [c_void_p, _dp, POINTER(c_longlong)]),
    'plx_last_sweep_clamped': (c_int, [c_void_p, POINTER(c_longlong)]),
    'plx_measure_i8_peak': (c_int, [c_void_p, _dp]),
    'plx_launch_count': (c_longlong, [c_void_p]),
    'plx_acq_refine': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_double, c_double, c_double, c_int,
                               c_void_p, c_void_p, c_void_p, c_void_p]),
    'plx_acq_value_grad': (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_double, c_double, c_void_p,
                                   c_void_p]),
    'plx_acq_sweep_noalloc': (c_int, [c_int, c_void_p, c_void_p, c_longlong, c_longlong, c_void_p, c_void_p,
                                      c_void_p, c_int, c_int, c_int, c_int, c_double, c_double, c_int,
                                      c_double, c_double, c_double, c_void_p, c_void_p]),
    'plx_lbfgsb1d_minimize': (c_int, [FG_CALLBACK, c_void_p, c_double, c_double, c_double, _dp, _dp,
                                      POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
}

_lib = None


class PlxError(RuntimeError):
    def __init__(self, code, message):
        super().__init__('plxbo error {}: {}'.format(code, message))
        self.code = code


def load():
    """Load the shared library (once) and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            '{} is missing: build the CUDA extension first (python -c "import __graft_entry__ as g; '
            'g.build()"). polyaxon_b200 has no CPU fallback.'.format(LIB_PATH))
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Handle(object):
    """One GP + sweep context on one GPU.  Thin, 1:1 with the C entry points."""

    def __init__(self, device=0):
        self.lib = load()
        self._h = c_void_p()
        rc = self.lib.plx_bo_create(int(device), ctypes.byref(self._h))
        if rc != PLX_OK:
            msg = self.lib.plx_last_error(None).decode()
            self._h = None
            raise PlxError(rc, msg)
        self.device = int(device)
        self.n = self.d = 0
        self.resident_key = None
        self.owner = None         # (id, fit count) of the DeviceGaussianProcess whose state this handle holds
        self.held = None          # host-side record of the last fit (DeviceGaussianProcess, PLX_BO_REFIT_EVERY)
        self.seed_cache = {}      # resident_key -> (x_seeds, generator state after the draws), UtilityFunction._warmup

    def close(self):
        if getattr(self, '_h', None):
            self.lib.plx_bo_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc == PLX_OK:
            return
        msg = self.lib.plx_last_error(self._h).decode()
        if rc == PLX_ERR_NOT_PD:
            raise np.linalg.LinAlgError(msg)
        raise PlxError(rc, msg)

    def set_stream(self, cuda_stream):
        self._check(self.lib.plx_bo_set_stream(self._h, c_void_p(cuda_stream or 0)))

    def set_option(self, name, value):
        self._check(self.lib.plx_bo_set_option(self._h, name.encode(), int(value)))

    def reserve(self, n, d, runs=1, seeds=0, candidates=0):
        """Allocate the workspaces of an (n, d) problem now, so the first call of a group carries no cudaMalloc."""
        self._check(self.lib.plx_bo_reserve(self._h, int(n), int(d), int(runs), int(seeds), int(candidates)))

    def set_data(self, X, y, kernel, nu=1.5, jitter=1e-10):
        X = _f64(X)
        y = _f64(y)
        if X.ndim != 2 or y.shape != (X.shape[0],):
            raise ValueError('X must be (n, d) and y (n,)')
        self.n, self.d = X.shape
        self.held = None
        self.owner = None
        self._check(self.lib.plx_gp_set_data(self._h, _ptr(X), _ptr(y), self.n, self.d, int(kernel),
                                             float(nu), float(jitter)))

    def lml(self, thetas, grad=True):
        thetas = _f64(np.atleast_1d(thetas))
        out = np.empty_like(thetas)
        g = np.empty_like(thetas) if grad else None
        self._check(self.lib.plx_gp_lml(self._h, _ptr(thetas), thetas.size, _ptr(out), _ptr(g)))
        return (out, g) if grad else out

    def fit(self, theta0, theta_restarts=(), bounds=(np.log(1e-5), np.log(1e5)), apply=True):
        """All optimiser runs in lock step.  apply=False returns the runs only and leaves the posterior state
        alone (a rank of a sharded fit: the state is built at the winner of all ranks' runs)."""
        self.set_option('fit_apply', 1 if apply else 0)
        rs = _f64(np.atleast_1d(np.asarray(theta_restarts, dtype=np.float64)))
        per_run = np.empty((rs.size + 1, 4))
        th, lml = c_double(), c_double()
        self._check(self.lib.plx_gp_fit(self._h, float(theta0), _ptr(rs) if rs.size else None, rs.size,
                                        float(bounds[0]), float(bounds[1]), ctypes.byref(th),
                                        ctypes.byref(lml), _ptr(per_run)))
        return th.value, lml.value, per_run

    def set_theta(self, theta):
        self.held = None
        self._check(self.lib.plx_gp_set_theta(self._h, float(theta)))

    def append(self, x_new, y_new):
        """One more observation at the current theta (bordered Cholesky, O(n^2))."""
        x_new = _f64(x_new).ravel()
        if x_new.shape[0] != self.d:
            raise ValueError('x_new must have {} entries'.format(self.d))
        self._check(self.lib.plx_gp_append(self._h, _ptr(x_new), float(y_new)))
        self.n += 1

    def get_state(self, want_L=True, want_Linv=False):
        L = np.empty((self.n, self.n)) if want_L else None
        Z = np.empty((self.n, self.n)) if want_Linv else None
        a = np.empty(self.n)
        self._check(self.lib.plx_gp_get_state(self._h, _ptr(L), _ptr(a), _ptr(Z)))
        return L, a, Z

    def sweep(self, Xc, acq, y_max, kappa=0.0, eps=0.0, idx_base=0, dump=False):
        """Host candidates (N x d).  Returns (best_value, best_index, best_x[, mean, std, acq])."""
        Xc = _f64(Xc)
        if Xc.ndim != 2 or Xc.shape[1] != self.d:
            raise ValueError('candidates must be (N, {})'.format(self.d))
        N = Xc.shape[0]
        bv, bi = c_double(), c_longlong()
        bx = np.empty(self.d)
        outs = [np.empty(N) for _ in range(3)] if dump else [None] * 3
        self._check(self.lib.plx_acq_sweep(self._h, _ptr(Xc), N, int(idx_base), int(acq), float(kappa),
                                           float(eps), float(y_max), ctypes.byref(bv), ctypes.byref(bi),
                                           _ptr(bx), _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2])))
        if dump:
            return bv.value, bi.value, bx, outs[0], outs[1], outs[2]
        return bv.value, bi.value, bx

    def sweep_dev(self, xc_ptr, N, acq, y_max, kappa=0.0, eps=0.0, idx_base=0):
        """Candidates resident on the device (raw pointer, e.g. torch tensor.data_ptr())."""
        bv, bi = c_double(), c_longlong()
        bx = np.empty(self.d)
        self._check(self.lib.plx_acq_sweep_dev(self._h, c_void_p(xc_ptr), int(N), int(idx_base), int(acq),
                                               float(kappa), float(eps), float(y_max), ctypes.byref(bv),
                                               ctypes.byref(bi), _ptr(bx)))
        return bv.value, bi.value, bx

    def refine(self, seeds, bounds, acq, y_max, kappa=0.0, eps=0.0, max_iter=0, want_nfev=False):
        """Batched bounded quasi-Newton refinement of every seed.  Returns (x[S,d], acq[S], success[S])."""
        seeds = _f64(np.atleast_2d(seeds))
        bounds = _f64(bounds)
        S = seeds.shape[0]
        if seeds.shape[1] != self.d or bounds.shape != (self.d, 2):
            raise ValueError('seeds must be (S, {0}) and bounds ({0}, 2)'.format(self.d))
        x = np.empty((S, self.d))
        vals = np.empty(S)
        ok = np.zeros(S, dtype=np.int32)
        nfev = np.zeros(S, dtype=np.int32)
        self._check(self.lib.plx_acq_refine(self._h, _ptr(seeds), S, _ptr(bounds), int(acq), float(kappa),
                                            float(eps), float(y_max), int(max_iter), _ptr(x), _ptr(vals),
                                            _ptr(ok), _ptr(nfev)))
        if want_nfev:
            return x, vals, ok.astype(bool), nfev
        return x, vals, ok.astype(bool)

    def acq_value_grad(self, points, acq, y_max, kappa=0.0, eps=0.0):
        pts = _f64(np.atleast_2d(points))
        vals = np.empty(pts.shape[0])
        grads = np.empty_like(pts)
        self._check(self.lib.plx_acq_value_grad(self._h, _ptr(pts), pts.shape[0], int(acq), float(kappa),
                                                float(eps), float(y_max), _ptr(vals), _ptr(grads)))
        return vals, grads

    def candidates_set(self, Xc, key=None):
        """Keep a device-resident copy of the candidate matrix; `key` is remembered for the caller."""
        Xc = _f64(Xc)
        self.resident_key = None
        self._check(self.lib.plx_candidates_set(self._h, _ptr(Xc), Xc.shape[0], Xc.shape[1]))
        self.resident_key = key

    def candidates_draw(self, rng, low, high, N, key=None, first=0, count=None):
        """Draw rng.uniform(low, high, size=(N, d)) on the device (bit-identical to numpy's MT19937
        stream) into the resident matrix and move `rng` past the WHOLE draw.  With first / count only
        rows [first, first + count) are produced and held (a rank's block of a sharded sweep).  Returns
        False -- nothing drawn, nothing changed -- when the generator is not an MT19937 at an even
        position."""
        state = rng.get_state()
        if state[0] != 'MT19937' or int(state[2]) & 1:
            return False
        low = _f64(low)
        rng_range = _f64(np.subtract(high, low))
        key_in = np.ascontiguousarray(state[1], dtype=np.uint32)
        key_out = np.empty(624, dtype=np.uint32)
        pos_out = c_int()
        self.resident_key = None
        count = int(N) - int(first) if count is None else int(count)
        self._check(self.lib.plx_candidates_draw_mt19937_range(self._h, _ptr(key_in), int(state[2]), _ptr(low),
                                                               _ptr(rng_range), int(N), low.size, int(first), count,
                                                               _ptr(key_out), ctypes.byref(pos_out)))
        rng.set_state((state[0], key_out, pos_out.value) + tuple(state[3:]))
        self.resident_key = key
        return True

    def candidates_get(self, first, count, d, out=None):
        """Rows [first, first + count) (global row numbers) of the resident block; `out` may be a
        C-contiguous float64 array to fill (e.g. the numpy view of a pinned buffer)."""
        if out is None:
            out = np.empty((count, d))
        if count == 0:
            return out
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.size == count * d
        self._check(self.lib.plx_candidates_get(self._h, _ptr(out), int(first), int(count)))
        return out

    def sweep_resident(self, acq, y_max, kappa=0.0, eps=0.0, idx_base=0):
        bv, bi = c_double(), c_longlong()
        bx = np.empty(self.d)
        self._check(self.lib.plx_acq_sweep_resident(self._h, int(idx_base), int(acq), float(kappa), float(eps),
                                                    float(y_max), ctypes.byref(bv), ctypes.byref(bi), _ptr(bx)))
        return bv.value, bi.value, bx

    def kernel_values(self, kernel, nu, r2):
        r2 = _f64(np.ravel(r2))
        out = np.empty_like(r2)
        self._check(self.lib.plx_kernel_values(self._h, int(kernel), float(nu), _ptr(r2), r2.size, _ptr(out)))
        return out

    def sweep_resident_range(self, first, count, acq, y_max, kappa=0.0, eps=0.0):
        bv, bi = c_double(), c_longlong()
        bx = np.empty(self.d)
        self._check(self.lib.plx_acq_sweep_resident_range(self._h, int(first), int(count), int(acq), float(kappa),
                                                          float(eps), float(y_max), ctypes.byref(bv),
                                                          ctypes.byref(bi), _ptr(bx)))
        return bv.value, bi.value, bx

    def last_sweep_stats(self):
        ms, nl = c_double(), c_longlong()
        self._check(self.lib.plx_last_sweep_stats(self._h, ctypes.byref(ms), ctypes.byref(nl)))
        return ms.value, nl.value

    def last_sweep_clamped(self):
        """Candidates of the last sweep whose negative variance was set to zero."""
        n = c_longlong()
        self._check(self.lib.plx_last_sweep_clamped(self._h, ctypes.byref(n)))
        return n.value

    def measure_i8_peak(self):
        tops = c_double()
        self._check(self.lib.plx_measure_i8_peak(self._h, ctypes.byref(tops)))
        return tops.value

    def launch_count(self):
        return int(self.lib.plx_launch_count(self._h))


def lbfgsb1d_minimize(fun, x0, lo, hi):
    """Host-only: run the library's one-variable L-BFGS-B on a Python callable returning (f, g)."""
    lib = load()

    def trampoline(x, fp, gp, _user):
        f, g = fun(x)
        fp[0] = f
        gp[0] = g

    cb = FG_CALLBACK(trampoline)
    x, f = c_double(), c_double()
    ne, ni, st = c_int(), c_int(), c_int()
    rc = lib.plx_lbfgsb1d_minimize(cb, None, float(x0), float(lo), float(hi), ctypes.byref(x), ctypes.byref(f),
                                   ctypes.byref(ne), ctypes.byref(ni), ctypes.byref(st))
    if rc != PLX_OK:
        raise PlxError(rc, 'lbfgsb1d_minimize: invalid arguments')
    return x.value, f.value, ne.value, ni.value, st.value


def sweep_noalloc(device, stream, xc_ptr, N, x_ptr, alpha_ptr, linv_ptr, ldz, n, d, kernel, nu, theta, acq, y_max,
                  best_value_ptr, best_index_ptr, kappa=0.0, eps=0.0, idx_base=0):
    """plx_acq_sweep_noalloc: every pointer is a raw device address owned by the caller; nothing is
    allocated and nothing is synchronised (results land in the two device cells on `stream`)."""
    lib = load()
    rc = lib.plx_acq_sweep_noalloc(int(device), c_void_p(stream or 0), c_void_p(xc_ptr), int(N), int(idx_base),
                                   c_void_p(x_ptr), c_void_p(alpha_ptr), c_void_p(linv_ptr), int(ldz), int(n),
                                   int(d), int(kernel), float(nu), float(theta), int(acq), float(kappa),
                                   float(eps), float(y_max), c_void_p(best_value_ptr), c_void_p(best_index_ptr))
    if rc != PLX_OK:
        raise PlxError(rc, lib.plx_last_error(None).decode())
