"""``UtilityFunction``: Gaussian-process surrogate + UCB / EI / POI acquisition, on the GPU.

Same constructor, attributes and methods as the reference class
(``.../bayesian_optimization/acquisition_function.py:17-115``); the arithmetic the reference hands
to scikit-learn / scipy runs in the CUDA library behind ``polyaxon_b200._lib``:

  reference                                              here
  -----------------------------------------------------  ------------------------------------------
  GaussianProcessRegressor(kernel, n_restarts, rng)      DeviceGaussianProcess (same attribute names)
  gp.fit(X, y)                    optimizer.py:21        plx_gp_set_data + plx_gp_fit
  gp.predict(x, return_std=True)  :49,53,58              plx_acq_sweep (mean / std dumps)
  compute(x, y_max)               :62-68                 plx_acq_sweep (acquisition dump)
  max_compute warm-up             :88-92                 plx_acq_sweep (arg-max on device)
  max_compute L-BFGS-B loop       :95-111                plx_acq_refine (all seeds batched)

Random numbers are drawn on the host from the reference's own ``RandomState`` in the reference's
order (restart thetas during fit -> x_tries -> x_seeds), so a seeded group sees bit-identical
candidate matrices.
"""
import hashlib
import os

import numpy as np

from polyaxon_b200 import _lib, device, dist as pdist
from polyaxon_b200.hpsearch.search_managers.utils import get_random_generator
from polyaxon_b200.schemas.hptuning import (
    AcquisitionFunctions,
    GaussianProcessConfig,
    GaussianProcessesKernels,
    UtilityFunctionConfig,
)

THETA_BOUNDS = (float(np.log(1e-5)), float(np.log(1e5)))      # sklearn's default length_scale_bounds


def _warn_clamped(handle):
    """sklearn warns when predicted variances come out negative and sets them to zero
    (_gpr.py:485-491, reached from acquisition_function.py:49); the kernels clamp and count, the warning is
    raised here with sklearn's text."""
    if handle.last_sweep_clamped() > 0:
        import warnings
        warnings.warn('Predicted variances smaller than 0. Setting those variances to 0.')


def _acq_id(name):
    if AcquisitionFunctions.is_ucb(name):
        return _lib.ACQ_UCB
    if AcquisitionFunctions.is_ei(name):
        return _lib.ACQ_EI
    if AcquisitionFunctions.is_poi(name):
        return _lib.ACQ_POI
    return None


class DeviceGaussianProcess(object):
    """The slice of ``sklearn.gaussian_process.GaussianProcessRegressor`` the reference touches:
    ``fit`` / ``predict`` plus the fitted attributes ``theta_``, ``length_scale_``,
    ``log_marginal_likelihood_value_``, ``L_``, ``alpha_``.  alpha (jitter) is sklearn's default
    1e-10 and y is not normalised, as in the reference (acquisition_function.py:42-46)."""

    alpha = 1e-10

    def __init__(self, kernel, length_scale, nu, n_restarts_optimizer, random_state, handle=None):
        self.kernel_id = _lib.KERNEL_RBF if GaussianProcessesKernels.is_rbf(kernel) else _lib.KERNEL_MATERN
        self.length_scale = float(length_scale)
        self.nu = float(nu) if nu is not None else 1.5
        self.n_restarts_optimizer = int(n_restarts_optimizer or 0)
        self.random_state = random_state
        self._handle = handle
        self.fitted = False
        self._fit_count = 0

    @property
    def handle(self):
        if self._handle is None:
            self._handle = device.get_handle()
        return self._handle

    def fit(self, X, y):
        X = np.asarray(X, dtype=np.float64)
        y = np.asarray(y, dtype=np.float64)
        if X.ndim != 2 or X.shape[0] != y.shape[0]:
            raise ValueError('Found input variables with inconsistent numbers of samples: '
                             '{} and {}'.format(X.shape, y.shape))
        # sklearn draws one theta per restart from the shared generator, before any candidate
        # (sklearn _gpr.py:329-333); same calls here so the stream stays aligned
        lo, hi = np.array([THETA_BOUNDS[0]]), np.array([THETA_BOUNDS[1]])
        restarts = [float(self.random_state.uniform(lo, hi)[0]) for _ in range(self.n_restarts_optimizer)]
        if self._append_instead_of_refit(X, y):
            return self
        self.handle.set_data(X, y, self.kernel_id, self.nu, self.alpha)
        theta, lml, per_run = self._fit_runs(np.log(self.length_scale), restarts)
        self.theta_ = theta
        self.length_scale_ = float(np.exp(theta))
        self.log_marginal_likelihood_value_ = lml
        self.optimizer_runs_ = per_run
        self.X_train_, self.y_train_ = X, y
        self.fitted = True
        self._stamp()
        self.handle.held = dict(X=X.copy(), y=y.copy(), kernel=(self.kernel_id, self.nu, self.length_scale,
                                                                self.n_restarts_optimizer),
                                theta=theta, lml=lml, since_fit=0)
        return self

    # The handle (observations, factor, posterior state on the GPU) is shared by every model of a thread
    # (device.get_handle); in the reference each UtilityFunction owns its GaussianProcessRegressor.  The stamp
    # says whose state the handle holds; a model that finds somebody else's puts its own back before it is used.
    def _stamp(self):
        self._fit_count += 1
        self.handle.owner = (id(self), self._fit_count)

    def ensure_state(self):
        """Make the handle hold THIS model's observations and posterior state (a no-op for the usual
        one-optimizer-per-call flow).  Raises when the model was never fitted."""
        if not self.fitted:
            raise RuntimeError('this Gaussian process has not been fitted: call fit(X, y) first')
        h = self.handle
        if getattr(h, 'owner', None) != (id(self), self._fit_count):
            h.set_data(self.X_train_, self.y_train_, self.kernel_id, self.nu, self.alpha)
            h.set_theta(self.theta_)
            h.owner = (id(self), self._fit_count)
        return h

    def _append_instead_of_refit(self, X, y):
        """Opt-in (PLX_BO_REFIT_EVERY = k > 1; the default 1 is the reference's behaviour, a full
        hyper-parameter fit on every call, optimizer.py:21): when this call's observations are the
        previous call's plus new rows at the end, and fewer than k calls have passed since the last
        fit, keep theta and extend the posterior state row by row on the device (plx_gp_append)."""
        every = int(os.environ.get('PLX_BO_REFIT_EVERY', '1') or 1)
        held = getattr(self.handle, 'held', None)
        if every <= 1 or held is None:          # `held` is dropped whenever anybody replaces the handle's data or theta
            return False
        n0 = held['X'].shape[0]
        same_model = held['kernel'] == (self.kernel_id, self.nu, self.length_scale, self.n_restarts_optimizer)
        if (not same_model or held['since_fit'] + 1 >= every or X.shape[0] <= n0 or X.shape[1] != held['X'].shape[1]
                or self.handle.n != n0 or not np.array_equal(X[:n0], held['X']) or not np.array_equal(y[:n0], held['y'])):
            return False
        try:
            for i in range(n0, X.shape[0]):
                self.handle.append(X[i], y[i])
        except np.linalg.LinAlgError:
            self.handle.held = None                  # partially extended: let the caller refit from scratch
            return False
        held.update(X=X.copy(), y=y.copy(), since_fit=held['since_fit'] + 1)
        self.theta_ = held['theta']
        self.length_scale_ = float(np.exp(held['theta']))
        self.log_marginal_likelihood_value_ = None   # not evaluated on this path
        self.optimizer_runs_ = None
        self.X_train_, self.y_train_ = X, y
        self.fitted = True
        self._stamp()
        return True

    def _fit_runs(self, theta0, restarts):
        """All optimiser runs (the configured length scale first, then the restarts).  Under
        torch.distributed the runs are dealt round-robin to the ranks, each rank advances its share in
        lock step on its GPU, one all-gather of {theta*, LML*} per run follows, and every rank rebuilds
        the posterior state at the winning theta locally (cheaper than shipping the n x n factor).
        The winner is the first maximum of LML in run order, as numpy.argmin over sklearn's list of
        -LML values (sklearn _gpr.py:336-337)."""
        rank, size = pdist.world()
        starts = [float(theta0)] + [float(t) for t in restarts]
        if size == 1 or len(starts) == 1 or os.environ.get('PLX_BO_DISTRIBUTED', '1') == '0':
            return self.handle.fit(theta0, restarts, THETA_BOUNDS)
        mine = starts[rank::size]
        per = (len(starts) + size - 1) // size
        pack = np.full((per, 4), np.nan)
        if mine:
            # this rank's runs only: the posterior state is built below, at the winner of ALL runs.  (With the
            # state built here a rank whose starts all give a non-positive-definite kernel matrix would raise
            # before the collective and leave the other ranks waiting in it.)
            part = self.handle.fit(mine[0], mine[1:], THETA_BOUNDS, apply=False)[2]
            pack[:len(part)] = part
        out = pdist.gather_rows(pack.ravel(), device=pdist.collective_device(self.handle.device))
        out = out.reshape(size, per, 4)
        per_run = np.array([out[i % size, i // size] for i in range(len(starts))])
        # sklearn: np.argmin over the list of -LML values (first minimum; -inf LML = +inf there, _gpr.py:336-337)
        best = int(np.argmin(-per_run[:, 2]))
        theta = float(per_run[best, 1])
        self.handle.set_theta(theta)            # LinAlgError on every rank alike when no run was positive definite
        return theta, float(per_run[best, 2]), per_run

    @property
    def L_(self):
        return self.handle.get_state(want_L=True)[0]

    @property
    def alpha_(self):
        return self.handle.get_state(want_L=False)[1]

    def predict(self, X, return_std=False):
        h = self.ensure_state()
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        _, _, _, mean, std, _ = h.sweep(X, _lib.ACQ_UCB, 0.0, kappa=0.0, dump=True)
        _warn_clamped(h)
        return (mean, std) if return_std else mean


class UtilityFunction(object):
    def __init__(self, config, seed=None):
        if not isinstance(config, UtilityFunctionConfig):
            raise ValueError('Received a non valid configuration.')
        self.config = config
        self.eps = config.eps
        self.kappa = config.kappa
        self.acquisition_function = config.acquisition_function
        self.random_generator = get_random_generator(seed=seed)
        self.gaussian_process = self.get_gaussian_process(config=config.gaussian_process,
                                                          random_generator=self.random_generator)

    @staticmethod
    def get_gaussian_process(config, random_generator):
        if not isinstance(config, GaussianProcessConfig):
            raise ValueError('Received a non valid configuration.')
        return DeviceGaussianProcess(kernel=config.kernel, length_scale=config.length_scale,
                                     nu=config.nu, n_restarts_optimizer=config.n_restarts_optimizer,
                                     random_state=random_generator)

    def _acq_args(self, y_max):
        acq = _acq_id(self.acquisition_function)
        return acq, dict(y_max=float(y_max if y_max is not None else 0.0),
                         kappa=float(self.kappa or 0.0), eps=float(self.eps or 0.0))

    def _warmup(self, h, acq, args, bounds, n_warmup, n_iter):
        """Sampling stage.  The draws depend only on the generator state on entry, the bounds and
        the two counts; a seeded group re-creates the same state on every call (manager.py:30), so
        the candidate matrix of the previous call is still in device memory and neither the
        numpy draw nor the host->device copy is repeated.  On a hit the generator is moved to the
        state it would have reached, so everything downstream sees the reference's stream.

        Under torch.distributed every rank has the same generator state; each draws and keeps only its
        own contiguous block of rows of that one matrix (the device jumps ahead in the MT19937 stream,
        csrc/mt19937.cuh) and all ranks leave with the state after the whole matrix."""
        rng = self.random_generator
        d = bounds.shape[0]
        rank, size = pdist.world()
        if os.environ.get('PLX_BO_DISTRIBUTED', '1') == '0':
            rank, size = 0, 1
        lo, hi = pdist.shard_range(n_warmup, rank, size)
        key = None
        if os.environ.get('PLX_BO_CANDIDATE_CACHE', '1') != '0' and hasattr(rng, 'get_state'):
            state = rng.get_state()
            digest = hashlib.sha1(state[1].tobytes())
            digest.update(repr((state[0], state[2], state[3], state[4], n_warmup, n_iter, rank, size)).encode())
            digest.update(bounds.tobytes())
            key = digest.hexdigest()
        cached = h.seed_cache.get(key) if key is not None else None
        if cached is not None and h.resident_key == key:
            x_seeds, state_after = cached
            rng.set_state(state_after)
            max_acq, x_max = self._sweep_resident(h, acq, args, lo, hi, size)
            return max_acq, x_max, x_seeds.copy()
        drawn_on_device = False
        if key is not None and os.environ.get('PLX_BO_DEVICE_RNG', '1') != '0':
            # numpy's MT19937 stream regenerated on the GPU, bit for bit; rng is moved past the whole matrix
            drawn_on_device = h.candidates_draw(rng, bounds[:, 0], bounds[:, 1], n_warmup, key=key,
                                                first=lo, count=hi - lo)
        if not drawn_on_device:
            x_tries = rng.uniform(bounds[:, 0], bounds[:, 1], size=(n_warmup, d))
        x_seeds = rng.uniform(bounds[:, 0], bounds[:, 1], size=(n_iter, d))
        if key is None:
            if size == 1:
                max_acq, _, x_max = h.sweep(x_tries, acq, **args)
            else:
                value, index, x = h.sweep(x_tries[lo:hi], acq, idx_base=lo, **args)
                max_acq, _, x_max = pdist.merge_best_x(value, index if hi > lo else -1, x,
                                                       device=pdist.collective_device(h.device))
            return max_acq, x_max, x_seeds
        if not drawn_on_device:
            h.candidates_set(x_tries, key=key)       # host draw (odd stream position): the whole matrix, rows lo..hi swept
        h.seed_cache.clear()                     # one resident matrix per handle
        h.seed_cache[key] = (x_seeds.copy(), rng.get_state())
        max_acq, x_max = self._sweep_resident(h, acq, args, lo, hi, size)
        return max_acq, x_max, x_seeds

    @staticmethod
    def _sweep_resident(h, acq, args, lo, hi, size):
        """Sweep of the resident rows.  With several ranks each sweeps its block and ONE all-gather of
        {value, global index, row} records merges the bests with numpy.argmax order on every rank."""
        if size == 1:
            value, _, x = h.sweep_resident_range(lo, hi - lo, acq, **args)
            return value, x
        value, index, x = h.sweep_resident_range(lo, hi - lo, acq, **args)
        value, _, x = pdist.merge_best_x(value, index if hi > lo else -1, x, device=pdist.collective_device(h.device))
        return value, x

    def compute(self, x, y_max):
        """Acquisition values of the rows of x (acquisition_function.py:62-68)."""
        acq, args = self._acq_args(y_max)
        if acq is None:
            return None
        x = np.atleast_2d(np.asarray(x, dtype=np.float64))
        h = self.gaussian_process.ensure_state()
        out = h.sweep(x, acq, dump=True, **args)
        _warn_clamped(h)
        return out[5]

    def max_compute(self, y_max, bounds, n_warmup=100000, n_iter=250):
        """arg max of the acquisition function (acquisition_function.py:70-115): `n_warmup` random
        candidates swept on the device, then `n_iter` bounded quasi-Newton refinements, batched."""
        acq, args = self._acq_args(y_max)
        bounds = np.asarray(bounds, dtype=np.float64)
        h = self.gaussian_process.ensure_state()
        max_acq, x_max, x_seeds = self._warmup(h, acq, args, bounds, n_warmup, n_iter)
        _warn_clamped(h)
        if n_iter > 0:
            rank, size = pdist.world()
            if size > 1 and os.environ.get('PLX_BO_DISTRIBUTED', '1') != '0':
                # seeds are dealt round-robin to the ranks, results gathered back in seed order
                xs, values, success = h.refine(x_seeds[rank::size], bounds, acq, **args)
                xs, values, success = pdist.gather_refined(xs, values, success, n_iter,
                                                           device=pdist.collective_device(h.device))
            else:
                xs, values, success = h.refine(x_seeds, bounds, acq, **args)
            for x_ref, value, ok in zip(xs, values, success):
                if not ok:
                    continue
                # `>=`: among equal values the later seed wins (acquisition_function.py:109)
                if max_acq is None or value >= max_acq:
                    x_max, max_acq = x_ref, value
        self.last_max_acq_ = max_acq
        return np.clip(x_max, bounds[:, 0], bounds[:, 1])
