"""CPU oracle, part 2: the reference's suggestion path end to end, on the libraries it calls.

TEST INFRASTRUCTURE ONLY (same rule as ``gp_numpy.py``): imported by ``tests/``, ``smoke()`` and the
CPU-baseline / ``--impl reference`` legs of ``bench.py``; never by ``polyaxon_b200/``.

This restates the 349 lines of Python glue in
``polyaxon/hpsearch/search_managers/bayesian_optimization/{optimizer,acquisition_function}.py`` and
runs them on the very libraries the reference delegates to: ``sklearn.gaussian_process`` and
``scipy.optimize`` / ``scipy.stats`` as installed in this image (sklearn 1.9.0, scipy 1.18.1; the
reference pins 0.20.0 / 1.1.0, which are not installable here).  Because the installed scipy rejects
two idioms of the reference (a 2-D ``x0`` at acquisition_function.py:100 and ``res.fun[0]`` on a
scalar at :109,111) the restatement ravels ``x0`` and reads ``res.fun`` through ``np.ravel`` - the
same 3-line patch ``make_goldens.py`` applies to the real module.  A falsy seed is mapped to the
global ``RandomState`` object instead of the ``np.random`` module (sklearn >= 1.2 refuses the module).

Parity status: PINNED - ``tests/test_oracle_golden.py`` compares ``suggest()`` here with the
outputs the real reference produced in the build container (``tests/golden/*.npz``).
"""
import numpy as np
from scipy.optimize import minimize
from scipy.stats import norm
from sklearn.gaussian_process import GaussianProcessRegressor
from sklearn.gaussian_process.kernels import RBF, Matern

from . import gp_numpy as G


def make_rng(seed):
    """utils.py:37-38 (``RandomState(seed) if seed else np.random``)."""
    return np.random.RandomState(seed) if seed else np.random.mtrand._rand


def make_gp(kernel, length_scale, nu, n_restarts, rng):
    """acquisition_function.py:31-46."""
    k = RBF(length_scale=length_scale) if kernel == G.KERNEL_RBF else \
        Matern(length_scale=length_scale, nu=nu)
    return GaussianProcessRegressor(kernel=k, n_restarts_optimizer=n_restarts, random_state=rng)


def acquisition(gp, x, acq, y_max, kappa, eps):
    """acquisition_function.py:48-68; one ``predict`` per call, as in the reference."""
    mean, std = gp.predict(x, return_std=True)
    if acq == G.ACQ_UCB:
        return mean + kappa * std
    with np.errstate(divide='ignore', invalid='ignore'):
        z = (mean - y_max - eps) / std
        if acq == G.ACQ_EI:
            return (mean - y_max - eps) * norm.cdf(z) + std * norm.pdf(z)
        return norm.cdf(z)


def warmup(gp, rng, bounds, n_warmup, acq, y_max, kappa, eps, chunk=None):
    """acquisition_function.py:88-92.  ``chunk`` bounds the size of K* (the reference evaluates
    all candidates in one call, which needs n_warmup x n_obs doubles)."""
    x_tries = rng.uniform(bounds[:, 0], bounds[:, 1], size=(n_warmup, bounds.shape[0]))
    if chunk is None or chunk >= n_warmup:
        ys = acquisition(gp, x_tries, acq, y_max, kappa, eps)
    else:
        ys = np.concatenate([acquisition(gp, x_tries[s:s + chunk], acq, y_max, kappa, eps)
                             for s in range(0, n_warmup, chunk)])
    return x_tries, ys


def refine(gp, x_seeds, bounds, acq, y_max, kappa, eps, x_max, max_acq):
    """acquisition_function.py:97-111: L-BFGS-B per seed on -acq with finite-difference gradients;
    failed runs skipped; ``>=`` lets a later seed win a tie."""
    log = []
    for x_try in x_seeds:
        res = minimize(lambda x: -acquisition(gp, x.reshape(1, -1), acq, y_max, kappa, eps)[0],
                       x_try.ravel(), bounds=bounds, method='L-BFGS-B')
        val = -float(np.ravel(res.fun)[0])
        log.append((res.x.copy(), val, bool(res.success), int(res.nfev)))
        if not res.success:
            continue
        if max_acq is None or val >= max_acq:
            x_max, max_acq = res.x, val
    return x_max, max_acq, log


def suggest(X, y, bounds, kernel, length_scale, nu, n_restarts, acq, kappa, eps, seed,
            n_warmup, n_iter, chunk=None, trace=None):
    """optimizer.py:16-25 (`_maximize`): fit, warm-up sweep, refinement, clip.

    ``trace`` (a dict) receives every intermediate the staged parity tests compare."""
    rng = make_rng(seed)
    gp = make_gp(kernel, length_scale, nu, n_restarts, rng)
    y_max = np.max(y)
    gp.fit(X, y)
    bounds = np.asarray(bounds, dtype=np.float64)
    x_tries, ys = warmup(gp, rng, bounds, n_warmup, acq, y_max, kappa, eps, chunk)
    idx = int(ys.argmax())
    x_max, max_acq = x_tries[idx], ys.max()
    x_seeds = rng.uniform(bounds[:, 0], bounds[:, 1], size=(n_iter, bounds.shape[0]))
    x_best, best_acq, log = refine(gp, x_seeds, bounds, acq, y_max, kappa, eps, x_max, max_acq)
    out = np.clip(x_best, bounds[:, 0], bounds[:, 1])
    if trace is not None:
        trace.update(theta=float(gp.kernel_.theta[0]), lml=float(gp.log_marginal_likelihood_value_),
                     L=gp.L_, alpha=gp.alpha_, x_tries=x_tries, ys=ys, argmax=idx,
                     warm_acq=float(max_acq), x_seeds=x_seeds, refine_log=log,
                     best_acq=float(best_acq), gp=gp)
    return out


def synthetic_problem(n, d, seed=1234):
    """The synthetic observation set of SURVEY.md section 8d / BASELINE.md section 2."""
    rs = np.random.RandomState(seed)
    X = rs.uniform(0.0, 1.0, (n, d))
    rest = X[:, 1:].mean(axis=1) if d > 1 else 0.0
    y = np.sin(3.0 * X[:, 0]) + rest + 0.01 * rs.randn(n)
    bounds = np.tile(np.array([[0.0, 1.0]]), (d, 1))
    return X, y, bounds
