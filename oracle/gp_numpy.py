"""CPU oracle, part 1: the Gaussian-process arithmetic of the suggestion path, restated in numpy.

TEST INFRASTRUCTURE ONLY.  Nothing in ``polyaxon_b200/`` may import this module; it is used by
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` as the checker, never as the product path.

The reference (``polyaxon/hpsearch/search_managers/bayesian_optimization/acquisition_function.py``)
contains no arithmetic of its own: it delegates to scikit-learn (pinned 0.20.0, installed 1.9.0)
and scipy (pinned 1.1.0, installed 1.18.1), neither of which is vendored under ``/root/reference``.
Every function below restates the published algorithm of the installed scikit-learn 1.9.0 (the
oracle of record) and cites the call site in the reference that reaches it.

Parity status: PINNED against outputs of the real reference run in the build container
(``oracle/make_goldens.py`` imports ``/root/reference`` unmodified apart from a 3-line scipy
compatibility patch and writes ``tests/golden/*.npz``); ``tests/test_oracle_golden.py`` checks every
function here against those files.  The reference's own tests pin no number on this path
(``tests/test_experiment_groups/test_search_managers.py:888-1070`` assert bounds membership only).
"""
import math

import numpy as np
from scipy.linalg import cho_solve, cholesky, solve_triangular
from scipy.spatial.distance import cdist, pdist, squareform
from scipy.special import erf, erfc, gamma, kv

KERNEL_RBF = 0
KERNEL_MATERN = 1

ACQ_UCB = 0
ACQ_EI = 1
ACQ_POI = 2

# sklearn defaults that the reference never overrides (acquisition_function.py:42-46)
GPR_ALPHA = 1e-10
THETA_BOUNDS = (math.log(1e-5), math.log(1e5))


def kernel_matrix(X, theta, kernel=KERNEL_MATERN, nu=1.5, Y=None, eval_gradient=False):
    """k(X, Y) for the isotropic RBF / Matern kernels, theta = log(length_scale).

    Follows sklearn ``kernels.py:1558-1587`` (RBF) and ``:1713-1786`` (Matern) as constructed at
    acquisition_function.py:36-40.  With ``eval_gradient`` also returns dK/dtheta (n x n).
    """
    ell = float(np.exp(theta))
    X = np.atleast_2d(np.asarray(X, dtype=np.float64))
    if Y is None:
        if kernel == KERNEL_RBF or (kernel == KERNEL_MATERN and nu == np.inf):
            d2 = squareform(pdist(X / ell, metric='sqeuclidean'))
            K = np.exp(-0.5 * d2)
            np.fill_diagonal(K, 1.0)
            if eval_gradient:
                return K, K * d2
            return K
        r = squareform(pdist(X / ell, metric='euclidean'))
    else:
        if eval_gradient:
            raise ValueError('Gradient can only be evaluated when Y is None.')
        Y = np.atleast_2d(np.asarray(Y, dtype=np.float64))
        if kernel == KERNEL_RBF or (kernel == KERNEL_MATERN and nu == np.inf):
            return np.exp(-0.5 * cdist(X / ell, Y / ell, metric='sqeuclidean'))
        r = cdist(X / ell, Y / ell, metric='euclidean')

    if nu == 0.5:
        K = np.exp(-r)
    elif nu == 1.5:
        t = r * math.sqrt(3)
        K = (1.0 + t) * np.exp(-t)
    elif nu == 2.5:
        t = r * math.sqrt(5)
        K = (1.0 + t + t ** 2 / 3.0) * np.exp(-t)
    else:
        t = r.copy()
        t[t == 0.0] += np.finfo(float).eps
        t = math.sqrt(2 * nu) * t
        K = (2 ** (1.0 - nu)) / gamma(nu) * t ** nu * kv(nu, t)
    if Y is None:
        np.fill_diagonal(K, 1.0)
    if not eval_gradient:
        return K

    d2 = r ** 2
    if nu == 0.5:
        G = np.zeros_like(K)
        nz = r != 0
        G[nz] = K[nz] * d2[nz] / r[nz]
    elif nu == 1.5:
        G = 3 * d2 * np.exp(-np.sqrt(3 * d2))
    elif nu == 2.5:
        t = np.sqrt(5 * d2)
        G = 5.0 / 3.0 * d2 * (t + 1) * np.exp(-t)
    else:
        # sklearn differences the kernel numerically for general nu (kernels.py:1774-1779):
        # forward difference with step 1e-10 around log(exp(theta)), base point re-evaluated
        h = 1e-10
        xk = float(np.log(ell))
        G = (kernel_matrix(X, xk + h, kernel, nu) - kernel_matrix(X, xk, kernel, nu)) / h
    return K, G


def log_marginal_likelihood(X, y, theta, kernel=KERNEL_MATERN, nu=1.5, alpha=GPR_ALPHA,
                            eval_gradient=True):
    """LML(theta) and dLML/dtheta, sklearn ``_gpr.py:583-654`` (reached from optimizer.py:21)."""
    y = np.asarray(y, dtype=np.float64)
    if eval_gradient:
        K, G = kernel_matrix(X, theta, kernel, nu, eval_gradient=True)
    else:
        K = kernel_matrix(X, theta, kernel, nu)
    K[np.diag_indices_from(K)] += alpha
    try:
        L = cholesky(K, lower=True, check_finite=False)
    except np.linalg.LinAlgError:
        return (-np.inf, 0.0) if eval_gradient else -np.inf
    a = cho_solve((L, True), y, check_finite=False)
    lml = -0.5 * float(y @ a) - float(np.log(np.diag(L)).sum()) - K.shape[0] / 2 * math.log(2 * math.pi)
    if not eval_gradient:
        return lml
    K_inv = cho_solve((L, True), np.eye(K.shape[0]), check_finite=False)
    grad = 0.5 * float(np.einsum('ij,ji->', np.outer(a, a) - K_inv, G))
    return lml, grad


def posterior_state(X, y, theta, kernel=KERNEL_MATERN, nu=1.5, alpha=GPR_ALPHA):
    """L = chol(K + alpha I) and alpha_ = K^-1 y at a fixed theta, sklearn ``_gpr.py:349-367``."""
    K = kernel_matrix(X, theta, kernel, nu)
    K[np.diag_indices_from(K)] += alpha
    L = cholesky(K, lower=True, check_finite=False)
    a = cho_solve((L, True), np.asarray(y, dtype=np.float64), check_finite=False)
    return L, a


def predict(Xc, X, L, a, theta, kernel=KERNEL_MATERN, nu=1.5):
    """Posterior mean and std at the candidates, sklearn ``_gpr.py:446-500``
    (reached from acquisition_function.py:49,53,58).  Variance by triangular solve, negatives
    clamped to zero, exactly as the installed sklearn does."""
    Ks = kernel_matrix(Xc, theta, kernel, nu, Y=X)
    mean = Ks @ a
    V = solve_triangular(L, Ks.T, lower=True, check_finite=False)
    var = np.ones(Ks.shape[0]) - np.einsum('ij,ji->i', V.T, V)
    var[var < 0] = 0.0
    return mean, np.sqrt(var)


def norm_cdf(z):
    """scipy.stats.norm.cdf == scipy.special.ndtr (Cephes): erf branch for |z|/sqrt2 < 1,
    erfc branch otherwise."""
    z = np.asarray(z, dtype=np.float64)
    x = z * math.sqrt(0.5)
    ax = np.abs(x)
    with np.errstate(invalid='ignore'):
        small = 0.5 + 0.5 * erf(x)
        big = 0.5 * erfc(ax)
        big = np.where(x > 0, 1.0 - big, big)
        out = np.where(ax < 1.0, small, big)
    return np.where(np.isnan(z), np.nan, out)


def norm_pdf(z):
    z = np.asarray(z, dtype=np.float64)
    return np.exp(-z ** 2 / 2.0) / math.sqrt(2 * math.pi)


def acquisition(mean, std, acq, y_max, kappa=0.0, eps=0.0):
    """UCB / EI / POI exactly as acquisition_function.py:48-61 (no guard for std == 0)."""
    if acq == ACQ_UCB:
        return mean + kappa * std
    with np.errstate(divide='ignore', invalid='ignore'):
        z = (mean - y_max - eps) / std
        if acq == ACQ_EI:
            return (mean - y_max - eps) * norm_cdf(z) + std * norm_pdf(z)
        return norm_cdf(z)


def argmax_first(ys):
    """numpy ``argmax`` semantics used at acquisition_function.py:91: first maximum, NaN wins."""
    return int(np.argmax(ys))


def sweep(Xc, X, L, a, theta, kernel, nu, acq, y_max, kappa=0.0, eps=0.0, chunk=65536):
    """Warm-up sweep of ``max_compute`` (acquisition_function.py:88-92) over a candidate matrix,
    chunked so K* never exceeds ``chunk`` x n.  Returns (best_index, best_value, ys)."""
    out = np.empty(Xc.shape[0])
    for s in range(0, Xc.shape[0], chunk):
        m, sd = predict(Xc[s:s + chunk], X, L, a, theta, kernel, nu)
        out[s:s + chunk] = acquisition(m, sd, acq, y_max, kappa, eps)
    if out.size == 0:
        return -1, np.nan, out
    idx = argmax_first(out)
    return idx, out[idx], out


def acquisition_value_and_grad(x, X, L, a, theta, kernel, nu, acq, y_max, kappa=0.0, eps=0.0):
    """Acquisition at one point with its analytic gradient (used to check the device gradient;
    the reference itself differences the value numerically, acquisition_function.py:99-102)."""
    ell = math.exp(theta)
    x = np.asarray(x, dtype=np.float64).ravel()
    diff = (x[None, :] - X) / ell                      # n x d
    r2 = np.einsum('ij,ij->i', diff, diff)
    r = np.sqrt(r2)
    if kernel == KERNEL_RBF or nu == np.inf:
        k = np.exp(-0.5 * r2)
        w = -k                                          # dk/d(r2/2)
    elif nu == 0.5:
        k = np.exp(-r)
        with np.errstate(divide='ignore', invalid='ignore'):
            w = np.where(r > 0, -k / r, 0.0)
    elif nu == 1.5:
        t = math.sqrt(3) * r
        k = (1 + t) * np.exp(-t)
        w = -3.0 * np.exp(-t)
    elif nu == 2.5:
        t = math.sqrt(5) * r
        k = (1 + t + t * t / 3.0) * np.exp(-t)
        w = -(5.0 / 3.0) * (1 + t) * np.exp(-t)
    else:
        raise NotImplementedError('analytic gradient only for closed-form kernels')
    dk = (w[:, None] * diff) / ell                      # n x d : dk_j/dx
    mean = float(k @ a)
    dmean = dk.T @ a
    v = solve_triangular(L, k, lower=True, check_finite=False)
    var = 1.0 - float(v @ v)
    u = solve_triangular(L, v, lower=True, trans='T', check_finite=False)   # K^-1 k
    dvar = -2.0 * (dk.T @ u)
    if var <= 0:
        var = 0.0
    std = math.sqrt(var)
    with np.errstate(divide='ignore', invalid='ignore'):
        dstd = dvar / (2.0 * std) if std > 0 else np.zeros_like(dvar)
    if acq == ACQ_UCB:
        return mean + kappa * std, dmean + kappa * dstd
    imp = mean - y_max - eps
    with np.errstate(divide='ignore', invalid='ignore'):
        z = imp / std
    cdf = float(norm_cdf(z))
    pdf = float(norm_pdf(z))
    if acq == ACQ_EI:
        return imp * cdf + std * pdf, dmean * cdf + dstd * pdf
    dz = (dmean - z * dstd) / std if std > 0 else np.zeros_like(dmean)
    return cdf, pdf * dz
