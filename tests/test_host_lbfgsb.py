"""Host logic: the one-variable L-BFGS-B of plx_gp_fit against scipy (the optimiser sklearn's
GaussianProcessRegressor.fit drives, _gpr.py:656-676) on the oracle's marginal likelihood."""
import numpy as np
import pytest
import scipy.optimize as so

from conftest import golden_names, load_golden
from oracle import gp_numpy as G
from polyaxon_b200 import _lib

LO, HI = G.THETA_BOUNDS


def _scipy(fun, x0):
    res = so.minimize(lambda x: fun(float(x[0])), np.array([x0]), jac=True, method='L-BFGS-B',
                      bounds=[(LO, HI)])
    return float(res.x[0]), float(res.fun), res.nfev, res.nit, res.status


@pytest.mark.parametrize('x0', [-3.0, 0.0, 2.5, 9.0, -11.0])
def test_analytic_functions_follow_scipy_step_for_step(x0):
    def fun(x):
        return ((x - 1.3) ** 2 * (1 + 0.1 * np.sin(5 * x)),
                2 * (x - 1.3) * (1 + 0.1 * np.sin(5 * x)) + (x - 1.3) ** 2 * 0.5 * np.cos(5 * x))
    mine = _lib.lbfgsb1d_minimize(fun, x0, LO, HI)
    ref = _scipy(fun, x0)
    assert mine[0] == pytest.approx(ref[0], abs=1e-12)
    assert mine[2:4] == ref[2:4]          # same number of evaluations and iterations


def test_minimum_on_the_bound():
    fun = lambda x: (np.exp(0.3 * x), 0.3 * np.exp(0.3 * x))
    mine = _lib.lbfgsb1d_minimize(fun, 1.0, LO, HI)
    ref = _scipy(fun, 1.0)
    assert mine[0] == ref[0] == LO
    assert mine[2:4] == ref[2:4]


SMALL = [n for n in golden_names() if '_m19_' not in n and 'n2048' not in n]


@pytest.mark.parametrize('name', SMALL)
def test_marginal_likelihood_runs_land_where_sklearn_lands(name):
    g = load_golden(name)
    kernel, nu = int(g['kernel']), float(g['nu'])

    def fun(theta):
        lml, grad = G.log_marginal_likelihood(g['X'], g['y'], theta, kernel, nu)
        return -lml, -grad

    for start, theta_ref, lml_ref in g['per_restart']:
        x, f, nfev, nit, status = _lib.lbfgsb1d_minimize(fun, float(start), LO, HI)
        sx, sf, snfev, snit, _ = _scipy(fun, float(start))
        assert x == pytest.approx(sx, abs=1e-9)
        assert (nfev, nit) == (snfev, snit)
        # and both agree with what sklearn recorded inside the real reference run
        assert x == pytest.approx(float(theta_ref), abs=1e-6)
        assert -f == pytest.approx(float(lml_ref), rel=1e-9, abs=1e-9)


def test_headline_shape_run_lands_where_sklearn_landed():
    """n_obs = 2048 (config C3 at its real shape): the first optimiser run on the oracle's marginal likelihood ends at
    the optimum sklearn recorded inside the real reference run.  (An evaluation costs the oracle seconds here, so only
    this run is replayed on the CPU, and step-for-step equality with scipy is not asked for: at this size the
    objective carries 1e-10 of rounding noise and one of the 17 runs takes 15 evaluations against scipy's 14 to the
    same optimum.  All 17 runs are held to the reference on the GPU, tests/test_gpu_parity.py.)"""
    g = load_golden('c3_d32_n2048_m15_poi_full')
    kernel, nu = int(g['kernel']), float(g['nu'])

    def fun(theta):
        lml, grad = G.log_marginal_likelihood(g['X'], g['y'], theta, kernel, nu)
        return -lml, -grad

    start, theta_ref, lml_ref = g['per_restart'][0]
    x, f, nfev, nit, status = _lib.lbfgsb1d_minimize(fun, float(start), LO, HI)
    assert status == 0
    assert x == pytest.approx(float(theta_ref), abs=1e-6)
    assert -f == pytest.approx(float(lml_ref), rel=1e-9, abs=1e-9)
