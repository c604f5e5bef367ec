"""plx_gp_append (bordered Cholesky, one observation at fixed theta) against a full rebuild on the
device and against the CPU oracle's posterior state.

Tolerances: the appended state and the rebuilt state are two orderings of the same float64
arithmetic; both carry an error of order eps * cond(K) relative to the largest entry, so
  * L and L^-1 rows: 100 eps cond(K) relative to the matrix's largest magnitude, floor 1e-11;
  * alpha, posterior mean: max(1e-9, 100 eps cond(K)) relative, as in test_gpu_parity.py;
  * posterior variance: 1e-9 absolute (k(x,x) = 1); arg-max index exact.
"""
import numpy as np
import pytest

from oracle import gp_numpy as G
from oracle.reference_path import synthetic_problem

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps


@pytest.fixture(scope='module')
def handles():
    from polyaxon_b200 import _lib
    a, b = _lib.Handle(0), _lib.Handle(0)
    yield a, b
    a.close()
    b.close()


def _compare_states(ha, hb, cond):
    La, aa, Za = ha.get_state(want_L=True, want_Linv=True)
    Lb, ab, Zb = hb.get_state(want_L=True, want_Linv=True)
    tol = max(1e-11, 100 * EPS * cond)
    assert np.max(np.abs(La - Lb)) <= tol * np.max(np.abs(Lb))
    assert np.max(np.abs(Za - Zb)) <= tol * np.max(np.abs(Zb))
    assert np.max(np.abs(aa - ab)) <= max(1e-9, 100 * EPS * cond) * np.max(np.abs(ab))
    return Lb, ab


@pytest.mark.parametrize('kernel,nu,theta', [(G.KERNEL_MATERN, 2.5, -0.3), (G.KERNEL_RBF, 0.0, 0.2),
                                             (G.KERNEL_MATERN, 1.9, 0.0)])
def test_append_equals_rebuild(handles, kernel, nu, theta):
    ha, hb = handles
    n0, n1, d = 120, 135, 5           # crosses the 128 padding step on the way
    X, y, _ = synthetic_problem(n1, d)
    ha.set_data(X[:n0], y[:n0], kernel, nu)
    ha.set_theta(theta)
    for i in range(n0, n1):
        ha.append(X[i], y[i])
        assert ha.n == i + 1
    hb.set_data(X, y, kernel, nu)
    hb.set_theta(theta)
    K = G.kernel_matrix(X, theta, kernel, nu) + G.GPR_ALPHA * np.eye(n1)
    cond = np.linalg.cond(K)
    _, alpha = _compare_states(ha, hb, cond)
    Lo, ao = G.posterior_state(X, y, theta, kernel, nu)
    assert np.max(np.abs(alpha - ao)) <= max(1e-9, 100 * EPS * cond) * np.max(np.abs(ao))

    rs = np.random.RandomState(5)
    Xc = rs.uniform(0, 1, size=(4096, d))
    va, ia, xa, ma, sa, qa = ha.sweep(Xc, G.ACQ_EI, float(y.max()), eps=0.0, dump=True)
    vb, ib, xb, mb, sb, qb = hb.sweep(Xc, G.ACQ_EI, float(y.max()), eps=0.0, dump=True)
    assert ia == ib
    scale = max(np.max(np.abs(mb)), np.max(np.abs(y)))
    assert np.max(np.abs(ma - mb)) <= max(1e-9, 100 * EPS * cond) * scale
    assert np.max(np.abs(sa ** 2 - sb ** 2)) <= 1e-9


def test_append_many_rows_large_state(handles):
    """n = 1000 -> 1030: several 128-blocks in the triangular mat-vecs, one padding step (1024)."""
    ha, hb = handles
    n0, n1, d = 1000, 1030, 16
    X, y, _ = synthetic_problem(n1, d)
    theta = 0.4
    ha.set_data(X[:n0], y[:n0], G.KERNEL_MATERN, 1.5)
    ha.set_theta(theta)
    for i in range(n0, n1):
        ha.append(X[i], y[i])
    hb.set_data(X, y, G.KERNEL_MATERN, 1.5)
    hb.set_theta(theta)
    K = G.kernel_matrix(X, theta, G.KERNEL_MATERN, 1.5) + G.GPR_ALPHA * np.eye(n1)
    _compare_states(ha, hb, np.linalg.cond(K))
    # the fit workspace still works on the grown data set
    la, ga = ha.lml([theta, 0.1])
    lb, gb = hb.lml([theta, 0.1])
    assert np.allclose(la, lb, rtol=1e-12) and np.allclose(ga, gb, rtol=1e-9)


def test_append_needs_state_and_checks_shape(handles):
    from polyaxon_b200 import _lib
    h = _lib.Handle(0)
    try:
        X, y, _ = synthetic_problem(20, 3)
        h.set_data(X, y, G.KERNEL_RBF, 0.0)
        with pytest.raises(_lib.PlxError):
            h.append(X[0], 0.0)                      # no theta yet
        h.set_theta(0.0)
        with pytest.raises(ValueError):
            h.append(np.zeros(4), 0.0)
        n = h.n
        # a duplicate of an existing point with no jitter: the pivot is zero up to rounding; the call
        # either refuses (numpy LinAlgError, state untouched) or extends the state -- never corrupts n
        h.set_data(X, y, G.KERNEL_RBF, 0.0, jitter=0.0)
        h.set_theta(0.0)
        try:
            h.append(X[3], y[3])
            assert h.n == n + 1
        except np.linalg.LinAlgError:
            assert h.n == n
            _, a, _ = h.get_state(want_L=False)
            assert a.shape == (n,) and np.all(np.isfinite(a))
    finally:
        h.close()
