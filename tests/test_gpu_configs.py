"""GPU parity at the shapes of BASELINE.json configs C4 (batched fit sweep, many restarts) and C5
(independent groups, replicas only), plus size-independent properties at full n_obs."""
import numpy as np
import pytest

from oracle import gp_numpy as G
from oracle.reference_path import synthetic_problem

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps


@pytest.fixture(scope='module')
def handle():
    from polyaxon_b200 import _lib
    h = _lib.Handle(0)
    yield h
    h.close()


@pytest.mark.parametrize('n', [64, 200, 1024])
def test_c4_batched_lml_over_33_thetas(handle, n):
    """C4: n_obs sweep, d = 32, 33 length scales evaluated at once (Cholesky + LML gradient)."""
    X, y, _ = synthetic_problem(n, 32)
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    thetas = np.linspace(-1.0, 3.0, 33)
    lml, grad = handle.lml(thetas)
    for i in (0, 7, 16, 25, 32):
        rl, rg = G.log_marginal_likelihood(X, y, float(thetas[i]), G.KERNEL_MATERN, 1.5)
        K = G.kernel_matrix(X, float(thetas[i]), G.KERNEL_MATERN, 1.5)
        tol = max(1e-10, 100 * EPS * np.linalg.cond(K + 1e-10 * np.eye(n)))
        assert lml[i] == pytest.approx(rl, rel=tol, abs=1e-8)
        assert grad[i] == pytest.approx(rg, rel=max(1e-8, tol), abs=max(1e-7, tol * abs(rl)))


def test_c4_largest_shape_against_the_oracle(handle):
    """n_obs = 4096, the top of C4's range: LML and its gradient against one oracle evaluation (a few seconds of
    LAPACK), the posterior on a candidate block against the oracle, and the size-independent properties: the
    gradient against a central difference of the device LML, and batch invariance."""
    X, y, _ = synthetic_problem(4096, 32)
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    th = 1.1
    h = 1e-4
    lml, grad = handle.lml([th - h, th, th + h])
    fd = (lml[2] - lml[0]) / (2 * h)
    assert grad[1] == pytest.approx(fd, rel=1e-5)
    alone, galone = handle.lml([th])
    assert alone[0] == lml[1] and galone[0] == grad[1]            # bit-identical whatever the batch
    rl, rg = G.log_marginal_likelihood(X, y, th, G.KERNEL_MATERN, 1.5)
    K = G.kernel_matrix(X, th, G.KERNEL_MATERN, 1.5)
    tol = max(1e-10, 100 * EPS * np.linalg.cond(K + 1e-10 * np.eye(4096)))
    assert lml[1] == pytest.approx(rl, rel=tol, abs=1e-8)
    assert grad[1] == pytest.approx(rg, rel=max(1e-8, tol), abs=max(1e-7, tol * abs(rl)))
    handle.set_theta(th)
    L, a = G.posterior_state(X, y, th, G.KERNEL_MATERN, 1.5)
    xc = np.random.RandomState(4).uniform(0, 1, (1500, 32))
    val, idx, _, mean, std, acq = handle.sweep(xc, G.ACQ_POI, float(y.max()), eps=0.0, dump=True)
    m, s = G.predict(xc, X, L, a, th, G.KERNEL_MATERN, 1.5)
    assert np.abs(mean - m).max() <= max(1e-9, tol) * max(1.0, np.abs(m).max())
    assert np.abs(std ** 2 - s ** 2).max() <= max(1e-9, tol)
    assert idx == int(np.argmax(G.acquisition(m, s, G.ACQ_POI, float(y.max()), 0.0, 0.0)))


def test_c5_independent_groups_on_one_device(handle):
    """C5: groups are independent GPs (d = 16, n_obs = 512); a rank works through its share one
    after the other (replicas only, no collective).  Each group's result must not depend on what
    ran before it on the same handle."""
    from polyaxon_b200 import dist as pd
    groups = list(range(6))
    lo, hi = pd.shard_range(len(groups), 1, 2)                    # what rank 1 of 2 would own
    assert (lo, hi) == (3, 6)
    results = {}
    for order in (groups, groups[::-1]):
        for gid in order:
            X, y, _ = synthetic_problem(512, 16, seed=1234 + gid)
            handle.set_data(X, y, G.KERNEL_MATERN, 2.5)
            theta, lml, _ = handle.fit(0.0)
            xc = np.random.RandomState(gid).uniform(0, 1, (3000, 16))
            val, idx, _ = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=2.576)
            if gid in results:
                assert results[gid] == (theta, lml, val, idx)
            results[gid] = (theta, lml, val, idx)
    X, y, _ = synthetic_problem(512, 16, seed=1234 + 2)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        from oracle import reference_path as R
        gp = R.make_gp(G.KERNEL_MATERN, 1.0, 2.5, 0, np.random.RandomState(1))
        gp.fit(X, y)
    assert results[2][0] == pytest.approx(float(gp.kernel_.theta[0]), abs=1e-6)
    assert results[2][1] == pytest.approx(float(gp.log_marginal_likelihood_value_), rel=1e-9)


def test_full_size_interpolation_property(handle):
    """n_obs = 2048, d = 32: at the observations themselves the posterior mean reproduces y up to
    the jitter and the variance collapses (size-independent property, no oracle needed)."""
    X, y, _ = synthetic_problem(2048, 32)
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    handle.set_theta(1.2)
    _, _, _, mean, std, _ = handle.sweep(X[:512], G.ACQ_UCB, 0.0, kappa=0.0, dump=True)
    _, alpha, _ = handle.get_state(want_L=False)
    assert np.abs(mean - (y[:512] - 1e-10 * alpha[:512])).max() <= 1e-9 * max(1.0, np.abs(y).max())
    assert std.max() <= 1e-4                                       # sqrt(jitter-level variance)


def test_sweep_beyond_2048_observations(handle):
    """n_obs = 3000 (not a multiple of the 128-row blocks), d = 16, against the oracle."""
    X, y, _ = synthetic_problem(3000, 16)
    theta = 0.9
    handle.set_data(X, y, G.KERNEL_MATERN, 2.5)
    handle.set_theta(theta)
    L, a = G.posterior_state(X, y, theta, G.KERNEL_MATERN, 2.5)
    xc = np.random.RandomState(8).uniform(0, 1, (777, 16))
    val, idx, _, mean, std, acq = handle.sweep(xc, G.ACQ_EI, float(y.max()), eps=0.0, dump=True)
    m, s = G.predict(xc, X, L, a, theta, G.KERNEL_MATERN, 2.5)
    ref = G.acquisition(m, s, G.ACQ_EI, float(y.max()), 0.0, 0.0)
    assert np.abs(mean - m).max() <= 1e-9 * max(1.0, np.abs(m).max())
    assert np.abs(std ** 2 - s ** 2).max() <= 1e-9
    assert idx == int(np.argmax(ref))


def test_host_sweep_across_chunk_boundaries_matches_resident_sweep(handle):
    """The host-buffer entry point streams candidates in double-buffered chunks (a whole number of
    128-candidate waves, ~64 MiB); N = chunk + 1 exercises the ragged last chunk and the running
    arg-max across chunks."""
    X = np.array([[0.1], [0.35], [0.5], [0.8], [0.95]])
    y = np.array([0.2, 0.9, 0.1, 0.7, 0.3])
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    handle.set_theta(np.log(0.2))
    wave = 128 * 148
    chunk = wave * ((64 << 20) // (8 * 1 * wave))
    N = chunk + 1
    xc = np.random.RandomState(2).uniform(0, 1, (N, 1))
    xc[-1, 0] = 0.35                      # the best candidate sits alone in the last chunk
    val, idx, x = handle.sweep(xc, G.ACQ_UCB, 0.9, kappa=0.0)
    handle.candidates_set(xc)
    val2, idx2, x2 = handle.sweep_resident(G.ACQ_UCB, 0.9, kappa=0.0)
    assert (val, idx) == (val2, idx2) and np.array_equal(x, x2)
    L, a = G.posterior_state(X, y, np.log(0.2), G.KERNEL_MATERN, 1.5)
    m, _ = G.predict(xc[-5:], X, L, a, np.log(0.2), G.KERNEL_MATERN, 1.5)
    assert val >= m.max() - 1e-12
