"""GPU parity: the CUDA path, called through the C ABI, against the goldens the real reference
produced (tests/golden/*.npz) and against the CPU oracle on seeded inputs.

Tolerances (float64 throughout; BASELINE.json: mean/std within 1e-9 relative, arg-max index exact):
  * L diagonal: 1e-11 relative.
  * alpha: |d_alpha| <= max(1e-9, 100 eps cond(K)) * max|alpha| -- alpha = K^-1 y inherits cond(K)
    from either solver (LAPACK in the oracle, blocked DMMA factorisation here).
  * mean: |d_mean| <= max(1e-9, 100 eps cond(K)) * max(|mean|, |y|_max).
  * variance: |d_var| <= 1e-9 * k(x,x) (k(x,x) = 1), i.e. absolute on the variance, and 1e-9 relative
    on std wherever var >= 1e-6.  std = sqrt(1 - |L^-1 k|^2) cancels catastrophically near
    observations, so a purely relative bound on std is not meaningful there.
  * acquisition: 1e-9 relative to max(1, |acq|_max), on candidates with var >= 1e-6.
  * arg-max index: exact.
"""
import warnings

import numpy as np
import pytest

from conftest import CLOSED_FORM, load_golden
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


def _setup(handle, g):
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    handle.set_theta(float(g['theta']))


def _acq_args(g):
    return dict(acq=int(g['acq']), y_max=float(g['y_max']), kappa=float(np.nan_to_num(g['kappa'])),
                eps=float(np.nan_to_num(g['eps'])))


def _x_tries(g):
    rs = np.random.RandomState(int(g['seed']))
    for _ in range(int(g['n_restarts'])):
        rs.uniform(*G.THETA_BOUNDS)
    b = g['bounds']
    return rs.uniform(b[:, 0], b[:, 1], size=(int(g['n_warmup']), b.shape[0]))


@pytest.mark.parametrize('name', CLOSED_FORM)
def test_posterior_state_matches_reference(handle, name):
    g = load_golden(name)
    _setup(handle, g)
    L, alpha, Linv = handle.get_state(want_L=True, want_Linv=True)
    cond = float(g['cond'])
    assert np.allclose(np.diag(L), g['L_diag'], rtol=1e-11, atol=0)
    tol = max(1e-9, 100 * EPS * cond)
    assert np.abs(alpha - g['alpha']).max() <= tol * np.abs(g['alpha']).max()
    if g['L'].size:
        assert np.allclose(L, g['L'], rtol=1e-10, atol=1e-12)
    n = L.shape[0]
    resid = np.abs(Linv @ L - np.eye(n)).max()
    assert resid <= 100 * EPS * max(1.0, np.sqrt(cond)) * n


@pytest.mark.parametrize('impl', [7, 6, 5, 4, 2, 1, 0])   # default rule (small sets: a thread per candidate), int8-sliced paired / cooperative / fused / two-kernel, FP64 DMMA + TMA ring, plain loads
@pytest.mark.parametrize('name', CLOSED_FORM)
def test_mean_std_acquisition_match_reference(handle, name, impl):
    g = load_golden(name)
    _setup(handle, g)
    handle.set_option('sweep_impl', impl)
    try:
        _, _, _, mean, std, acq = handle.sweep(g['x_tries_head'], dump=True, **_acq_args(g))
    finally:
        handle.set_option('sweep_impl', 3)                # back to the default (chosen per GP state)
    cond = float(g['cond'])
    mtol = max(1e-9, 100 * EPS * cond) * max(np.abs(g['mean']).max(), np.abs(g['y']).max())
    assert np.abs(mean - g['mean']).max() <= mtol
    var, var_ref = std ** 2, g['std'] ** 2
    vtol = max(1e-9, 100 * EPS * cond)
    assert np.abs(var - var_ref).max() <= vtol
    big = var_ref >= 1e-6
    if cond < 1e8:
        assert np.allclose(std[big], g['std'][big], rtol=1e-9, atol=0)
        ref = g['ys_head']
        ok = big & np.isfinite(ref)
        assert np.allclose(acq[ok], ref[ok], rtol=0, atol=1e-9 * max(1.0, np.abs(ref[ok]).max()))


@pytest.mark.parametrize('impl', [7, 6, 5, 4, 2, 1, 0])   # default rule (small sets: a thread per candidate), int8-sliced paired / cooperative / fused / two-kernel, FP64 DMMA + TMA ring, plain loads
@pytest.mark.parametrize('name', CLOSED_FORM)
def test_sweep_argmax_is_bit_exact(handle, name, impl, record_property):
    g = load_golden(name)
    _setup(handle, g)
    xt = _x_tries(g)
    handle.set_option('sweep_impl', impl)
    try:
        val, idx, x = handle.sweep(xt, **_acq_args(g))
    finally:
        handle.set_option('sweep_impl', 3)                # back to the default (chosen per GP state)
    if float(g['cond']) < 1e8:
        assert idx == int(g['argmax'])
        assert np.array_equal(x, g['x_warm'])
        assert val == pytest.approx(float(g['warm_acq']), rel=1e-9)
    else:
        # ill-conditioned reference fixture (ref14: duplicated rows + 1e-10 jitter, cond 2e10): two FP64
        # factorisations of that matrix differ by far more than the gaps between candidates, so only the winning
        # value is held to the reference; whether the index happened to agree is recorded, not hidden
        # ("known deviations")
        matched = idx == int(g['argmax'])
        record_property('argmax_index_matched_reference', matched)
        print('%s impl %d: cond %.1e, arg-max index %s the reference (%d vs %d)' % (
            name, impl, float(g['cond']), 'equals' if matched else 'differs from', idx, int(g['argmax'])))
        assert val == pytest.approx(float(g['warm_acq']), rel=1e-6)


@pytest.mark.parametrize('name', CLOSED_FORM)
def test_lml_and_gradient_batched(handle, name):
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    lml, grad = handle.lml(g['fixed_thetas'])
    for i, th in enumerate(g['fixed_thetas']):
        K = G.kernel_matrix(g['X'], float(th), int(g['kernel']), float(g['nu']))
        cond = np.linalg.cond(K + 1e-10 * np.eye(K.shape[0]))
        tol = max(1e-10, 100 * EPS * cond)
        assert lml[i] == pytest.approx(float(g['lml_fixed'][i]), rel=tol, abs=1e-9)
        assert grad[i] == pytest.approx(float(g['grad_fixed'][i]), rel=max(1e-8, tol),
                                        abs=max(1e-7, tol * abs(float(g['lml_fixed'][i]))))


def test_lml_batches_split_into_passes_give_identical_values(handle):
    """A batch that does not fit in device memory is evaluated in several passes; forced here with the
    `lml_batch_limit` option.  Every theta is computed independently, so the results are bit-identical."""
    g = load_golden('c3_d32_n512_m15_poi')
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    thetas = np.linspace(-1.0, 2.0, 7)
    whole = handle.lml(thetas)
    try:
        handle.set_option('lml_batch_limit', 3)
        parts = handle.lml(thetas)
        fit_parts = handle.fit(0.0, [-1.0, 1.0, 2.0, 0.5])
    finally:
        handle.set_option('lml_batch_limit', 0)
    assert np.array_equal(whole[0], parts[0]) and np.array_equal(whole[1], parts[1])
    fit_whole = handle.fit(0.0, [-1.0, 1.0, 2.0, 0.5])
    assert fit_parts[0] == fit_whole[0] and fit_parts[1] == fit_whole[1]
    assert np.array_equal(fit_parts[2], fit_whole[2])


@pytest.mark.parametrize('name', CLOSED_FORM)
def test_fit_reaches_the_reference_optimum(handle, name):
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    theta, lml, per_run = handle.fit(np.log(float(g['length_scale'])), g['theta_inits'])
    assert theta == pytest.approx(float(g['theta']), abs=1e-6)
    assert lml == pytest.approx(float(g['lml']), rel=1e-9, abs=1e-8)
    ref = g['per_restart']
    assert np.allclose(per_run[:, 0], ref[:, 0])
    assert np.allclose(per_run[:, 1], ref[:, 1], atol=1e-5)
    assert np.allclose(per_run[:, 2], ref[:, 2], rtol=1e-8, atol=1e-7)


@pytest.mark.parametrize('name', [n for n in CLOSED_FORM if load_golden(n)['X'].shape[0] <= 128])
def test_one_launch_fit_equals_the_batched_launch_sequence(handle, name):
    """Small observation sets (by default up to 64, here forced up to the kernel's capacity of 128) are fitted with
    every optimiser run from start to convergence inside one kernel (fit_small.cuh: factorisation, gradient and the
    L-BFGS-B state machine on the device); larger sets use the batched launch sequence driven from the host.  Same optimiser code, the factorisation differs in summation order only."""
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    starts = np.concatenate([g['theta_inits'], [-3.0, 0.4, 2.5, 9.0, -9.0]])
    out = {}
    try:
        for flag in (2, 0):
            handle.set_option('fit_small', flag)
            out[flag] = handle.fit(np.log(float(g['length_scale'])), starts)
    finally:
        handle.set_option('fit_small', 1)
    well = float(g['cond']) < 1e8
    assert out[2][0] == pytest.approx(out[0][0], abs=1e-6 if well else 1e-3)
    assert out[2][1] == pytest.approx(out[0][1], rel=1e-9 if well else 1e-6, abs=1e-8)
    a, b = out[2][2], out[0][2]
    assert np.array_equal(a[:, 0], b[:, 0])
    finite = np.isfinite(b[:, 2])
    assert np.array_equal(np.isfinite(a[:, 2]), finite)
    assert np.allclose(a[finite, 2], b[finite, 2], rtol=1e-8 if well else 1e-5, atol=1e-7)
    if well:
        assert np.allclose(a[:, 1], b[:, 1], atol=1e-5)
        assert np.array_equal(a[:, 3], b[:, 3])                      # the same number of evaluations per run


@pytest.mark.parametrize('n_cand', [0, 1, 127, 128, 129, 1000])
def test_ragged_candidate_counts(handle, n_cand):
    g = load_golden('c5_d16_n128_m25_ei')
    _setup(handle, g)
    xt = _x_tries(g)[:n_cand]
    val, idx, x = handle.sweep(xt, **_acq_args(g))
    if n_cand == 0:
        assert idx == -1 and np.isnan(val)
        return
    L, a = G.posterior_state(g['X'], g['y'], float(g['theta']), int(g['kernel']), float(g['nu']))
    ridx, rval, _ = G.sweep(xt, g['X'], L, a, float(g['theta']), int(g['kernel']), float(g['nu']),
                            int(g['acq']), float(g['y_max']), 0.0, float(g['eps']))
    assert idx == ridx
    assert val == pytest.approx(rval, rel=1e-9)
    assert np.array_equal(x, xt[ridx])


@pytest.mark.parametrize('n,d', [(1, 1), (2, 3), (17, 5), (130, 40), (300, 70)])
def test_odd_shapes_against_oracle(handle, n, d):
    X, y, bounds = synthetic_problem(n, d, seed=77 + n)
    theta = 0.3
    handle.set_data(X, y, G.KERNEL_MATERN, 2.5)
    handle.set_theta(theta)
    L, a = G.posterior_state(X, y, theta, G.KERNEL_MATERN, 2.5)
    xc = np.random.RandomState(n).uniform(0, 1, (700, d))
    val, idx, _, mean, std, acq = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=2.0, dump=True)
    m, s = G.predict(xc, X, L, a, theta, G.KERNEL_MATERN, 2.5)
    assert np.abs(mean - m).max() <= 1e-9 * max(1.0, np.abs(m).max())
    assert np.abs(std ** 2 - s ** 2).max() <= 1e-9
    ridx = int(np.argmax(m + 2.0 * s))
    assert idx == ridx


def test_nan_wins_and_first_index_on_ties(handle):
    # EI/POI with eps=0 at an observed point that is also the incumbent: std -> 0 and mean - y_max -> 0
    # gives z = 0/0 = NaN in the reference; numpy.argmax returns the first NaN.
    X = np.array([[0.1], [0.5], [0.9]])
    y = np.array([0.0, 1.0, 0.2])
    handle.set_data(X, y, G.KERNEL_RBF, 1.5, jitter=0.0)
    handle.set_theta(np.log(0.05))          # essentially uncorrelated: K ~ I, alpha ~ y
    xc = np.array([[0.3], [0.5], [0.7], [0.5]])
    L, a = G.posterior_state(X, y, np.log(0.05), G.KERNEL_RBF, 1.5, alpha=0.0)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        ridx, rval, racq = G.sweep(xc, X, L, a, np.log(0.05), G.KERNEL_RBF, 1.5, G.ACQ_POI, 1.0)
    val, idx, _, _, _, acq = handle.sweep(xc, G.ACQ_POI, 1.0, dump=True)
    assert np.array_equal(np.isnan(acq), np.isnan(racq))
    assert idx == ridx
    # exact ties: duplicated candidates -> the lower index wins
    xc2 = np.tile(np.array([[0.42]]), (300, 1))
    val, idx, _ = handle.sweep(xc2, G.ACQ_UCB, 1.0, kappa=1.0)
    assert idx == 0


def test_not_positive_definite_raises_linalgerror(handle):
    X = np.array([[0.2, 0.3], [0.2, 0.3], [0.7, 0.1]])       # duplicated row, no jitter
    handle.set_data(X, np.array([1.0, 1.0, 0.0]), G.KERNEL_RBF, 1.5, jitter=0.0)
    with pytest.raises(np.linalg.LinAlgError):
        handle.set_theta(0.0)
    lml, grad = handle.lml([0.0])
    assert lml[0] == -np.inf and grad[0] == 0.0              # sklearn _gpr.py:592-593


def test_unsupported_kernel_arguments_fail_loudly(handle):
    from polyaxon_b200._lib import PlxError
    with pytest.raises(PlxError):
        handle.set_data(np.zeros((2, 1)), np.zeros(2), G.KERNEL_MATERN, -1.0)
    with pytest.raises(PlxError):
        handle.set_data(np.zeros((2, 1)), np.zeros(2), 7, 1.5)


def test_large_problem_against_oracle_and_shard_merge(handle):
    """n_obs = 2048, d = 32 (the headline shape): posterior on a block against the oracle, and the
    size-independent property that sharding the candidates and merging the per-shard bests
    (value, global index) reproduces the single sweep."""
    X, y, _ = synthetic_problem(2048, 32)
    theta = 1.2
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    handle.set_theta(theta)
    rs = np.random.RandomState(5)
    xc = rs.uniform(0, 1, (40000, 32))
    y_max = float(y.max())
    val, idx, xb, mean, std, acq = handle.sweep(xc, G.ACQ_POI, y_max, eps=0.0, dump=True)
    L, a = G.posterior_state(X, y, theta, G.KERNEL_MATERN, 1.5)
    m, s = G.predict(xc[:2048], X, L, a, theta, G.KERNEL_MATERN, 1.5)
    assert np.abs(mean[:2048] - m).max() <= 1e-9 * max(1.0, np.abs(m).max())
    assert np.allclose(std[:2048], s, rtol=1e-9, atol=0)
    # arg-max against the oracle's own sweep of all 40 000 candidates (not against the device's dump)
    ridx, rval, racq = G.sweep(xc, X, L, a, theta, G.KERNEL_MATERN, 1.5, G.ACQ_POI, y_max, 0.0, 0.0, chunk=8192)
    assert idx == ridx
    assert val == pytest.approx(rval, rel=1e-9)
    assert val == acq[idx]
    assert np.array_equal(xb, xc[idx])
    best = (-np.inf, -1)
    for lo in range(0, 40000, 13337):
        v, i, _ = handle.sweep(xc[lo:lo + 13337], G.ACQ_POI, y_max, eps=0.0, idx_base=lo)
        if v > best[0] or (v == best[0] and i < best[1]):
            best = (v, i)
    assert best == (val, idx)


@pytest.mark.parametrize('kernel,nu', [(0, 1.5), (1, 0.5), (1, 1.5), (1, 2.5)])
def test_inner_loop_kernel_values_to_a_few_ulp(handle, kernel, nu):
    """The sweep's branch-free sqrt / exp against numpy on squared distances over the whole range the
    path can produce (0, denormal-small, typical, far beyond underflow)."""
    rs = np.random.RandomState(0)
    r2 = np.concatenate([[0.0, 1e-300, 1e-200, 1e-30, 1e-8, 0.5, 1.0, 2.0, 700.0, 5e5, 1e12],
                         10.0 ** rs.uniform(-12, 6, 20000), rs.uniform(0, 60, 20000)])
    got = handle.kernel_values(kernel, nu, r2)
    r = np.sqrt(r2)
    if kernel == 0:
        want = np.exp(-0.5 * r2)
    elif nu == 0.5:
        want = np.exp(-r)
    elif nu == 1.5:
        t = r * np.sqrt(3)
        want = (1 + t) * np.exp(-t)
    else:
        t = r * np.sqrt(5)
        want = (1 + t + t * t / 3.0) * np.exp(-t)
    tiny = want < 1e-290                       # flushed to zero below the normal range
    assert np.all(got[tiny] <= 1e-290)
    ok = ~tiny
    ulp = np.abs(got[ok] - want[ok]) / np.spacing(want[ok])
    assert ulp.max() <= 4.0, ulp.max()
    assert got[0] == 1.0
    # the range checks read the argument's high word: -0.0 is a zero distance, NaN stays NaN (a NaN candidate
    # coordinate must reach the acquisition value, where numpy's arg-max picks it)
    special = handle.kernel_values(kernel, nu, np.array([-0.0, np.nan]))
    assert special[0] == 1.0 and np.isnan(special[1])


def test_reserved_workspaces_change_nothing_but_the_first_call():
    """plx_bo_reserve allocates the workspaces of a shape ahead of time (no cudaMalloc inside the first call of a
    group); results are those of an unreserved handle, and it is refused once data has been set."""
    from polyaxon_b200 import _lib
    g = load_golden('c2_d8_n256_rbf_ei')
    out = []
    for reserve in (True, False):
        h = _lib.Handle(0)
        try:
            if reserve:
                h.reserve(256, 8, runs=3, seeds=4, candidates=8192)
            h.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
            theta, lml, _ = h.fit(0.0, [1.0, -1.0])
            val, idx, x = h.sweep(_x_tries(g), **_acq_args(g))
            xs, vals, ok = h.refine(g['x_seeds'], g['bounds'], **_acq_args(g))
            out.append((theta, lml, val, idx, vals.tolist()))
            with pytest.raises(_lib.PlxError):
                h.reserve(64, 2)
        finally:
            h.close()
    assert out[0] == out[1]


def test_negative_variances_are_counted_and_warned_about(handle):
    """sklearn sets negative predicted variances to zero and warns (_gpr.py:485-491).  The kernels clamp and count;
    the drop-in predict raises the same warning.  Candidates ON observations with a tiny jitter make 1 - |L^-1 k|^2
    round below zero for some of them."""
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import DeviceGaussianProcess
    X, y, _ = synthetic_problem(300, 3, seed=11)
    gp = DeviceGaussianProcess('rbf', 1.0, 1.5, 0, np.random.RandomState(1), handle=handle)
    gp.alpha = 1e-14
    gp.n_restarts_optimizer = 0
    handle.set_data(X, y, G.KERNEL_RBF, 1.5, jitter=1e-14)
    handle.set_theta(-0.5)
    for impl in (3, 1):
        handle.set_option('sweep_impl', impl)
        _, _, _, mean, std, _ = handle.sweep(np.repeat(X, 4, axis=0), G.ACQ_UCB, 0.0, kappa=0.0, dump=True)
        clamped = handle.last_sweep_clamped()
        assert clamped == int(np.sum(std == 0.0))
    handle.set_option('sweep_impl', 3)
    gp.X_train_, gp.y_train_, gp.theta_, gp.fitted = X, y, -0.5, True
    gp._stamp()
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter('always')
        gp.predict(np.repeat(X, 4, axis=0), return_std=True)
    assert (clamped > 0) == any('Predicted variances smaller than 0' in str(w.message) for w in caught)
    _, _, _ = handle.sweep(np.random.RandomState(0).uniform(2, 3, (500, 3)), G.ACQ_UCB, 0.0, kappa=1.0)
    assert handle.last_sweep_clamped() == 0                      # far from every observation: nothing to clamp
