"""The int8-sliced sweeps (sweep_i8*.cuh: paired cooperative = 6, cooperative = 5, fused = 4, two kernels = 2) at sizes and in corners the
goldens do not reach, against the CPU oracle and against the FP64 DMMA kernel (1).

Tolerances as in test_gpu_parity.py: posterior variance 1e-9 absolute (k(x,x) = 1) -- max(1e-9, 100 eps cond(K)) on
the ill-conditioned state --, posterior mean max(1e-9, 100 eps cond(K)) relative to its scale, arg-max index exact.
"""
import threading
import warnings

import numpy as np
import pytest

from oracle import gp_numpy as G
from oracle.reference_path import synthetic_problem

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps
INT8_IMPLS = (6, 5, 4, 2)


@pytest.fixture(scope='module')
def handle():
    from polyaxon_b200 import _lib
    h = _lib.Handle(0)
    yield h
    h.close()


def _oracle(X, y, theta, kernel, nu, xc):
    L, a = G.posterior_state(X, y, theta, kernel, nu)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return G.predict(xc, X, L, a, theta, kernel, nu)


@pytest.mark.parametrize('n,d,kernel,nu', [(1600, 7, G.KERNEL_MATERN, 1.5), (1000, 33, G.KERNEL_RBF, 0.0),
                                           (257, 3, G.KERNEL_MATERN, 2.5), (70, 2, G.KERNEL_MATERN, 0.5)])
def test_all_variants_against_oracle_at_ragged_sizes(handle, n, d, kernel, nu):
    """n not a multiple of 64 (ragged last half-block), d above 32 (two dimension chunks), fewer than four
    half-blocks (idle CTAs in a cooperative group), N not a multiple of 128."""
    X, y, _ = synthetic_problem(n, d)
    theta = 0.3
    handle.set_data(X, y, kernel, nu)
    handle.set_theta(theta)
    xc = np.random.RandomState(n).uniform(0, 1, (128 * 37 * 3 + 77, d))
    m, s = _oracle(X, y, theta, kernel, nu, xc)
    cond = np.linalg.cond(G.kernel_matrix(X, theta, kernel, nu) + G.GPR_ALPHA * np.eye(n))
    mtol = max(1e-9, 100 * EPS * cond) * max(np.abs(m).max(), np.abs(y).max())
    vtol = max(1e-9, 100 * EPS * cond)
    results = {}
    try:
        for impl in INT8_IMPLS + (1,):
            handle.set_option('sweep_impl', impl)
            val, idx, x, mean, std, acq = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=1.5, dump=True)
            assert np.abs(mean - m).max() <= mtol, impl
            assert np.abs(std ** 2 - s ** 2).max() <= vtol, impl
            results[impl] = (idx, std ** 2)
    finally:
        handle.set_option('sweep_impl', 3)
    for impl in INT8_IMPLS:
        assert results[impl][0] == results[1][0]
        # exact slices against the FP64 tensor pipe: both are roundings of the same sum
        assert np.abs(results[impl][1] - results[1][1]).max() <= max(1e-12, 10 * EPS * cond)


def test_ill_conditioned_state_and_candidates_on_observations(handle):
    """Long RBF length scale: cond(K) ~ 1e10, rows of L^-1 up to 1e5.  Candidates that coincide with observations
    give k = 1 exactly (leading digit 64) and a variance that cancels to ~0."""
    n, d = 300, 4
    X, y, _ = synthetic_problem(n, d)
    theta = 0.9
    handle.set_data(X, y, G.KERNEL_RBF, 0.0)
    handle.set_theta(theta)
    rs = np.random.RandomState(9)
    xc = np.vstack([X[:50], rs.uniform(0, 1, (4000, d)), X[-50:]])
    m, s = _oracle(X, y, theta, G.KERNEL_RBF, 0.0, xc)
    cond = np.linalg.cond(G.kernel_matrix(X, theta, G.KERNEL_RBF, 0.0) + G.GPR_ALPHA * np.eye(n))
    assert cond > 1e8
    vtol = max(1e-9, 100 * EPS * cond)
    ref = None
    try:
        for impl in (1,) + INT8_IMPLS:
            handle.set_option('sweep_impl', impl)
            _, _, _, mean, std, _ = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=0.0, dump=True)
            var = std ** 2
            assert np.all(np.isfinite(var)) and np.all(var >= 0.0)
            assert np.abs(var - s ** 2).max() <= vtol, impl
            assert var[:50].max() <= vtol and var[-50:].max() <= vtol       # on the observations
            if ref is None:
                ref = var
            else:
                assert np.abs(var - ref).max() <= vtol
    finally:
        handle.set_option('sweep_impl', 3)


def test_two_cooperative_sweeps_from_two_handles_at_once():
    """The cooperative kernel spins on counters shared by the CTAs of a group, which needs the whole grid resident;
    it is started with a cooperative launch so that sweeps from two handles (two worker threads of
    suggest_groups) queue up instead of starving each other."""
    from polyaxon_b200 import _lib
    n, d = 512, 6
    X, y, _ = synthetic_problem(n, d)
    xc = np.random.RandomState(4).uniform(0, 1, (128 * 148 * 2, d))
    handles = [_lib.Handle(0), _lib.Handle(0)]
    out = [None, None]

    def work(k):
        h = handles[k]
        h.set_data(X, y, G.KERNEL_MATERN, 2.5)
        h.set_theta(0.1)
        h.set_option('sweep_impl', 5)
        res = []
        for _ in range(6):
            res.append(h.sweep(xc, G.ACQ_EI, float(y.max()), eps=0.0)[:2])
        out[k] = res

    threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    try:
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=60)
        assert not any(t.is_alive() for t in threads), 'concurrent cooperative sweeps did not finish'
        assert out[0] == out[1] and len(set(out[0])) == 1
    finally:
        if not any(t.is_alive() for t in threads):
            for h in handles:
                h.close()


@pytest.mark.parametrize('impl', INT8_IMPLS)
@pytest.mark.parametrize('n', [70, 600, 2048])
def test_a_planes_through_tensor_memory_equal_shared_memory_operands(handle, impl, n):
    """The candidate tile's digit planes reach the tensor core through tensor memory (tcgen05.cp into a ring of
    8-column slots, option i8_operand = 0, the default) or straight from shared memory (1).  Both are the same exact
    integer products in the same accumulation classes, so every output must be bit-identical -- also after many
    stages, when every slot of the ring has been overwritten often."""
    d = 9
    X, y, _ = synthetic_problem(n, d, seed=5 + n)
    handle.set_data(X, y, G.KERNEL_MATERN, 2.5)
    handle.set_theta(0.2)
    xc = np.random.RandomState(n + impl).uniform(0, 1, (128 * 148 * 2 + 31, d))
    out = {}
    try:
        handle.set_option('sweep_impl', impl)
        for operand in (0, 1):
            handle.set_option('i8_operand', operand)
            out[operand] = handle.sweep(xc, G.ACQ_EI, float(y.max()), eps=0.01, dump=True)
    finally:
        handle.set_option('i8_operand', 0)
        handle.set_option('sweep_impl', 3)
    assert out[0][1] == out[1][1] and out[0][0] == out[1][0]
    for k in (3, 4, 5):
        assert np.array_equal(out[0][k], out[1][k], equal_nan=True)
    m, s = _oracle(X, y, 0.2, G.KERNEL_MATERN, 2.5, xc[:4096])
    assert np.abs(out[0][4][:4096] ** 2 - s ** 2).max() <= 1e-9


@pytest.mark.parametrize('impl', (5, 4))
@pytest.mark.parametrize('n,d,kernel,nu', [(1600, 32, G.KERNEL_MATERN, 0.5), (700, 8, G.KERNEL_MATERN, 1.5),
                                           (300, 1, G.KERNEL_RBF, 0.0), (2048, 31, G.KERNEL_MATERN, 2.5)])
def test_tensor_path_distances_equal_direct_differences(handle, impl, n, d, kernel, nu):
    """Option i8_distance: the builders form |c - x|^2 as |u|^2 + |v|^2 - 2 u.v with the cross term on the FP64
    tensor path (1; the default of the cooperative form) or by direct differences (0).  Candidates that are
    observations, or sit 1e-12 .. 1e-2 away from one, are where the expansion cancels: those pairs must be recomputed
    directly (exactly zero / a relative-accurate small distance -- nu = 0.5 takes its square root)."""
    X, y, _ = synthetic_problem(n, d, seed=11 + d)
    theta = 0.4
    handle.set_data(X, y, kernel, nu)
    handle.set_theta(theta)
    rs = np.random.RandomState(d)
    near = [X[:40]] + [X[40 * (k + 1):40 * (k + 2)] + 10.0 ** (-12 + k) * rs.standard_normal((40, d)) for k in range(6)]
    xc = np.vstack(near + [rs.uniform(0, 1, (128 * 37 + 5, d))])
    m, s = _oracle(X, y, theta, kernel, nu, xc)
    cond = np.linalg.cond(G.kernel_matrix(X, theta, kernel, nu) + G.GPR_ALPHA * np.eye(n))
    mtol = max(1e-9, 100 * EPS * cond) * max(np.abs(m).max(), np.abs(y).max())
    vtol = max(1e-9, 100 * EPS * cond)
    out = {}
    try:
        handle.set_option('sweep_impl', impl)
        for dist in (0, 1):
            handle.set_option('i8_distance', dist)
            out[dist] = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=1.0, dump=True)
            assert np.abs(out[dist][3] - m).max() <= mtol, dist
            assert np.abs(out[dist][4] ** 2 - s ** 2).max() <= vtol, dist
    finally:
        handle.set_option('i8_distance', 2)
        handle.set_option('sweep_impl', 3)
    if cond < 1e8:                                             # beyond that the runner-up may win within the tolerance
        assert out[0][1] == out[1][1]
    assert np.abs(out[0][4] ** 2 - out[1][4] ** 2).max() <= max(1e-10, EPS * cond)
    assert np.abs(out[0][3] - out[1][3]).max() <= max(1e-10, EPS * cond) * max(np.abs(m).max(), 1.0)
    assert (out[1][4][:40] ** 2).max() <= vtol                 # on the observations themselves


@pytest.mark.parametrize('n,d', [(1100, 20), (1100, 8), (1500, 32)])
def test_default_form_between_1024_and_1536_observations(handle, n, d):
    """The default (sweep_impl 3) picks the cooperative form from 1024 observations when the tensor-path builders apply
    (16 to 32 dimensions) and the fused form otherwise; either way the posterior is the oracle's."""
    X, y, _ = synthetic_problem(n, d, seed=n + d)
    theta = 0.5
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    handle.set_theta(theta)
    xc = np.random.RandomState(n).uniform(0, 1, (128 * 148 + 9, d))
    m, s = _oracle(X, y, theta, G.KERNEL_MATERN, 1.5, xc)
    cond = np.linalg.cond(G.kernel_matrix(X, theta, G.KERNEL_MATERN, 1.5) + G.GPR_ALPHA * np.eye(n))
    handle.set_option('sweep_impl', 3)
    val, idx, x, mean, std, acq = handle.sweep(xc, G.ACQ_EI, float(y.max()), eps=0.0, dump=True)
    assert np.abs(mean - m).max() <= max(1e-9, 100 * EPS * cond) * max(np.abs(m).max(), np.abs(y).max())
    assert np.abs(std ** 2 - s ** 2).max() <= max(1e-9, 100 * EPS * cond)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        want = G.acquisition(m, s, G.ACQ_EI, float(y.max()), 0.0, 0.0)
    if cond < 1e8:
        assert idx == G.argmax_first(want)


@pytest.mark.parametrize('impl', (3, 4, 1))
def test_nan_candidate_wins_the_arg_max_like_numpy(handle, impl):
    """A candidate with a NaN coordinate has a NaN acquisition value in the reference (sklearn propagates it through
    cdist and the kernel), and numpy's arg-max returns the first NaN (acquisition_function.py:91).  The builders'
    range checks work on bit patterns and the tensor-path distances go through DMMA: the NaN must survive both."""
    n, d = 1600, 20
    X, y, _ = synthetic_problem(n, d, seed=3)
    handle.set_data(X, y, G.KERNEL_MATERN, 1.5)
    handle.set_theta(0.3)
    xc = np.random.RandomState(8).uniform(0, 1, (128 * 40 + 3, d))
    xc[777, 5] = np.nan
    xc[3000, 0] = np.nan
    try:
        handle.set_option('sweep_impl', impl)
        val, idx, x, mean, std, acq = handle.sweep(xc, G.ACQ_UCB, float(y.max()), kappa=1.0, dump=True)
    finally:
        handle.set_option('sweep_impl', 3)
    assert idx == 777 and np.isnan(val)
    assert np.isnan(acq[777]) and np.isnan(acq[3000])
    assert np.isfinite(np.delete(acq, [777, 3000])).all()
