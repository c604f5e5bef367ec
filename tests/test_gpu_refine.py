"""GPU: analytic acquisition gradient and the batched refinement (acquisition_function.py:95-111).

Contract (BASELINE.json): the refined suggestion must reach an acquisition value within 1e-6 of
scipy's.  The golden files hold what scipy's L-BFGS-B (finite-difference gradients) reached from
every seed inside the real reference run; the device optimiser starts from the same seeds.
Values claimed by the device are re-evaluated with the CPU oracle before they are compared.
"""
import warnings

import numpy as np
import pytest

from conftest import CLOSED_FORM, load_golden
from oracle import gp_numpy as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def handle():
    from polyaxon_b200 import _lib
    h = _lib.Handle(0)
    yield h
    h.close()


def _args(g):
    return dict(acq=int(g['acq']), y_max=float(g['y_max']), kappa=float(np.nan_to_num(g['kappa'])),
                eps=float(np.nan_to_num(g['eps'])))


def _oracle_acq(g, L, a, x):
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        m, s = G.predict(np.atleast_2d(x), g['X'], L, a, float(g['theta']), int(g['kernel']), float(g['nu']))
        return G.acquisition(m, s, int(g['acq']), float(g['y_max']), float(np.nan_to_num(g['kappa'])),
                             float(np.nan_to_num(g['eps'])))


WELL = [n for n in CLOSED_FORM if not n.startswith('ref1')]


@pytest.mark.parametrize('name', WELL)
def test_value_and_gradient_against_oracle(handle, name):
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    handle.set_theta(float(g['theta']))
    L, a = G.posterior_state(g['X'], g['y'], float(g['theta']), int(g['kernel']), float(g['nu']))
    pts = np.vstack([g['x_seeds'], g['x_tries_head'][:5]])
    vals, grads = handle.acq_value_grad(pts, **_args(g))
    for x, v, gr in zip(pts, vals, grads):
        rv, rg = G.acquisition_value_and_grad(x, g['X'], L, a, float(g['theta']), int(g['kernel']),
                                              float(g['nu']), int(g['acq']), float(g['y_max']),
                                              float(np.nan_to_num(g['kappa'])), float(np.nan_to_num(g['eps'])))
        assert v == pytest.approx(rv, rel=1e-9, abs=1e-13)
        assert np.allclose(gr, rg, rtol=1e-7, atol=1e-9 * max(1.0, np.abs(rg).max()))


@pytest.mark.parametrize('name', CLOSED_FORM)
def test_refinement_reaches_scipys_value(handle, name):
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    handle.set_theta(float(g['theta']))
    L, a = G.posterior_state(g['X'], g['y'], float(g['theta']), int(g['kernel']), float(g['nu']))
    b = g['bounds']
    xs, vals, ok, nfev = handle.refine(g['x_seeds'], b, want_nfev=True, **_args(g))
    assert np.all(xs >= b[:, 0] - 1e-15) and np.all(xs <= b[:, 1] + 1e-15)
    # what the device reports is what the oracle computes at that point
    for x, v, success in zip(xs, vals, ok):
        if success and float(g['cond']) < 1e8:
            assert v == pytest.approx(float(_oracle_acq(g, L, a, x)[0]), rel=1e-8, abs=1e-12)

    # acceptance rule of acquisition_function.py:105-111 on both sides
    def accept(start, values, success):
        best = start
        for v, s in zip(values, success):
            if s and v >= best:
                best = v
        return best
    mine = accept(float(g['warm_acq']), vals, ok)
    ref = accept(float(g['warm_acq']), g['ref_acq'], g['ref_success'])
    assert mine >= ref - 1e-6 * max(1.0, abs(ref))
    # per seed: where both converged the analytic-gradient run is at least as good as scipy's
    both = ok & g['ref_success']
    if float(g['cond']) < 1e8:
        assert np.all(vals[both] >= g['ref_acq'][both] - 1e-6 * np.maximum(1.0, np.abs(g['ref_acq'][both])))


def test_refine_edge_cases(handle):
    g = load_golden('c1_d2_n15_m25_ucb')
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    handle.set_theta(float(g['theta']))
    b = g['bounds']
    xs, vals, ok = handle.refine(np.zeros((0, 2)), b, **_args(g))
    assert xs.shape == (0, 2)
    # seeds outside the box are clipped first, like scipy does with x0
    xs, vals, ok = handle.refine(np.array([[-3.0, 9.0], [0.5, 0.5]]), b, **_args(g))
    assert np.all(xs >= b[:, 0]) and np.all(xs <= b[:, 1])
    # a degenerate box (lo == hi) pins the variable
    b2 = b.copy()
    b2[0] = [0.3, 0.3]
    xs, vals, ok = handle.refine(np.array([[0.3, 0.9]]), b2, **_args(g))
    assert xs[0, 0] == 0.3 and ok[0]
    # many seeds at once (more than one register block) agree with the same seeds run alone
    rs = np.random.RandomState(4)
    seeds = rs.uniform(0, 1, (37, 2))
    xa, va, oa = handle.refine(seeds, b, **_args(g))
    xb, vb, ob = handle.refine(seeds[20:21], b, **_args(g))
    assert np.array_equal(xa[20], xb[0]) and va[20] == vb[0]


def test_wide_one_hot_space_is_refined(handle):
    """A search space whose encoded width is far beyond 128 (categoricals widen it quickly: one column per
    choice).  The reference's scipy loop has no width limit (acquisition_function.py:95-111); the device
    refinement keeps its per-seed gradient in 16 d bytes of shared memory and takes d up to 2048."""
    from oracle.reference_path import synthetic_problem
    d, n = 300, 40
    X, y, bounds = synthetic_problem(n, d, seed=8)
    theta = 1.5
    handle.set_data(X, y, G.KERNEL_MATERN, 2.5)
    handle.set_theta(theta)
    L, a = G.posterior_state(X, y, theta, G.KERNEL_MATERN, 2.5)
    seeds = np.random.RandomState(2).uniform(0, 1, (6, d))
    y_max = float(y.max())
    vals, grads = handle.acq_value_grad(seeds, G.ACQ_EI, y_max, eps=0.0)
    for x, v, gr in zip(seeds, vals, grads):
        rv, rg = G.acquisition_value_and_grad(x, X, L, a, theta, G.KERNEL_MATERN, 2.5, G.ACQ_EI, y_max, 0.0, 0.0)
        assert v == pytest.approx(rv, rel=1e-9, abs=1e-13)
        assert np.allclose(gr, rg, rtol=1e-7, atol=1e-9 * max(1.0, np.abs(rg).max()))
    xs, got, ok = handle.refine(seeds, bounds, G.ACQ_EI, y_max, eps=0.0)
    assert ok.all()
    assert np.all(xs >= 0.0) and np.all(xs <= 1.0)
    for x, v, v0 in zip(xs, got, vals):
        m, s = G.predict(x[None, :], X, L, a, theta, G.KERNEL_MATERN, 2.5)
        assert v == pytest.approx(G.acquisition(m, s, G.ACQ_EI, y_max, 0.0, 0.0)[0], rel=1e-9, abs=1e-13)
        assert v >= v0 - 1e-12                     # never worse than where it started


@pytest.mark.parametrize('name', [n for n in WELL if 'n2048' not in n and 'n512' not in n])
def test_one_launch_refinement_equals_round_by_round(handle, name):
    """Up to 256 observations every seed is refined from start to finish in one launch (refine_persist_kernel);
    larger sets take four launches per round.  Same state machine, same arithmetic up to the summation order of
    one matrix-vector product: the two must land on the same optimum."""
    g = load_golden(name)
    handle.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
    handle.set_theta(float(g['theta']))
    seeds = np.vstack([g['x_seeds'], g['x_tries_head'][:6]])
    out = {}
    try:
        for flag in (1, 0):
            handle.set_option('refine_persist', flag)
            out[flag] = handle.refine(seeds, g['bounds'], want_nfev=True, **_args(g))
    finally:
        handle.set_option('refine_persist', 1)
    assert np.array_equal(out[1][2], out[0][2])                       # success flags
    assert np.allclose(out[1][1], out[0][1], rtol=1e-8, atol=1e-10)   # values reached
    assert np.allclose(out[1][0], out[0][0], atol=1e-5)               # points reached
