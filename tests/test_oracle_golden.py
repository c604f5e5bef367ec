"""The CPU oracle against the outputs of the real reference (tests/golden/*.npz, written by
oracle/make_goldens.py from /root/reference in the build container)."""
import warnings

import numpy as np
import pytest

from conftest import golden_names, load_golden
from oracle import gp_numpy as G
from oracle import reference_path as R


@pytest.mark.parametrize('name', golden_names())
def test_lml_and_gradient_at_fixed_thetas(name):
    g = load_golden(name)
    for th, lml_ref, grad_ref in zip(g['fixed_thetas'], g['lml_fixed'], g['grad_fixed']):
        lml, grad = G.log_marginal_likelihood(g['X'], g['y'], float(th), int(g['kernel']),
                                              float(g['nu']))
        assert lml == pytest.approx(float(lml_ref), rel=1e-10, abs=1e-9)
        if '_m19_' in name:
            # sklearn differences K numerically with h=1e-10 here; only the sign and
            # magnitude are meaningful
            assert grad == pytest.approx(float(grad_ref), rel=1e-2, abs=1e-2 * abs(float(lml_ref)))
        else:
            assert grad == pytest.approx(float(grad_ref), rel=1e-8, abs=1e-7)


@pytest.mark.parametrize('name', golden_names())
def test_posterior_state_mean_std_acq(name):
    g = load_golden(name)
    kernel, nu, theta = int(g['kernel']), float(g['nu']), float(g['theta'])
    L, a = G.posterior_state(g['X'], g['y'], theta, kernel, nu)
    scale = np.abs(g['alpha']).max()
    assert np.allclose(a, g['alpha'], rtol=1e-9, atol=1e-9 * scale)
    assert np.allclose(np.diag(L), g['L_diag'], rtol=1e-12, atol=0)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        mean, std = G.predict(g['x_tries_head'], g['X'], L, a, theta, kernel, nu)
    assert np.allclose(mean, g['mean'], rtol=1e-9, atol=1e-9 * np.abs(g['mean']).max())
    assert np.allclose(std ** 2, g['std'] ** 2, rtol=1e-9, atol=1e-12)
    acq = G.acquisition(mean, std, int(g['acq']), float(g['y_max']),
                        float(np.nan_to_num(g['kappa'])), float(np.nan_to_num(g['eps'])))
    ok = np.isfinite(g['ys_head'])
    assert np.array_equal(np.isnan(acq), np.isnan(g['ys_head']))
    assert np.allclose(acq[ok], g['ys_head'][ok], rtol=1e-9,
                       atol=1e-9 * max(1.0, np.abs(g['ys_head'][ok]).max()))


@pytest.mark.parametrize('name', golden_names())
def test_full_sweep_argmax_matches_reference(name):
    g = load_golden(name)
    kernel, nu, theta = int(g['kernel']), float(g['nu']), float(g['theta'])
    rs = np.random.RandomState(int(g['seed']))
    for _ in range(int(g['n_restarts'])):
        rs.uniform(*G.THETA_BOUNDS)
    b = g['bounds']
    x_tries = rs.uniform(b[:, 0], b[:, 1], size=(int(g['n_warmup']), b.shape[0]))
    assert x_tries.sum() == float(g['x_tries_sum'])          # RNG draw-order contract
    assert np.array_equal(x_tries[:len(g['x_tries_head'])], g['x_tries_head'])
    L, a = G.posterior_state(g['X'], g['y'], theta, kernel, nu)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        idx, best, _ = G.sweep(x_tries, g['X'], L, a, theta, kernel, nu, int(g['acq']),
                               float(g['y_max']), float(np.nan_to_num(g['kappa'])),
                               float(np.nan_to_num(g['eps'])), chunk=3000)
    assert idx == int(g['argmax'])
    assert best == pytest.approx(float(g['warm_acq']), rel=1e-9)
    x_seeds = rs.uniform(b[:, 0], b[:, 1], size=(int(g['n_iter']), b.shape[0]))
    assert np.array_equal(x_seeds, g['x_seeds'])


@pytest.mark.parametrize('name', ['c1_d2_n15_m25_ucb', 'c5_d16_n128_m25_ei', 'ref10_m15_ucb'])
def test_reference_path_port_reproduces_reference_suggestion(name):
    g = load_golden(name)
    trace = {}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        x = R.suggest(g['X'], g['y'], g['bounds'], int(g['kernel']), float(g['length_scale']),
                      float(g['nu']), int(g['n_restarts']), int(g['acq']),
                      float(np.nan_to_num(g['kappa'])), float(np.nan_to_num(g['eps'])),
                      int(g['seed']), int(g['n_warmup']), int(g['n_iter']), trace=trace)
    assert trace['theta'] == pytest.approx(float(g['theta']), abs=1e-12)
    assert trace['argmax'] == int(g['argmax'])
    assert np.array_equal(x, g['x_star'])
    ref_acq = np.array([r[1] for r in trace['refine_log']])
    assert np.array_equal(ref_acq, g['ref_acq'])


@pytest.mark.parametrize('name', ['c1_d2_n15_m25_ucb', 'c2_d8_n256_rbf_ei', 'c3_d32_n512_m15_poi',
                                  'c5_d16_n128_m05_ucb', 'c5_d16_n128_m25_ei'])
def test_analytic_gradient_against_central_differences(name):
    g = load_golden(name)
    kernel, nu, theta = int(g['kernel']), float(g['nu']), float(g['theta'])
    L, a = G.posterior_state(g['X'], g['y'], theta, kernel, nu)
    args = (g['X'], L, a, theta, kernel, nu, int(g['acq']), float(g['y_max']),
            float(np.nan_to_num(g['kappa'])), float(np.nan_to_num(g['eps'])))
    for x in g['x_seeds'][:2]:
        val, grad = G.acquisition_value_and_grad(x, *args)
        m, s = G.predict(x[None, :], g['X'], L, a, theta, kernel, nu)
        ref = G.acquisition(m, s, int(g['acq']), float(g['y_max']),
                            float(np.nan_to_num(g['kappa'])), float(np.nan_to_num(g['eps'])))[0]
        assert val == pytest.approx(ref, rel=1e-10, abs=1e-14)
        h = 1e-6
        fd = np.empty_like(x)
        for i in range(x.size):
            e = np.zeros_like(x)
            e[i] = h
            fd[i] = (G.acquisition_value_and_grad(x + e, *args)[0]
                     - G.acquisition_value_and_grad(x - e, *args)[0]) / (2 * h)
        assert np.allclose(grad, fd, rtol=2e-5, atol=1e-7 * max(1.0, np.abs(fd).max()))
