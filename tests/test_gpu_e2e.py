"""GPU: the drop-in classes end to end (BOSearchManager -> BOOptimizer -> UtilityFunction) against
what the real reference suggested for the same config, observations and seed (tests/golden)."""
import os
import warnings

import numpy as np
import pytest

from conftest import load_golden
from oracle import gp_numpy as G

pytestmark = pytest.mark.gpu


def _config(g, matrix, optimization='maximize'):
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    acq = {0: 'ucb', 1: 'ei', 2: 'poi'}[int(g['acq'])]
    uf = {'acquisition_function': acq,
          'gaussian_process': {'kernel': 'rbf' if int(g['kernel']) == 0 else 'matern',
                               'length_scale': float(g['length_scale']), 'nu': float(g['nu']),
                               'n_restarts_optimizer': int(g['n_restarts'])},
          'n_warmup': int(g['n_warmup']), 'n_iter': int(g['n_iter'])}
    if acq == 'ucb':
        uf['kappa'] = float(g['kappa'])
    else:
        uf['eps'] = float(g['eps'])
    return HPTuningConfig.from_dict({
        'seed': int(g['seed']), 'concurrency': 2,
        'bo': {'n_iterations': 5, 'n_initial_trials': 5, 'metric': {'name': 'm', 'optimization': optimization},
               'utility_function': uf},
        'matrix': matrix})


def _acq_at(g, theta, x):
    L, a = G.posterior_state(g['X'], g['y'], theta, int(g['kernel']), float(g['nu']))
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        m, s = G.predict(np.atleast_2d(x), g['X'], L, a, theta, int(g['kernel']), float(g['nu']))
    return float(G.acquisition(m, s, int(g['acq']), float(g['y_max']), float(np.nan_to_num(g['kappa'])),
                               float(np.nan_to_num(g['eps'])))[0])


@pytest.mark.parametrize('name', ['c1_d2_n15_m25_ucb', 'c3_d32_n512_m15_poi', 'c5_d16_n128_m05_ucb',
                                  'c5_d16_n128_m25_ei'])
def test_continuous_spaces_through_the_manager(name):
    from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
    from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager
    g = load_golden(name)
    d = g['X'].shape[1]
    keys = ['x%03d' % i for i in range(d)]
    cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
    manager = get_search_algorithm_manager(cfg)
    n = g['X'].shape[0]
    iteration = BOIterationConfig.from_dict({
        'iteration': 1, 'experiment_ids': list(range(n)),
        'experiments_configs': [[i, dict(zip(keys, row.tolist()))] for i, row in enumerate(g['X'])],
        'experiments_metrics': [[i, float(v)] for i, v in enumerate(g['y'])]})
    out = manager.get_suggestions(iteration)
    assert isinstance(out, list) and len(out) == 1
    suggestion = out[0]
    assert sorted(suggestion) == keys
    x = np.array([suggestion[k] for k in keys])
    assert all(isinstance(suggestion[k], np.float64) for k in keys)
    assert np.all(x >= 0) and np.all(x <= 1)
    # same GP (theta* within optimiser tolerance) and an acquisition value no worse than the reference's
    theta = float(g['theta'])
    ref_best = _acq_at(g, theta, g['x_star'])
    mine = _acq_at(g, theta, x)
    assert mine >= ref_best - 1e-6 * max(1.0, abs(ref_best))


def test_mixed_space_decodes_like_the_reference():
    """C2: 3 continuous + 2 discrete + one 3-way categorical, RBF, EI."""
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('c2_d8_n256_rbf_ei')
    acts = ['relu', 'tanh', 'gelu']
    matrix = {'a_lr': {'uniform': [0, 1]}, 'b_mom': {'uniform': [0, 1]}, 'c_wd': {'uniform': [0, 1]},
              'd_layers': {'range': [1, 9, 1]}, 'e_width': {'linspace': [0, 1, 11]},
              'f_act': {'values': acts}}
    cfg = _config(g, matrix)
    configs = [{'a_lr': r[0], 'b_mom': r[1], 'c_wd': r[2], 'd_layers': int(r[3]), 'e_width': float(r[4]),
                'f_act': acts[int(np.argmax(r[5:8]))]} for r in g['X']]
    optimizer = BOOptimizer(hptuning_config=cfg)
    optimizer.add_observations(configs=configs, metrics=g['y'].tolist())
    assert np.array_equal(optimizer.space.x, g['X'])
    assert np.array_equal(optimizer.space.bounds, g['bounds'])
    x = optimizer._maximize()
    gp = optimizer.utility_function.gaussian_process
    assert gp.theta_ == pytest.approx(float(g['theta']), abs=1e-6)
    assert gp.log_marginal_likelihood_value_ == pytest.approx(float(g['lml']), rel=1e-9)
    theta = float(g['theta'])
    assert _acq_at(g, theta, x) >= _acq_at(g, theta, g['x_star']) - 1e-6
    suggestion = optimizer.space.get_suggestion(x)
    assert suggestion['f_act'] in acts
    assert suggestion['d_layers'] in list(range(1, 9))
    assert any(abs(suggestion['e_width'] - v) < 1e-12 for v in np.linspace(0, 1, 11))
    assert 0 <= suggestion['a_lr'] <= 1


def test_reference_fixture_sets_run_and_stay_in_bounds():
    """The observation sets of test_search_managers.py:894-1063 (closed-form nu instead of 1.9):
    same assertions as the reference's own tests -- bounds membership."""
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('ref10_m15_ucb')
    matrix = {'learning_rate': {'uniform': [0.001, 0.01]}, 'dropout': {'values': [0.25, 0.3]},
              'activation': {'values': ['relu', 'sigmoid']}}
    cfg = _config(g, matrix, optimization='minimize')
    # columns: activation one-hot (relu, sigmoid), dropout, learning_rate
    configs = [{'activation': ['relu', 'sigmoid'][int(np.argmax(r[0:2]))], 'dropout': float(r[2]),
                'learning_rate': float(r[3])} for r in g['X']]
    optimizer = BOOptimizer(hptuning_config=cfg)
    optimizer.add_observations(configs=configs, metrics=(-g['y']).tolist())
    assert np.array_equal(optimizer.space.x, g['X'])
    suggestion = optimizer.get_suggestion()
    assert 0.001 <= suggestion['learning_rate'] <= 0.01
    assert suggestion['dropout'] in [0.25, 0.3]
    assert suggestion['activation'] in ['relu', 'sigmoid']
    assert optimizer.utility_function.gaussian_process.theta_ == pytest.approx(float(g['theta']), abs=1e-6)


def test_resident_candidate_cache_is_transparent(monkeypatch):
    """Second call of a seeded group reuses the device-resident candidate matrix: same suggestion,
    same generator state afterwards, as with the cache switched off."""
    from polyaxon_b200 import device
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('c5_d16_n128_m25_ei')
    d = g['X'].shape[1]
    keys = ['x%03d' % i for i in range(d)]
    cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
    configs = [dict(zip(keys, row.tolist())) for row in g['X']]

    def run():
        opt = BOOptimizer(hptuning_config=cfg)
        opt.add_observations(configs=configs, metrics=g['y'].tolist())
        s = opt.get_suggestion()
        return s, opt.utility_function.random_generator.get_state()[1].copy()

    device.get_handle().seed_cache.clear()
    device.get_handle().resident_key = None
    first, state1 = run()
    key = device.get_handle().resident_key
    assert key is not None and key in device.get_handle().seed_cache
    uploads = {'n': 0}
    from polyaxon_b200 import _lib
    real_set = _lib.Handle.candidates_set

    def counting_set(self, *a, **k):
        uploads['n'] += 1
        return real_set(self, *a, **k)
    monkeypatch.setattr(_lib.Handle, 'candidates_set', counting_set)
    second, state2 = run()
    assert uploads['n'] == 0                           # neither drawn nor copied again
    assert device.get_handle().resident_key == key
    monkeypatch.setenv('PLX_BO_CANDIDATE_CACHE', '0')
    third, state3 = run()
    assert first == second == third
    assert np.array_equal(state1, state2) and np.array_equal(state1, state3)


@pytest.mark.parametrize('d,n_cand,pre_draws', [(1, 1, 0), (2, 1000, 0), (7, 12345, 3), (32, 40000, 16), (5, 311, 1)])
def test_device_mt19937_reproduces_numpy_bit_for_bit(d, n_cand, pre_draws):
    """x_tries drawn on the GPU == RandomState.uniform(low, high, size=(N, d)) (acquisition_function.py:88-89),
    and the host generator is left in the state numpy would have reached."""
    from polyaxon_b200 import _lib
    h = _lib.Handle(0)
    try:
        lo = np.linspace(-3.0, 2.0, d)
        hi = lo + np.linspace(0.5, 7.0, d)
        a = np.random.RandomState(4242)
        b = np.random.RandomState(4242)
        for _ in range(pre_draws):                        # e.g. restart thetas drawn during fit
            a.uniform(-11.5, 11.5)
            b.uniform(-11.5, 11.5)
        want = a.uniform(lo, hi, size=(n_cand, d))
        assert h.candidates_draw(b, lo, hi, n_cand)
        got = h.candidates_get(0, n_cand, d)
        assert np.array_equal(got, want)
        sa, sb = a.get_state(), b.get_state()
        assert np.array_equal(sa[1], sb[1]) and sa[2] == sb[2]
        assert np.array_equal(a.uniform(size=5), b.uniform(size=5))
        # an odd stream position (a 32-bit draw happened) is refused: the caller draws on the host
        c = np.random.RandomState(1)
        c.randint(0, 10)
        if c.get_state()[2] & 1:
            assert not h.candidates_draw(c, lo, hi, n_cand)
    finally:
        h.close()


@pytest.mark.parametrize('d,n_cand,pre_draws,world', [(32, 2000000, 16, 2), (3, 5000000, 0, 1), (32, 300000, 2, 3),
                                                      (1, 700, 1, 2), (7, 9, 0, 4), (2, 2, 0, 4)])
def test_device_mt19937_draws_any_block_of_rows(d, n_cand, pre_draws, world):
    """The sharded draw (plx_candidates_draw_mt19937_range): every rank of a multi-GPU job draws only its block of
    rows of the SAME numpy matrix by jumping ahead in the MT19937 stream (csrc/mt19937.cuh), the blocks are cut
    into per-CTA chunks that jump as well (the first two shapes are large enough for that), and every rank ends with
    the generator state numpy reaches after the whole matrix."""
    from polyaxon_b200 import _lib, dist as pdist
    h = _lib.Handle(0)
    try:
        lo = np.linspace(-3.0, 2.0, d)
        hi = lo + np.linspace(0.5, 7.0, d)
        a = np.random.RandomState(99)
        for _ in range(pre_draws):
            a.uniform(-11.5, 11.5)
        start = a.get_state()
        want = a.uniform(lo, hi, size=(n_cand, d))
        end = a.get_state()
        for rank in range(world):
            first, stop = pdist.shard_range(n_cand, rank, world)
            b = np.random.RandomState()
            b.set_state(start)
            assert h.candidates_draw(b, lo, hi, n_cand, first=first, count=stop - first)
            got = h.candidates_get(first, stop - first, d)
            assert np.array_equal(got, want[first:stop]), rank
            sb = b.get_state()
            assert np.array_equal(sb[1], end[1]) and sb[2] == end[2], rank
            if stop > first:
                with pytest.raises(_lib.PlxError):
                    h.candidates_get(stop, 1, d)              # outside the block this rank holds
    finally:
        h.close()


def test_device_rng_and_host_rng_give_the_same_suggestion(monkeypatch):
    from polyaxon_b200 import device
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('c1_d2_n15_m25_ucb')
    keys = ['x000', 'x001']
    cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
    configs = [dict(zip(keys, row.tolist())) for row in g['X']]
    out = []
    for flag in ('1', '0'):
        monkeypatch.setenv('PLX_BO_DEVICE_RNG', flag)
        device.get_handle().seed_cache.clear()
        opt = BOOptimizer(hptuning_config=cfg)
        opt.add_observations(configs=configs, metrics=g['y'].tolist())
        out.append((opt.get_suggestion(), opt.utility_function.random_generator.get_state()[1].copy()))
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1], out[1][1])


def test_suggest_groups_runs_independent_groups_in_order():
    """C5 shape (scaled down): independent groups through the batch helper equal the one-by-one calls."""
    from polyaxon_b200.groups import suggest_groups
    from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
    from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager
    from oracle.reference_path import synthetic_problem
    cfgs, its = [], []
    g = load_golden('c5_d16_n128_m25_ei')
    keys = ['x%03d' % i for i in range(16)]
    for gid in range(3):
        X, y, _ = synthetic_problem(64, 16, seed=1234 + gid)
        cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
        cfg.seed = 100 + gid
        cfgs.append(cfg)
        its.append(BOIterationConfig.from_dict({
            'iteration': 1, 'experiment_ids': list(range(64)),
            'experiments_configs': [[i, dict(zip(keys, row.tolist()))] for i, row in enumerate(X)],
            'experiments_metrics': [[i, float(v)] for i, v in enumerate(y)]}))
    together = suggest_groups(cfgs, its)                 # three worker threads, a handle each
    # BOIterationConfig.combined_* extends its lists in place (as the reference's does), so rebuild them
    its2 = [BOIterationConfig.from_dict({'iteration': 1, 'experiment_ids': it.experiment_ids,
                                         'experiments_configs': it.experiments_configs,
                                         'experiments_metrics': it.experiments_metrics}) for it in its]
    alone = [get_search_algorithm_manager(c).get_suggestions(i) for c, i in zip(cfgs, its2)]
    assert together == alone
    assert len({tuple(sorted(s[0].items())) for s in together}) == 3
    its3 = [BOIterationConfig.from_dict({'iteration': 1, 'experiment_ids': it.experiment_ids,
                                         'experiments_configs': it.experiments_configs,
                                         'experiments_metrics': it.experiments_metrics}) for it in its2]
    assert suggest_groups(cfgs, its3, workers=1) == alone


def test_hold_theta_between_refits_is_opt_in(monkeypatch):
    """PLX_BO_REFIT_EVERY=k > 1: calls whose observations extend the previous call's keep theta and
    extend the posterior state on the device; the default refits every call like the reference."""
    from polyaxon_b200 import _lib, device
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('c5_d16_n128_m25_ei')
    d = g['X'].shape[1]
    keys = ['x%03d' % i for i in range(d)]
    cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
    configs = [dict(zip(keys, row.tolist())) for row in g['X']]
    metrics = g['y'].tolist()
    fits = {'n': 0}
    real_fit = _lib.Handle.fit

    def counting_fit(self, *a, **k):
        fits['n'] += 1
        return real_fit(self, *a, **k)
    monkeypatch.setattr(_lib.Handle, 'fit', counting_fit)

    def run(n):
        opt = BOOptimizer(hptuning_config=cfg)
        opt.add_observations(configs=configs[:n], metrics=metrics[:n])
        s = opt.get_suggestion()
        return s, opt.utility_function.gaussian_process

    n = len(configs)
    run(n - 3); run(n - 2)
    assert fits['n'] == 2                              # default: a fit per call
    monkeypatch.setenv('PLX_BO_REFIT_EVERY', '3')
    fits['n'] = 0
    _, gp0 = run(n - 3)                                # fit
    s1, gp1 = run(n - 2)                               # +1 row, theta held
    s2, gp2 = run(n)                                   # +2 rows, theta held
    assert fits['n'] == 1
    assert gp1.theta_ == gp0.theta_ == gp2.theta_
    h = device.get_handle()
    assert h.n == n
    _, alpha, Z = h.get_state(want_L=False, want_Linv=True)
    ref = _lib.Handle(0)
    try:
        ref.set_data(g['X'], g['y'], int(g['kernel']), float(g['nu']))
        ref.set_theta(gp0.theta_)
        _, alpha_ref, Z_ref = ref.get_state(want_L=False, want_Linv=True)
    finally:
        ref.close()
    tol = max(1e-9, 100 * np.finfo(float).eps * float(g['cond']))
    assert np.max(np.abs(alpha - alpha_ref)) <= tol * np.max(np.abs(alpha_ref))
    assert np.max(np.abs(Z - Z_ref)) <= tol * np.max(np.abs(Z_ref))
    for s in (s1, s2):
        assert set(s) == set(keys) and all(0.0 <= v <= 1.0 for v in s.values())
    _, gp3 = run(n)                                    # third call since the fit: refit
    assert fits['n'] == 2
    # a different observation history never takes the shortcut
    fits['n'] = 0
    run(n - 1)
    assert fits['n'] == 1


def test_unseeded_group_uses_and_advances_the_global_numpy_generator():
    """seed=None: the reference draws from the numpy module itself (utils.py:37-38).  The device draw
    must consume exactly the numbers numpy would have produced and leave the global state where the
    host draw leaves it."""
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    g = load_golden('c1_d2_n15_m25_ucb')
    keys = ['x000', 'x001']
    cfg = _config(g, {k: {'uniform': [0, 1]} for k in keys})
    cfg.seed = None
    configs = [dict(zip(keys, row.tolist())) for row in g['X']]
    out = []
    for flag in ('1', '0'):
        os.environ['PLX_BO_DEVICE_RNG'] = flag
        try:
            np.random.seed(4242)
            opt = BOOptimizer(hptuning_config=cfg)
            assert opt.utility_function.random_generator is np.random
            opt.add_observations(configs=configs, metrics=g['y'].tolist())
            out.append((opt.get_suggestion(), np.random.get_state()[1].copy(), np.random.get_state()[2]))
        finally:
            del os.environ['PLX_BO_DEVICE_RNG']
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1], out[1][1]) and out[0][2] == out[1][2]
