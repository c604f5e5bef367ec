"""Host layer: the config stand-in used when polyaxon_schemas is not installed -- the attribute
surface the suggestion path reads, pinned where the reference's tests pin it
(tests/test_experiment_groups/test_search_managers.py:715-782)."""
import numpy as np
import pytest

from polyaxon_b200.schemas import hptuning as H

pytestmark = pytest.mark.skipif(H.HAVE_POLYAXON_SCHEMAS, reason='real polyaxon_schemas in use')


def test_matrix_kinds_and_bounds():
    M = H.MatrixConfig.from_dict
    v = M({'values': [1, 2, 3]})
    assert v.is_discrete and not v.is_categorical and not v.is_continuous and (v.min, v.max) == (1, 3)
    assert v.to_numpy() == [1, 2, 3]
    c = M({'values': ['a', 'b', 'c']})
    assert c.is_categorical and c.is_discrete and c.min is None and c.max is None
    r = M({'range': [1, 5, 1]})
    assert np.array_equal(r.to_numpy(), [1, 2, 3, 4]) and (r.min, r.max) == (1, 5)      # declared stop, not last value
    ls = M({'linspace': [1, 2, 5]})
    assert np.array_equal(ls.to_numpy(), [1., 1.25, 1.5, 1.75, 2.]) and (ls.min, ls.max) == (1, 2)
    u = M({'uniform': [0.001, 0.01]})
    assert u.is_uniform and u.is_continuous and not u.is_discrete and (u.min, u.max) == (0.001, 0.01)
    lu = M({'loguniform': {'low': -3, 'high': 0}})
    assert lu.is_continuous and not lu.is_uniform and lu.min is None
    with pytest.raises(H.ConfigError):
        lu.to_numpy()
    pv = M({'pvalues': [['relu', 0.1], ['sigmoid', 0.8]]})
    assert pv.is_categorical and pv.to_numpy() == ['relu', 'sigmoid']
    assert M({'range': '0:10:2'}).to_dict() == {'range': {'start': 0.0, 'stop': 10.0, 'step': 2.0}}
    with pytest.raises(H.ConfigError):
        M({'values': [1], 'uniform': [0, 1]})
    with pytest.raises(H.ConfigError):
        M({'nonsense': [1]})


def test_matrix_sampling_uses_the_given_generator():
    M = H.MatrixConfig.from_dict
    a = np.random.RandomState(3)
    b = np.random.RandomState(3)
    assert M({'uniform': [0, 1]}).sample(rand_generator=a) == b.uniform(0, 1)
    assert M({'values': [5, 6, 7]}).sample(rand_generator=a) == b.choice([5, 6, 7])
    assert M({'range': [0, 4, 1]}).sample(rand_generator=a) == b.choice(np.arange(0, 4, 1))
    got = M({'pvalues': [['x', 0.25], ['y', 0.75]]}).sample(rand_generator=a)
    assert got == b.choice(['x', 'y'], p=[0.25, 0.75])
    assert M({'loguniform': [-2, 0]}).sample(rand_generator=a) == np.exp(b.uniform(-2, 0))
    assert M({'normal': [0, 1]}).sample(rand_generator=a) == b.normal(0, 1)


def test_utility_function_validation_and_defaults():
    with pytest.raises(H.ConfigError, match='kappa'):
        H.UtilityFunctionConfig(acquisition_function='ucb')
    with pytest.raises(H.ConfigError, match='eps'):
        H.UtilityFunctionConfig(acquisition_function='ei')
    with pytest.raises(H.ConfigError):
        H.UtilityFunctionConfig(acquisition_function='thompson', kappa=1.0)
    uf = H.UtilityFunctionConfig(acquisition_function='poi', eps=0.0)
    gp = uf.gaussian_process
    assert (gp.kernel, gp.length_scale, gp.nu, gp.n_restarts_optimizer) == ('matern', 1.0, 1.5, 0)
    assert uf.n_warmup is None and uf.n_iter is None
    with pytest.raises(H.ConfigError):
        H.GaussianProcessConfig(kernel='periodic')
    assert uf.to_dict()['gaussian_process']['kernel'] == 'matern'


def test_hptuning_round_trip_and_algorithm_selection():
    d = {'seed': 4, 'concurrency': 2,
         'bo': {'n_iterations': 3, 'n_initial_trials': 2, 'metric': {'name': 'loss', 'optimization': 'minimize'},
                'utility_function': {'acquisition_function': 'ucb', 'kappa': 2.5,
                                     'gaussian_process': {'kernel': 'rbf', 'length_scale': 2.0, 'nu': 1.5,
                                                          'n_restarts_optimizer': 1}}},
         'matrix': {'lr': {'uniform': {'low': 0.0, 'high': 1.0}}, 'act': {'values': ['a', 'b']}}}
    cfg = H.HPTuningConfig.from_dict(d)
    assert cfg.search_algorithm == H.SearchAlgorithms.BO and H.SearchAlgorithms.is_bo(cfg.search_algorithm)
    assert H.Optimization.minimize(cfg.bo.metric.optimization)
    assert cfg.to_dict() == d
    assert H.HPTuningConfig.from_dict({'matrix': {'x': {'values': [1]}}}).search_algorithm == H.SearchAlgorithms.GRID
    with pytest.raises(H.ConfigError):
        H.SearchMetricConfig(name='m', optimization='sideways')
    assert H.AcquisitionFunctions.is_ucb('ucb') and H.AcquisitionFunctions.is_ei('ei') and H.AcquisitionFunctions.is_poi('poi')
    assert H.GaussianProcessesKernels.is_rbf('rbf') and H.GaussianProcessesKernels.is_mattern('matern')
