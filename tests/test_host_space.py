"""Host layer: the reference's own API-shape tests for the Bayesian-optimisation managers, restated
(tests/test_experiment_groups/test_search_managers.py:631-879, test_schemas.py:70-86).  No GPU: the
arithmetic entry points are patched out exactly where the reference's tests patch them."""
from unittest.mock import patch

import numpy as np
import pytest

from polyaxon_b200.hpsearch.schemas import get_iteration_config
from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import UtilityFunction
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.manager import BOSearchManager
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.space import SearchSpace
from polyaxon_b200.hpsearch.search_managers.utils import Suggestion, get_random_suggestions
from polyaxon_b200.schemas.hptuning import HPTuningConfig, SearchAlgorithms


def config1():
    return HPTuningConfig.from_dict({
        'concurrency': 2,
        'bo': {'n_iterations': 5, 'n_initial_trials': 5,
               'metric': {'name': 'loss', 'optimization': 'minimize'},
               'utility_function': {'acquisition_function': 'ucb', 'kappa': 1.2,
                                    'gaussian_process': {'kernel': 'matern', 'length_scale': 1.0,
                                                         'nu': 1.9, 'n_restarts_optimizer': 0}}},
        'matrix': {'feature1': {'values': [1, 2, 3]}, 'feature2': {'linspace': [1, 2, 5]},
                   'feature3': {'range': [1, 5, 1]}}})


def config2():
    return HPTuningConfig.from_dict({
        'concurrency': 2,
        'bo': {'n_iterations': 4, 'n_initial_trials': 4,
               'metric': {'name': 'accuracy', 'optimization': 'maximize'},
               'utility_function': {'acquisition_function': 'ei', 'eps': 1.2,
                                    'gaussian_process': {'kernel': 'matern', 'length_scale': 1.0,
                                                         'nu': 1.9, 'n_restarts_optimizer': 0}}},
        'matrix': {'feature1': {'values': [1, 2, 3, 4, 5]}, 'feature2': {'linspace': [1, 5, 5]},
                   'feature3': {'range': [1, 6, 1]}, 'feature4': {'uniform': [1, 5]},
                   'feature5': {'values': ['a', 'b', 'c']}}})


def test_manager_dispatch_and_algorithm_check():
    manager = get_search_algorithm_manager(config1())
    assert isinstance(manager, BOSearchManager) and manager.NAME == SearchAlgorithms.BO
    assert get_search_algorithm_manager(None) is None
    grid = HPTuningConfig.from_dict({'matrix': {'lr': {'values': [1, 2]}}})
    assert get_search_algorithm_manager(grid).NAME == SearchAlgorithms.GRID
    with pytest.raises(ValueError):
        BOSearchManager(hptuning_config=grid)
    assert manager.should_reschedule(4) and not manager.should_reschedule(5)


def test_first_get_suggestions_returns_initial_random_suggestion():
    assert len(BOSearchManager(config1()).get_suggestions()) == 5
    assert len(BOSearchManager(config2()).get_suggestions()) == 4


def test_random_suggestions_are_distinct_seeded_and_capped():
    cfg = config1()
    cfg.seed = 42
    a = get_random_suggestions(cfg.matrix, 5, seed=42)
    b = get_random_suggestions(cfg.matrix, 5, seed=42)
    assert a == b and len({Suggestion(p) for p in a}) == 5
    small = HPTuningConfig.from_dict({'matrix': {'f': {'values': [1, 2, 3]}}})
    assert len(get_random_suggestions(small.matrix, 10)) == 3       # all-discrete space of size 3
    with pytest.raises(ValueError):
        get_random_suggestions(small.matrix, 0)


def test_iteration_suggestions_calls_optimizer():
    iteration_config = BOIterationConfig.from_dict({
        'iteration': 2,
        'old_experiment_ids': [1, 2, 3],
        'old_experiments_configs': [[1, {'feature1': 1, 'feature2': 1, 'feature3': 1}],
                                    [2, {'feature1': 2, 'feature2': 1.2, 'feature3': 2}],
                                    [3, {'feature1': 3, 'feature2': 1.3, 'feature3': 3}]],
        'old_experiments_metrics': [[1, 1], [2, 2], [3, 3]],
        'experiment_ids': [4],
        'experiments_configs': [[4, {'feature1': 2, 'feature2': 1.5, 'feature3': 4}]],
        'experiments_metrics': [[4, 4]]})
    with patch.object(BOOptimizer, 'get_suggestion') as get_suggestion_mock:
        BOSearchManager(config1()).get_suggestions(iteration_config)
    assert get_suggestion_mock.call_count == 1


def test_only_experiments_with_metrics_become_observations():
    iteration_config = BOIterationConfig.from_dict({
        'iteration': 1, 'experiment_ids': [7, 8, 9],
        'experiments_configs': [[7, {'feature1': 1, 'feature2': 1, 'feature3': 1}],
                                [8, {'feature1': 2, 'feature2': 1.25, 'feature3': 2}],
                                [9, {'feature1': 3, 'feature2': 2, 'feature3': 4}]],
        'experiments_metrics': [[9, 0.5], [7, 0.1]]})
    seen = {}

    def record(self, configs, metrics):
        seen['configs'], seen['metrics'] = configs, metrics

    with patch.object(BOOptimizer, 'add_observations', record), \
            patch.object(BOOptimizer, 'get_suggestion', return_value={'feature1': 1}):
        out = BOSearchManager(config1()).get_suggestions(iteration_config)
    assert out == [{'feature1': 1}]
    assert seen['metrics'] == [0.5, 0.1]
    assert [c['feature1'] for c in seen['configs']] == [3, 1]


def test_space_search():
    space1 = SearchSpace(hptuning_config=config1())
    assert space1.dim == 3 and len(space1.bounds) == 3
    assert len(space1.discrete_features) == 3 and len(space1.categorical_features) == 0
    assert space1.features == ['feature1', 'feature2', 'feature3']
    assert np.all(space1.bounds == [[1, 3], [1, 2], [1, 5]])
    assert np.all(space1.discrete_features['feature1']['values'] == [1, 2, 3])
    assert np.all(space1.discrete_features['feature2']['values'] == np.asarray([1., 1.25, 1.5, 1.75, 2.]))
    assert np.all(space1.discrete_features['feature3']['values'] == np.asarray([1, 2, 3, 4]))

    space2 = SearchSpace(hptuning_config=config2())
    assert space2.dim == 7 and len(space2.bounds) == 7 and len(space2.features) == 5
    assert len(space2.discrete_features) == 3 and len(space2.categorical_features) == 1
    assert np.all(space2.bounds == [[1, 5], [1, 5], [1, 6], [1, 5], [0, 1], [0, 1], [0, 1]])
    assert space2.categorical_features == {'feature5': {'values': ['a', 'b', 'c'], 'number': 3}}
    assert space2.discrete_features['feature1']['values'] == [1, 2, 3, 4, 5]
    assert np.all(space2.discrete_features['feature2']['values'] == np.asarray([1, 2, 3, 4, 5]))
    assert np.all(space2.discrete_features['feature3']['values'] == np.asarray([1, 2, 3, 4, 5]))


def test_add_observation_to_space_search():
    space1 = SearchSpace(hptuning_config=config1())
    assert space1.x == [] and space1.y == []
    assert not space1.is_observations_valid()
    space1.add_observations(configs=[{'feature1': 1, 'feature2': 1, 'feature3': 1},
                                     {'feature1': 2, 'feature2': 1.2, 'feature3': 2},
                                     {'feature1': 3, 'feature2': 1.3, 'feature3': 3}], metrics=[1, 2, 3])
    assert space1.is_observations_valid()
    assert np.all(space1.x == [[1, 1, 1], [2, 1.2, 2], [3, 1.3, 3]])
    assert np.all(space1.y == np.array([-1, -2, -3]))            # minimise -> negated

    space2 = SearchSpace(hptuning_config=config2())
    space2.add_observations(
        configs=[{'feature1': 1, 'feature2': 1, 'feature3': 1, 'feature4': 1, 'feature5': 'a'},
                 {'feature1': 2, 'feature2': 1.2, 'feature3': 2, 'feature4': 4, 'feature5': 'b'},
                 {'feature1': 3, 'feature2': 1.3, 'feature3': 3, 'feature4': 3, 'feature5': 'a'}],
        metrics=[1, 2, 3])
    assert np.all(space2.x == [[1, 1, 1, 1, 1, 0, 0], [2, 1.2, 2, 4, 0, 1, 0], [3, 1.3, 3, 3, 1, 0, 0]])
    assert np.all(space2.y == np.array([1, 2, 3]))


def test_space_get_suggestion():
    space1 = SearchSpace(hptuning_config=config1())
    assert space1.get_suggestion(None) is None
    assert space1.get_suggestion([1, 1, 1]) == {'feature1': 1, 'feature2': 1, 'feature3': 1}
    assert space1.get_suggestion([1, 1.2, 2]) == {'feature1': 1, 'feature2': 1.25, 'feature3': 2}
    assert space1.get_suggestion([1, 1.5, 3]) == {'feature1': 1, 'feature2': 1.5, 'feature3': 3}
    space2 = SearchSpace(hptuning_config=config2())
    assert space2.get_suggestion([1, 1, 1, 1, 1, 0, 0]) == {
        'feature1': 1, 'feature2': 1, 'feature3': 1, 'feature4': 1, 'feature5': 'a'}
    assert space2.get_suggestion([1, 1.2, 2, 3, 0, 0, 1]) == {
        'feature1': 1, 'feature2': 1, 'feature3': 2, 'feature4': 3, 'feature5': 'c'}
    assert space2.get_suggestion([1, 1.8, 3, 3, 0, 1, 0]) == {
        'feature1': 1, 'feature2': 2, 'feature3': 3, 'feature4': 3, 'feature5': 'b'}


def test_optimizer_defaults_and_add_observations_delegation():
    optimizer = BOOptimizer(hptuning_config=config1())
    assert (optimizer.n_warmup, optimizer.n_iter) == (5, 10)      # optimizer.py:13-14
    with patch.object(SearchSpace, 'add_observations') as add_observations_mock:
        optimizer.add_observations(configs=[], metrics=[])
    assert add_observations_mock.call_count == 1
    assert optimizer.get_suggestion() is None                     # no observations -> None


def test_utility_function_rejects_foreign_configs():
    with pytest.raises(ValueError, match='Received a non valid configuration.'):
        UtilityFunction(config={'acquisition_function': 'ucb'})
    with pytest.raises(ValueError, match='Received a non valid configuration.'):
        UtilityFunction.get_gaussian_process(config={}, random_generator=None)
    uf = UtilityFunction(config=config1().bo.utility_function, seed=3)
    assert uf.kappa == 1.2 and uf.eps is None and uf.acquisition_function == 'ucb'
    assert uf.gaussian_process.random_state is uf.random_generator       # shared RNG object
    assert UtilityFunction(config=config1().bo.utility_function, seed=0).random_generator is np.random


def test_bo_iteration_config_round_trip():
    cfg = {'iteration': 1, 'old_experiment_ids': [1, 2, 3],
           'old_experiments_configs': [[1, {'a': 1}], [2, {'a': 2}], [3, {'a': 3}]],
           'old_experiments_metrics': [[1, 0.1], [2, 0.2], [3, 0.3]],
           'experiment_ids': [4, 5], 'experiments_configs': [[4, {'a': 4}], [5, {'a': 5}]],
           'experiments_metrics': [[4, 0.4], [5, 0.5]]}
    assert BOIterationConfig.from_dict(cfg).to_dict() == cfg
    bare = BOIterationConfig.from_dict({'iteration': 0}).to_dict()      # None fields are dumped too
    assert bare['iteration'] == 0 and len(bare) == 7 and all(v is None for k, v in bare.items() if k != 'iteration')
    it = get_iteration_config(SearchAlgorithms.BO, cfg)
    assert it.combined_experiment_ids == [1, 2, 3, 4, 5]
    assert dict(it.combined_experiments_metrics)[5] == 0.5
    with pytest.raises(ValueError):
        get_iteration_config(SearchAlgorithms.BO, None)
    with pytest.raises(ValueError):
        BOIterationConfig.from_dict({'iteration': 1, 'experiments_metrics': [[1, 2, 3]]})


def test_models_sharing_a_handle_put_their_own_state_back():
    """every model of a thread shares one C-ABI handle; in the reference each UtilityFunction owns its
    GaussianProcessRegressor.  A model that finds another model's state on the handle re-establishes its own
    before predicting, and an unfitted model refuses."""
    import numpy as np
    import pytest
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import DeviceGaussianProcess

    class Handle(object):
        device = 0

        def __init__(self):
            self.calls, self.owner, self.held = [], None, None

        def set_data(self, X, y, kernel, nu, jitter):
            self.calls.append(('set_data', float(X[0, 0])))
            self.owner = None

        def fit(self, theta0, restarts, bounds, apply=True):
            self.calls.append(('fit',))
            return 0.5, -1.0, np.array([[theta0, 0.5, -1.0, 3.0]])

        def set_theta(self, theta):
            self.calls.append(('set_theta', theta))

        def sweep(self, X, acq, y_max, kappa=0.0, dump=False):
            self.calls.append(('sweep',))
            return 0.0, 0, X[0], np.zeros(len(X)), np.ones(len(X)), np.zeros(len(X))

        def last_sweep_clamped(self):
            return 0

    h = Handle()
    a = DeviceGaussianProcess('rbf', 1.0, 1.5, 0, np.random.RandomState(1), handle=h)
    b = DeviceGaussianProcess('rbf', 1.0, 1.5, 0, np.random.RandomState(1), handle=h)
    with pytest.raises(RuntimeError):
        a.predict(np.zeros((1, 2)))
    a.fit(np.full((3, 2), 1.0), np.arange(3.0))
    b.fit(np.full((3, 2), 2.0), np.arange(3.0))
    del h.calls[:]
    b.predict(np.zeros((1, 2)))                            # b's state is on the handle: nothing to restore
    assert h.calls == [('sweep',)]
    a.predict(np.zeros((1, 2)))                            # a's is not: data and theta go back first
    assert h.calls[1:] == [('set_data', 1.0), ('set_theta', 0.5), ('sweep',)]
    del h.calls[:]
    a.predict(np.zeros((1, 2)))
    assert h.calls == [('sweep',)]
