"""Grid / random / Hyperband managers and the iteration managers: host bookkeeping that shares the
``get_search_algorithm_manager`` entry point with the Bayesian-optimisation manager.  Expectations are
the reference's own (tests/test_experiment_groups/test_search_managers.py:59-629,
tests/test_experiment_groups/test_iteration_managers.py), with a plain object standing in for the
Django experiment group."""
from unittest.mock import patch

import pytest

from polyaxon_b200.hpsearch.iteration_managers import (
    BOIterationManager,
    HyperbandIterationManager,
    get_search_iteration_manager,
)
from polyaxon_b200.hpsearch.schemas import HyperbandIterationConfig, get_iteration_config
from polyaxon_b200.hpsearch.search_managers import (
    BOSearchManager,
    GridSearchManager,
    HyperbandSearchManager,
    RandomSearchManager,
    get_search_algorithm_manager,
)
from polyaxon_b200.schemas.hptuning import HPTuningConfig, MatrixConfig

MATRIX3 = {'feature1': {'values': [1, 2, 3]}, 'feature2': {'linspace': [1, 2, 5]}, 'feature3': {'range': [1, 5, 1]}}


def _hyperband(max_iter, resource, matrix):
    return HPTuningConfig.from_dict({
        'concurrency': 2,
        'hyperband': {'max_iter': max_iter, 'eta': 3, 'resource': resource, 'resume': False,
                      'metric': {'name': 'loss', 'optimization': 'minimize'}},
        'matrix': matrix})


@pytest.fixture
def manager1():
    return HyperbandSearchManager(_hyperband(10, {'name': 'steps', 'type': 'float'}, MATRIX3))


@pytest.fixture
def manager2():
    return HyperbandSearchManager(_hyperband(81, {'name': 'size', 'type': 'int'},
                                             dict(MATRIX3, feature4={'range': [1, 5, 1]})))


def near(value, expected):
    assert expected - 0.02 <= value <= expected + 0.02


def test_entry_point_picks_the_manager():
    base = {'concurrency': 2, 'matrix': {'feature': {'values': [1, 2, 3]}}}
    assert isinstance(get_search_algorithm_manager(HPTuningConfig.from_dict(base)), GridSearchManager)
    assert isinstance(get_search_algorithm_manager(
        HPTuningConfig.from_dict(dict(base, random_search={'n_experiments': 3}))), RandomSearchManager)
    assert isinstance(get_search_algorithm_manager(_hyperband(10, {'name': 'steps', 'type': 'float'}, MATRIX3)),
                      HyperbandSearchManager)
    bo = dict(base, matrix={'x': {'uniform': [0, 1]}},
              bo={'n_iterations': 2, 'n_initial_trials': 2, 'metric': {'name': 'loss', 'optimization': 'minimize'},
                  'utility_function': {'acquisition_function': 'ucb', 'kappa': 1.0,
                                       'gaussian_process': {'kernel': 'matern'}}})
    assert isinstance(get_search_algorithm_manager(HPTuningConfig.from_dict(bo)), BOSearchManager)
    assert get_search_algorithm_manager(None) is None
    with pytest.raises(ValueError):
        RandomSearchManager(HPTuningConfig.from_dict(base))        # config of another algorithm


def test_grid_search_suggestions():
    one = HPTuningConfig.from_dict({'concurrency': 2, 'grid_search': {'n_experiments': 10},
                                    'matrix': {'feature': {'values': [1, 2, 3]}}})
    assert GridSearchManager(one).get_suggestions() == [{'feature': 1}, {'feature': 2}, {'feature': 3}]
    three = HPTuningConfig.from_dict({'concurrency': 2, 'grid_search': {'n_experiments': 10}, 'matrix': MATRIX3})
    assert len(GridSearchManager(three).get_suggestions()) == 10          # 3 * 5 * 4 = 60, cut to 10
    full = HPTuningConfig.from_dict({'concurrency': 2, 'matrix': MATRIX3})
    assert len(GridSearchManager(full).get_suggestions()) == 60
    two = HPTuningConfig.from_dict({'concurrency': 2, 'grid_search': {'n_experiments': 10},
                                    'matrix': {'feature1': {'values': [1, 2, 3]}, 'feature2': {'logspace': '0.01:0.1:5'}}})
    with patch.object(MatrixConfig, 'to_numpy', return_value=[1]) as to_numpy:
        GridSearchManager(two).get_suggestions()
    assert to_numpy.call_count == 2


def test_random_search_suggestions():
    def count(matrix, n=10):
        cfg = HPTuningConfig.from_dict({'concurrency': 2, 'random_search': {'n_experiments': n}, 'matrix': matrix})
        return RandomSearchManager(cfg).get_suggestions()
    assert len(count({'feature1': {'values': [1, 2]}, 'feature3': {'range': [1, 3, 1]}})) == 4   # space of 4
    assert len(count(MATRIX3)) == 10
    mixed = {'feature1': {'pvalues': [(1, 0.3), (2, 0.3), (3, 0.3)]}, 'feature2': {'uniform': [0, 1]},
             'feature3': {'qlognormal': [0, 0.5, 0.51]}}
    got = count(mixed)
    assert len(got) == 10 and all(set(s) == set(mixed) for s in got)
    cfg = HPTuningConfig.from_dict({'concurrency': 2, 'random_search': {'n_experiments': 1},
                                    'matrix': dict(mixed, feature4={'range': [1, 5, 1]})})
    with patch.object(MatrixConfig, 'sample', return_value=1) as sample:
        RandomSearchManager(cfg).get_suggestions()
    assert sample.call_count == 4
    seeded = HPTuningConfig.from_dict({'seed': 7, 'concurrency': 2, 'random_search': {'n_experiments': 5},
                                       'matrix': MATRIX3})
    assert RandomSearchManager(seeded).get_suggestions() == RandomSearchManager(seeded).get_suggestions()


def test_hyperband_geometry(manager1, manager2):
    assert (manager1.max_iter, manager1.eta, manager1.s_max, manager1.B) == (10, 3, 2, 30)
    assert (manager2.max_iter, manager2.eta, manager2.s_max, manager2.B) == (81, 3, 4, 405)
    assert [manager1.get_bracket(iteration=i) for i in range(3)] == [2, 1, 0]
    assert [manager2.get_bracket(iteration=i) for i in range(5)] == [4, 3, 2, 1, 0]
    assert [manager1.get_n_configs(bracket=b) for b in (2, 1, 0)] == [9, 5, 3]
    assert [manager2.get_n_configs(bracket=b) for b in (4, 3, 2, 1, 0)] == [81, 34, 15, 8, 5]
    for b, r in ((2, 1.11), (1, 3.33), (0, 10)):
        near(manager1.get_resources(bracket=b), r)
        near(manager1.get_resources_for_iteration(iteration=2 - b), r)
    for b, r in ((4, 1), (3, 3), (2, 9), (1, 27), (0, 81)):
        near(manager2.get_resources(bracket=b), r)
        near(manager2.get_resources_for_iteration(iteration=4 - b), r)


def test_hyperband_halving_counts(manager1, manager2):
    keep = manager1.get_n_config_to_keep
    assert [keep(n_suggestions=9, bracket_iteration=j) for j in range(3)] == [3, 1, 0]
    assert [keep(n_suggestions=5, bracket_iteration=j) for j in range(2)] == [1, 0]
    assert keep(n_suggestions=3, bracket_iteration=0) == 1
    per = manager1.get_n_config_to_keep_for_iteration
    assert [per(iteration=0, bracket_iteration=j) for j in range(3)] == [3, 1, 0]
    assert [per(iteration=1, bracket_iteration=j) for j in range(2)] == [1, 0]
    assert per(iteration=2, bracket_iteration=0) == 1
    keep2, per2 = manager2.get_n_config_to_keep, manager2.get_n_config_to_keep_for_iteration
    assert [keep2(n_suggestions=81, bracket_iteration=j) for j in range(5)] == [27, 9, 3, 1, 0]
    assert [per2(iteration=0, bracket_iteration=j) for j in range(5)] == [27, 9, 3, 1, 0]
    assert [keep2(n_suggestions=34, bracket_iteration=j) for j in range(4)] == [11, 3, 1, 0]   # reference: 27 -> 34 configs
    assert [per2(iteration=1, bracket_iteration=j) for j in range(4)] == [11, 3, 1, 0]
    assert [per2(iteration=2, bracket_iteration=j) for j in range(3)] == [5, 1, 0]
    assert [per2(iteration=3, bracket_iteration=j) for j in range(2)] == [2, 0]
    assert per2(iteration=4, bracket_iteration=0) == 1


def test_hyperband_resources_per_round(manager1, manager2):
    for j, r in enumerate((1.11, 3.33, 9.99)):
        near(manager1.get_n_resources(n_resources=1.11, bracket_iteration=j), r)
        near(manager1.get_n_resources_for_iteration(iteration=0, bracket_iteration=j), r)
    for j, r in enumerate((3.33, 9.99)):
        near(manager1.get_n_resources_for_iteration(iteration=1, bracket_iteration=j), r)
    near(manager1.get_n_resources_for_iteration(iteration=2, bracket_iteration=0), 9.99)
    for it in range(5):
        for j in range(5 - it):
            near(manager2.get_n_resources_for_iteration(iteration=it, bracket_iteration=j), 3 ** (it + j))


def test_hyperband_scheduling_decisions(manager1):
    resched = {(0, 0): False, (0, 1): False, (0, 2): True, (0, 3): True, (1, 0): False, (1, 1): True,
               (1, 2): True, (2, 0): False, (2, 1): False, (5, 0): False}
    for (it, j), want in resched.items():
        assert manager1.should_reschedule(iteration=it, bracket_iteration=j) is want
    reduce_ = {(0, 0): True, (0, 1): True, (0, 2): False, (0, 3): False, (1, 0): True, (1, 1): False,
               (1, 2): False, (2, 0): True, (2, 1): False, (5, 0): False}
    for (it, j), want in reduce_.items():
        assert manager1.should_reduce_configs(iteration=it, bracket_iteration=j) is want


def test_hyperband_suggestions(manager1, manager2):
    for bad in (None, 1):
        with pytest.raises(ValueError):
            manager1.get_suggestions(bad)
    cases = [(manager1, 0, 2, 9, 'steps', 9.99), (manager1, 1, 0, 5, 'steps', 3.33), (manager1, 2, 0, 3, 'steps', 9.99),
             (manager2, 2, 0, 15, 'size', 9), (manager2, 4, 0, 5, 'size', 81)]
    for manager, it, j, count, resource, budget in cases:
        record = get_iteration_config('hyperband', {'iteration': it, 'bracket_iteration': j})
        assert isinstance(record, HyperbandIterationConfig)
        got = manager.get_suggestions(iteration_config=record)
        assert len(got) == count
        for s in got:
            near(s[resource], budget)
            assert set(manager.hptuning_config.matrix) <= set(s)
    assert isinstance(manager2.get_suggestions(record)[0]['size'], int)


class FakeGroup(object):
    """What the iteration managers touch of the reference's ExperimentGroup model."""
    id = 1

    def __init__(self, hptuning_config):
        self.hptuning_config = hptuning_config
        self.search_algorithm = hptuning_config.search_algorithm
        self.search_manager = get_search_algorithm_manager(hptuning_config)
        self.rows, self.metrics = [], {}

    class Row(object):
        def __init__(self, data):
            self.data, self.saves = data, 0

        def save(self):
            self.saves += 1

    def store(self, group, data):
        assert group is self
        self.rows.append(self.Row(data))
        return self.rows[-1]

    @property
    def iteration(self):
        return self.rows[-1] if self.rows else None

    @property
    def iteration_config(self):
        return get_iteration_config(self.search_algorithm, self.rows[-1].data) if self.rows else None

    def get_experiments_metrics(self, experiment_ids, metric):
        return [[i, self.metrics.get(i)] for i in reversed(experiment_ids or [])]


def test_iteration_manager_factory():
    grid = FakeGroup(HPTuningConfig.from_dict({'concurrency': 2, 'matrix': {'f': {'values': [1, 2]}}}))
    assert get_search_iteration_manager(grid) is None
    hb = FakeGroup(_hyperband(10, {'name': 'steps', 'type': 'float'}, MATRIX3))
    assert isinstance(get_search_iteration_manager(hb, store=hb.store), HyperbandIterationManager)


def test_hyperband_iteration_manager_rounds():
    group = FakeGroup(_hyperband(10, {'name': 'steps', 'type': 'float'}, MATRIX3))
    manager = HyperbandIterationManager(group, store=group.store)
    manager.update_iteration()                       # nothing to update yet: a warning, no row
    assert group.rows == []
    row = manager.create_iteration()
    assert row.data == {'iteration': 0, 'bracket_iteration': 0, 'experiment_ids': None, 'experiments_metrics': None}
    manager.add_iteration_experiments(experiment_ids=[11, 12, 13, 14])
    assert group.iteration.data['experiment_ids'] == [11, 12, 13, 14] and group.iteration.saves == 1
    manager.update_iteration()
    assert group.iteration.data['experiments_metrics'] == []
    assert manager.get_reduced_configs() == []
    group.metrics = {11: 0.9, 12: 0.2, 13: 0.5, 14: 0.7}
    manager.update_iteration()
    assert group.iteration.data['experiments_metrics'] == [[14, 0.7], [13, 0.5], [12, 0.2], [11, 0.9]]
    assert manager.get_reduced_configs() == [12, 13, 14]          # minimise: three smallest of bracket 2, round 0
    # rounds: (0,0) -> (0,1) -> (0,2) -> bracket 1 (1,0) -> (1,1) -> bracket 0 (2,0) -> stop
    seen = []
    for _ in range(5):
        seen.append((lambda d: (d['iteration'], d['bracket_iteration']))(manager.create_iteration([1]).data))
    assert seen == [(0, 1), (0, 2), (1, 0), (1, 1), (2, 0)]
    manager.create_iteration()                        # (2, 0) still has a halving to one config
    assert group.iteration.data['iteration'] == 2 and group.iteration.data['bracket_iteration'] == 1
    with pytest.raises(ValueError):
        manager.create_iteration()


def test_bo_iteration_manager_carries_history():
    cfg = HPTuningConfig.from_dict({
        'concurrency': 2, 'matrix': {'x': {'uniform': [0, 1]}},
        'bo': {'n_iterations': 3, 'n_initial_trials': 2, 'metric': {'name': 'loss', 'optimization': 'minimize'},
               'utility_function': {'acquisition_function': 'ucb', 'kappa': 1.0,
                                    'gaussian_process': {'kernel': 'matern'}}}})
    group = FakeGroup(cfg)
    manager = BOIterationManager(group, store=group.store)
    assert manager.get_metric_name() == 'loss'
    manager.update_iteration()
    assert group.rows == []
    ids1, cfg1 = [1, 2], [[1, {'i': 0}], [2, {'i': 1}]]
    row = manager.create_iteration(experiment_ids=ids1, experiments_configs=cfg1)
    assert row.data == {'iteration': 0, 'old_experiment_ids': None, 'old_experiments_configs': None,
                        'old_experiments_metrics': None, 'experiment_ids': ids1, 'experiments_configs': cfg1,
                        'experiments_metrics': None}
    group.metrics = {1: 0.8, 2: 0.8}
    manager.update_iteration()
    met1 = [[2, 0.8], [1, 0.8]]
    assert group.iteration.data['experiments_metrics'] == met1
    ids2, cfg2 = [3, 4], [[3, {'i': 0}], [4, {'i': 1}]]
    row = manager.create_iteration(experiment_ids=ids2, experiments_configs=cfg2)
    assert row.data == {'iteration': 1, 'old_experiment_ids': ids1, 'old_experiments_configs': cfg1,
                        'old_experiments_metrics': met1, 'experiment_ids': ids2, 'experiments_configs': cfg2,
                        'experiments_metrics': None}
    group.metrics.update({3: 0.9, 4: None})          # experiment 4 has not reported: dropped
    manager.update_iteration()
    assert group.iteration.data['experiments_metrics'] == [[3, 0.9]]
    row = manager.create_iteration(experiment_ids=[5], experiments_configs=[[5, {'i': 0}]])
    assert row.data['iteration'] == 2
    assert row.data['old_experiment_ids'] == ids1 + ids2
    assert row.data['old_experiments_configs'] == cfg1 + cfg2
    assert row.data['old_experiments_metrics'] == met1 + [[3, 0.9]]
