from polyaxon_b200.hpsearch.search_managers.base import BaseSearchAlgorithmManager
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
from polyaxon_b200.hpsearch.search_managers.utils import get_random_suggestions
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


class BOSearchManager(BaseSearchAlgorithmManager):
    """Entry point of the suggestion path (``.../bayesian_optimization/manager.py:7-37``)."""

    NAME = SearchAlgorithms.BO

    def __init__(self, hptuning_config):
        super().__init__(hptuning_config=hptuning_config)
        self.n_initial_trials = hptuning_config.bo.n_initial_trials
        self.n_iterations = hptuning_config.bo.n_iterations

    def get_suggestions(self, iteration_config=None):
        if not iteration_config:
            # iteration 0: random initial trials, no GP involved
            return get_random_suggestions(matrix=self.hptuning_config.matrix,
                                          n_suggestions=self.n_initial_trials,
                                          seed=self.hptuning_config.seed)
        # observations = experiments that reported a metric, in metric order
        by_id = dict(iteration_config.combined_experiments_configs)
        metrics = dict(iteration_config.combined_experiments_metrics)
        configs = [by_id[experiment_id] for experiment_id in metrics]
        optimizer = BOOptimizer(hptuning_config=self.hptuning_config)
        optimizer.add_observations(configs=configs, metrics=list(metrics.values()))
        suggestion = optimizer.get_suggestion()
        return [suggestion] if suggestion else None

    def should_reschedule(self, iteration):
        """Whether another iteration has to be scheduled."""
        return iteration < self.n_iterations
