from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import UtilityFunction
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.space import SearchSpace


class BOOptimizer(object):
    """``.../bayesian_optimization/optimizer.py:5-33``: owns the search space and the utility
    function; one suggestion = GP fit on the observations + arg max of the acquisition."""

    def __init__(self, hptuning_config):
        self.hptuning_config = hptuning_config
        bo = hptuning_config.bo
        self.n_initial_trials = bo.n_initial_trials
        self.space = SearchSpace(hptuning_config=hptuning_config)
        self.utility_function = UtilityFunction(config=bo.utility_function, seed=hptuning_config.seed)
        self.n_warmup = bo.utility_function.n_warmup or 5        # effective defaults: optimizer.py:13-14
        self.n_iter = bo.utility_function.n_iter or 10

    def _maximize(self):
        """Find argmax of the acquisition function."""
        if not self.space.is_observations_valid():
            return None
        y_max = self.space.y.max()
        self.utility_function.gaussian_process.fit(self.space.x, self.space.y)
        return self.utility_function.max_compute(y_max=y_max, bounds=self.space.bounds,
                                                 n_warmup=self.n_warmup, n_iter=self.n_iter)

    def add_observations(self, configs, metrics):
        self.space.add_observations(configs=configs, metrics=metrics)

    def get_suggestion(self):
        return self.space.get_suggestion(self._maximize())
