"""Rounds of a Bayesian-optimisation group (hpsearch/iteration_managers/bayesian_optimization.py:5-37):
every new round carries all earlier experiments, configs and metrics in its ``old_*`` fields, which is
what ``BOSearchManager.get_suggestions`` reads as its observation set."""
from polyaxon_b200.hpsearch.iteration_managers.base import BaseIterationManager
from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig


class BOIterationManager(BaseIterationManager):
    def get_metric_name(self):
        return self.experiment_group.hptuning_config.bo.metric.name

    def create_iteration(self, experiment_ids, experiments_configs):
        previous = self.experiment_group.iteration_config
        history = {}
        if previous is not None:
            history = dict(old_experiment_ids=previous.combined_experiment_ids,
                           old_experiments_configs=previous.combined_experiments_configs,
                           old_experiments_metrics=previous.combined_experiments_metrics)
        config = BOIterationConfig(iteration=0 if previous is None else previous.iteration + 1,
                                   experiment_ids=experiment_ids, experiments_configs=experiments_configs,
                                   **history)
        return self.store(self.experiment_group, config.to_dict())
