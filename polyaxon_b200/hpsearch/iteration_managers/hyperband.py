"""Rounds of a Hyperband group (hpsearch/iteration_managers/hyperband.py:10-113): either the next
halving inside the current bracket or the first round of the next bracket, and which experiments
survive a halving.  Restarting / resuming the survivors (``reduce_configs`` in the reference,
:74-113) is scheduler work and stays with the caller."""
import logging

from polyaxon_b200.hpsearch.iteration_managers.base import BaseIterationManager
from polyaxon_b200.hpsearch.schemas.hyperband import HyperbandIterationConfig
from polyaxon_b200.schemas.hptuning import Optimization

_logger = logging.getLogger('polyaxon.hpsearch.iteration_manager')


class HyperbandIterationManager(BaseIterationManager):
    def get_metric_name(self):
        return self.experiment_group.hptuning_config.hyperband.metric.name

    def create_iteration(self, experiment_ids=None):
        manager = self.experiment_group.search_manager
        previous = self.experiment_group.iteration_config
        if previous is None:
            position = (0, 0)
        elif manager.should_reschedule(iteration=previous.iteration,
                                       bracket_iteration=previous.bracket_iteration):
            position = (previous.iteration + 1, 0)
        elif manager.should_reduce_configs(iteration=previous.iteration,
                                           bracket_iteration=previous.bracket_iteration):
            position = (previous.iteration, previous.bracket_iteration + 1)
        else:
            raise ValueError('Hyperband create iteration failed for `{}`, '
                             'could not reschedule ot reduce configs'.format(self.experiment_group.id))
        config = HyperbandIterationConfig(iteration=position[0], bracket_iteration=position[1],
                                          experiment_ids=experiment_ids or None)
        return self.store(self.experiment_group, config.to_dict())

    def get_reduced_configs(self):
        """Ids of the experiments that go on to the next halving: the best
        ``get_n_config_to_keep_for_iteration`` by metric (stable order among equals)."""
        config = self.experiment_group.iteration_config
        if config is None:
            _logger.error('Experiment group `%s` attempt to update iteration, but has no iteration',
                          self.experiment_group.id)
            return None
        keep = self.experiment_group.search_manager.get_n_config_to_keep_for_iteration(
            iteration=config.iteration, bracket_iteration=config.bracket_iteration)
        descending = Optimization.maximize(self.experiment_group.hptuning_config.hyperband.metric.optimization)
        ranked = sorted(config.experiments_metrics, key=lambda pair: pair[1], reverse=descending)
        return [pair[0] for pair in ranked[:keep]]
