"""Iteration bookkeeping shared by the algorithms that run in rounds
(hpsearch/iteration_managers/base.py:6-45).

The reference keeps one ``ExperimentGroupIteration`` row per round and reads / writes it through
the Django model.  Here the group object only has to offer what those methods touch --
``iteration_config``, ``iteration`` (a record with ``.data`` and ``.save()``), ``hptuning_config``,
``search_manager``, ``get_experiments_metrics(experiment_ids=, metric=)`` -- and new rows are made by
``store(experiment_group, data)``; the default store is the reference's model, imported on use.
"""
import logging

_logger = logging.getLogger('polyaxon.hpsearch.iteration_manager')


def _model_store(experiment_group, data):
    from db.models.experiment_groups import ExperimentGroupIteration   # the reference's Django model
    return ExperimentGroupIteration.objects.create(experiment_group=experiment_group, data=data)


class BaseIterationManager(object):
    def __init__(self, experiment_group, store=None):
        self.experiment_group = experiment_group
        self.store = store or _model_store

    def get_metric_name(self):
        return None

    def get_iteration_config(self):
        config = self.experiment_group.iteration_config
        if config is None:
            _logger.warning('Experiment group `%s` attempted to update iteration, but has no iteration',
                            self.experiment_group.id)
        return config

    def _update_config(self, iteration_config):
        record = self.experiment_group.iteration
        record.data = iteration_config.to_dict()
        record.save()

    def update_iteration(self):
        """Write the finished experiments' metrics into the current round (entries without a
        metric are dropped, base.py:36)."""
        config = self.get_iteration_config()
        if not config:
            return
        metrics = self.experiment_group.get_experiments_metrics(experiment_ids=config.experiment_ids,
                                                                metric=self.get_metric_name())
        config.experiments_metrics = [pair for pair in metrics if pair[1] is not None]
        self._update_config(config)

    def add_iteration_experiments(self, experiment_ids):
        config = self.get_iteration_config()
        if not config:
            return
        config.experiment_ids = experiment_ids
        self._update_config(config)


BaseIterationManger = BaseIterationManager      # the reference's spelling (base.py:6)
