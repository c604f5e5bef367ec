"""Wire format of the Bayesian-optimisation state between iterations.

Mirrors ``polyaxon/hpsearch/schemas/bayesian_optimization.py:6-70`` (a marshmallow schema in the
reference; marshmallow is not needed for a flat record, so validation is done by hand): the JSON
stored in ``ExperimentGroupIteration.data``.  Round-trip ``from_dict(cfg).to_dict() == cfg`` is
pinned by the reference's ``tests/test_experiment_groups/test_schemas.py:70-86``.
"""

_FIELDS = ('iteration', 'old_experiment_ids', 'old_experiments_configs', 'old_experiments_metrics',
           'experiment_ids', 'experiments_configs', 'experiments_metrics')
_PAIR_FIELDS = ('old_experiments_configs', 'old_experiments_metrics', 'experiments_configs',
                'experiments_metrics')


class BOIterationConfig(object):
    def __init__(self, iteration, old_experiment_ids=None, old_experiments_metrics=None,
                 old_experiments_configs=None, experiment_ids=None, experiments_metrics=None,
                 experiments_configs=None):
        self.iteration = iteration
        self.old_experiment_ids = old_experiment_ids
        self.old_experiments_metrics = old_experiments_metrics
        self.old_experiments_configs = old_experiments_configs
        self.experiment_ids = experiment_ids
        self.experiments_configs = experiments_configs
        self.experiments_metrics = experiments_metrics

    @classmethod
    def from_dict(cls, value):
        unknown = set(value) - set(_FIELDS)
        if unknown:
            raise ValueError('Unknown iteration fields: {}'.format(sorted(unknown)))
        if 'iteration' not in value:
            raise ValueError('`iteration` is required')
        data = dict(value)
        data['iteration'] = int(data['iteration'])
        for key in ('old_experiment_ids', 'experiment_ids'):
            if data.get(key) is not None:
                data[key] = [int(v) for v in data[key]]
        for key in _PAIR_FIELDS:
            if data.get(key) is not None:
                pairs = [list(p) for p in data[key]]
                if any(len(p) != 2 for p in pairs):
                    raise ValueError('`{}` must hold [experiment_id, value] pairs'.format(key))
                data[key] = pairs
        return cls(**data)

    def to_dict(self):
        # every field is dumped, None included: the stored rows the reference's tests compare against
        # (tests/test_experiment_groups/test_iteration_managers.py:152-160) carry `'old_experiment_ids': None`
        return {k: getattr(self, k) for k in _FIELDS}

    def _combined(self, old, new):
        # the reference extends the stored `old_*` list in place when it is non-empty; keep that
        out = getattr(self, old) or []
        out += getattr(self, new) or []
        return out

    @property
    def combined_experiment_ids(self):
        return self._combined('old_experiment_ids', 'experiment_ids')

    @property
    def combined_experiments_configs(self):
        return self._combined('old_experiments_configs', 'experiments_configs')

    @property
    def combined_experiments_metrics(self):
        return self._combined('old_experiments_metrics', 'experiments_metrics')
