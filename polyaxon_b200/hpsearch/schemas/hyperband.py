"""Wire format of a Hyperband group iteration (hpsearch/schemas/hyperband.py:6-32): which bracket
(``iteration``) and which halving round inside it (``bracket_iteration``), plus the experiments of
the round and their metrics.  ``to_dict`` dumps every field, None included, as the rows compared in
the reference's tests/test_experiment_groups/test_iteration_managers.py:62-67 show."""

_FIELDS = ('iteration', 'bracket_iteration', 'experiment_ids', 'experiments_metrics')


class HyperbandIterationConfig(object):
    def __init__(self, iteration, bracket_iteration, experiment_ids=None, experiments_metrics=None):
        self.iteration = iteration
        self.bracket_iteration = bracket_iteration
        self.experiment_ids = experiment_ids
        self.experiments_metrics = experiments_metrics

    @classmethod
    def from_dict(cls, value):
        unknown = set(value) - set(_FIELDS)
        if unknown:
            raise ValueError('Unknown iteration fields: {}'.format(sorted(unknown)))
        for key in ('iteration', 'bracket_iteration'):
            if key not in value:
                raise ValueError('`{}` is required'.format(key))
        data = dict(value)
        data['iteration'] = int(data['iteration'])
        data['bracket_iteration'] = int(data['bracket_iteration'])
        if data.get('experiment_ids') is not None:
            data['experiment_ids'] = [int(v) for v in data['experiment_ids']]
        if data.get('experiments_metrics') is not None:
            pairs = [list(p) for p in data['experiments_metrics']]
            if any(len(p) != 2 for p in pairs):
                raise ValueError('`experiments_metrics` must hold [experiment_id, value] pairs')
            data['experiments_metrics'] = pairs
        return cls(**data)

    def to_dict(self):
        return {k: getattr(self, k) for k in _FIELDS}
