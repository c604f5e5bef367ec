"""Encoding between hyper-parameter dicts and the optimiser's vector space.

Host-side string/dict work; mirrors ``SearchSpace`` of the reference
(``.../bayesian_optimization/space.py:9-164``) attribute for attribute: features in sorted key
order, categorical -> one-hot block with bounds (0, 1), discrete -> one column bounded by the
declared min/max, ``uniform`` -> one column bounded by (low, high).  As in the reference, other
distributions get a feature but no bound (space.py:90 tests ``is_uniform`` only).
"""
import logging

import numpy as np

from polyaxon_b200.schemas.hptuning import Optimization

_logger = logging.getLogger('polyaxon.hpsearch.search_managers')


class SearchSpace(object):
    def __init__(self, hptuning_config):
        self.hptuning_config = hptuning_config
        self._dim = 0
        self._bounds = []
        self._features = []
        self._discrete_features = {}
        self._categorical_features = {}
        self._x = []
        self._y = []
        self.set_bounds()

    x = property(lambda self: self._x)
    y = property(lambda self: self._y)
    dim = property(lambda self: self._dim)
    features = property(lambda self: self._features)
    discrete_features = property(lambda self: self._discrete_features)
    categorical_features = property(lambda self: self._categorical_features)
    bounds = property(lambda self: self._bounds)

    def is_observations_valid(self):
        n_x, n_y = len(self.x), len(self.y)
        if n_x != n_y:
            _logger.warning("X and Y observations don't have the same size.")
            return False
        if n_x == 0 or n_y == 0:
            _logger.warning("Space has no observations.")
            return False
        return True

    def set_bounds(self):
        rows = []
        matrix = self.hptuning_config.matrix
        for name in sorted(matrix.keys()):
            spec = matrix[name]
            self._features.append(name)
            if spec.is_categorical:
                choices = spec.to_numpy()
                rows.extend([(0, 1)] * len(choices))
                self._categorical_features[name] = {'values': choices, 'number': len(choices)}
                self._dim += len(choices)
            elif spec.is_discrete:
                rows.append((spec.min, spec.max))
                self._discrete_features[name] = {'values': spec.to_numpy()}
                self._dim += 1
            elif spec.is_uniform:
                rows.append((float(spec.min), float(spec.max)))
                self._dim += 1
        self._bounds = np.asarray(rows)

    def parse_y(self, metrics):
        if not metrics:
            return metrics
        sign = 1.0 if Optimization.maximize(self.hptuning_config.bo.metric.optimization) else -1.0
        return np.array([sign * float(m) for m in metrics])

    def parse_x(self, configs):
        if not configs:
            return configs
        rows = []
        for config in configs:
            row = []
            for name in self._features:
                one_hot = self._categorical_features.get(name)
                if one_hot is not None:
                    row.extend(1 if choice == config[name] else 0 for choice in one_hot['values'])
                else:
                    row.append(config[name])
            rows.append(row)
        return np.array(rows)

    def add_observations(self, configs, metrics):
        self._x = self.parse_x(configs=configs)
        self._y = self.parse_y(metrics=metrics)

    def _nearest_feasible(self, name, value):
        feasible = self._discrete_features[name]['values']
        return feasible[int(np.argmin(np.absolute(np.subtract(feasible, value))))]

    def get_suggestion(self, suggestion):
        """Vector -> {feature: value}: snap discrete columns, arg-max one-hot blocks."""
        if suggestion is None:
            return suggestion
        pos, out = 0, []
        for name in self._features:
            if name in self._discrete_features:
                out.append(self._nearest_feasible(name, suggestion[pos]))
                pos += 1
            elif name in self._categorical_features:
                width = self._categorical_features[name]['number']
                out.append(self._categorical_features[name]['values'][
                    int(np.argmax(suggestion[pos:pos + width]))])
                pos += width
            else:
                out.append(suggestion[pos])
                pos += 1
        return dict(zip(self._features, out))
