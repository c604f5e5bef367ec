"""Configuration objects read by the Bayesian-optimisation suggestion path.

The reference gets these from the un-vendored ``polyaxon_schemas==0.2.8`` package
(re-exported at ``polyaxon/schemas/hptuning.py:1-19``).  When that package is
importable it is used as is; otherwise the classes below provide exactly the
attribute surface the hot path reads:

* ``HPTuningConfig.{seed, matrix, search_algorithm, bo, concurrency}``
* ``BOConfig.{n_initial_trials, n_iterations, metric, utility_function}``
* ``UtilityFunctionConfig.{acquisition_function, kappa, eps, n_warmup, n_iter, gaussian_process}``
* ``GaussianProcessConfig.{kernel, length_scale, nu, n_restarts_optimizer}``
* ``MatrixConfig.{is_categorical, is_discrete, is_uniform, is_continuous, min, max, to_numpy(), sample()}``
* enums ``AcquisitionFunctions``, ``GaussianProcessesKernels``, ``Optimization``, ``SearchAlgorithms``

Semantics of every ``MatrixConfig`` property are pinned by the reference's own
tests (``tests/test_experiment_groups/test_search_managers.py:715-879``); those
tests are restated in ``tests/test_host_space.py``.

This is configuration plumbing only: no arithmetic of the hot path lives here.
"""
import copy

import numpy as np

try:  # pragma: no cover - the real package is absent from this image
    from polyaxon_schemas.hptuning import (  # noqa: F401
        BOConfig,
        GaussianProcessConfig,
        HPTuningConfig,
        SearchMetricConfig,
        UtilityFunctionConfig,
    )
    from polyaxon_schemas.matrix import MatrixConfig  # noqa: F401
    from polyaxon_schemas.utils import (  # noqa: F401
        AcquisitionFunctions,
        GaussianProcessesKernels,
        Optimization,
        SearchAlgorithms,
    )
    HAVE_POLYAXON_SCHEMAS = True
except ImportError:
    HAVE_POLYAXON_SCHEMAS = False


class ConfigError(ValueError):
    """Raised for malformed hptuning sections (marshmallow's ValidationError in the reference)."""


if not HAVE_POLYAXON_SCHEMAS:

    class _Enum(object):
        VALUES = ()

        @classmethod
        def check(cls, value, what):
            if value not in cls.VALUES:
                raise ConfigError('`{}` is not a valid {}; expected one of {}'.format(
                    value, what, sorted(cls.VALUES)))
            return value

    class SearchAlgorithms(_Enum):
        GRID = 'grid'
        RANDOM = 'random'
        HYPERBAND = 'hyperband'
        BO = 'bo'
        VALUES = {GRID, RANDOM, HYPERBAND, BO}

        @classmethod
        def is_grid(cls, value):
            return value == cls.GRID

        @classmethod
        def is_random(cls, value):
            return value == cls.RANDOM

        @classmethod
        def is_hyperband(cls, value):
            return value == cls.HYPERBAND

        @classmethod
        def is_bo(cls, value):
            return value == cls.BO

    class Optimization(_Enum):
        MAXIMIZE = 'maximize'
        MINIMIZE = 'minimize'
        VALUES = {MAXIMIZE, MINIMIZE}

        @classmethod
        def maximize(cls, value):
            return value == cls.MAXIMIZE

        @classmethod
        def minimize(cls, value):
            return value == cls.MINIMIZE

    class AcquisitionFunctions(_Enum):
        UCB = 'ucb'
        EI = 'ei'
        POI = 'poi'
        VALUES = {UCB, EI, POI}

        @classmethod
        def is_ucb(cls, value):
            return value == cls.UCB

        @classmethod
        def is_ei(cls, value):
            return value == cls.EI

        @classmethod
        def is_poi(cls, value):
            return value == cls.POI

    class GaussianProcessesKernels(_Enum):
        RBF = 'rbf'
        MATERN = 'matern'
        VALUES = {RBF, MATERN}

        @classmethod
        def is_rbf(cls, value):
            return value == cls.RBF

        @classmethod
        def is_mattern(cls, value):
            return value == cls.MATERN

    _NUMERIC = (int, float, complex, np.integer, np.floating)

    class MatrixConfig(object):
        """One hyper-parameter declaration: a dict with exactly one option key."""

        RANGES = ('range', 'linspace', 'logspace', 'geomspace')
        ENUMS = ('values', 'pvalues')
        DISTRIBUTIONS = ('uniform', 'quniform', 'loguniform', 'qloguniform',
                         'normal', 'qnormal', 'lognormal', 'qlognormal')
        CONTINUOUS = DISTRIBUTIONS
        _RANGE_ARGS = {'range': ('start', 'stop', 'step'),
                       'linspace': ('start', 'stop', 'num'),
                       'logspace': ('start', 'stop', 'num'),
                       'geomspace': ('start', 'stop', 'num')}
        _DIST_ARGS = {'uniform': ('low', 'high'), 'quniform': ('low', 'high', 'q'),
                      'loguniform': ('low', 'high'), 'qloguniform': ('low', 'high', 'q'),
                      'normal': ('loc', 'scale'), 'qnormal': ('loc', 'scale', 'q'),
                      'lognormal': ('loc', 'scale'), 'qlognormal': ('loc', 'scale', 'q')}
        _RANGE_FUNCS = {'range': np.arange, 'linspace': np.linspace,
                        'logspace': np.logspace, 'geomspace': np.geomspace}

        def __init__(self, **options):
            options = {k: v for k, v in options.items() if v is not None}
            if len(options) != 1:
                raise ConfigError('Matrix entries take exactly one option, got {}'.format(
                    sorted(options)))
            (self.key, value), = options.items()
            if self.key in self._RANGE_ARGS:
                value = self._named(value, self._RANGE_ARGS[self.key])
            elif self.key in self._DIST_ARGS:
                value = self._named(value, self._DIST_ARGS[self.key])
            elif self.key == 'pvalues':
                value = [tuple(v) for v in value]
                total = sum(v[1] for v in value)
                if abs(total - 1.0) > 1e-9 and total > 1.0:
                    raise ConfigError('pvalues probabilities must sum to at most 1.')
            elif self.key != 'values':
                raise ConfigError('Unknown matrix option `{}`'.format(self.key))
            self.value = value

        @staticmethod
        def _named(value, names):
            if isinstance(value, dict):
                missing = [n for n in names if n not in value]
                if missing:
                    raise ConfigError('Matrix option is missing {}'.format(missing))
                return {n: value[n] for n in names}
            if isinstance(value, str):
                value = [float(v) for v in value.split(':')]
            value = list(value)
            if len(value) != len(names):
                raise ConfigError('Matrix option expects {} got {}'.format(names, value))
            return dict(zip(names, value))

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {self.key: copy.deepcopy(self.value)}

        @property
        def is_distribution(self):
            return self.key in self.DISTRIBUTIONS

        @property
        def is_continuous(self):
            return self.key in self.CONTINUOUS

        @property
        def is_discrete(self):
            return not self.is_continuous

        @property
        def is_range(self):
            return self.key in self.RANGES

        @property
        def is_uniform(self):
            return self.key == 'uniform'

        @property
        def is_categorical(self):
            if self.key == 'values':
                return any(not isinstance(v, _NUMERIC) or isinstance(v, bool)
                           for v in self.value)
            if self.key == 'pvalues':
                return any(not isinstance(v[0], _NUMERIC) or isinstance(v[0], bool)
                           for v in self.value)
            return False

        @property
        def min(self):
            if self.is_categorical:
                return None
            if self.is_range:
                return self.value['start']
            if self.is_uniform:
                return self.value['low']
            if self.is_discrete:
                return min(self.to_numpy())
            return None

        @property
        def max(self):
            if self.is_categorical:
                return None
            if self.is_range:
                return self.value['stop']
            if self.is_uniform:
                return self.value['high']
            if self.is_discrete:
                return max(self.to_numpy())
            return None

        @property
        def length(self):
            if self.is_distribution:
                raise ConfigError('Distributions have no length.')
            return len(self.to_numpy())

        def to_numpy(self):
            if self.is_distribution:
                raise ConfigError(
                    'Distribution should not call `to_numpy`, instead it should call `sample`.')
            if self.is_range:
                return self._RANGE_FUNCS[self.key](**self.value)
            if self.key == 'pvalues':
                return [v[0] for v in self.value]
            return self.value

        def sample(self, size=None, rand_generator=None):
            size = None if size == 1 else size
            rng = rand_generator or np.random
            key, value = self.key, copy.deepcopy(self.value)
            if key == 'values' or self.is_range:
                return rng.choice(self.to_numpy(), size=size)
            if key == 'pvalues':
                # the reference's own test matrix has weights summing to 0.9
                # (test_search_managers.py:138); they are used as relative weights here
                weights = np.asarray([v[1] for v in value], dtype=float)
                return rng.choice([v[0] for v in value], size=size, p=weights / weights.sum())
            if key == 'uniform':
                return rng.uniform(value['low'], value['high'], size=size)
            if key == 'quniform':
                draw = rng.uniform(value['low'], value['high'], size=size)
                return np.round(draw / value['q']) * value['q']
            if key == 'loguniform':
                return np.exp(rng.uniform(value['low'], value['high'], size=size))
            if key == 'qloguniform':
                draw = np.exp(rng.uniform(value['low'], value['high'], size=size))
                return np.round(draw / value['q']) * value['q']
            if key == 'normal':
                return rng.normal(value['loc'], value['scale'], size=size)
            if key == 'qnormal':
                draw = rng.normal(value['loc'], value['scale'], size=size)
                return np.round(draw / value['q']) * value['q']
            if key == 'lognormal':
                return rng.lognormal(value['loc'], value['scale'], size=size)
            draw = rng.lognormal(value['loc'], value['scale'], size=size)
            return np.round(draw / value['q']) * value['q']

    class SearchMetricConfig(object):
        def __init__(self, name, optimization=Optimization.MAXIMIZE):
            self.name = name
            self.optimization = Optimization.check(optimization, 'optimization')

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {'name': self.name, 'optimization': self.optimization}

    class GaussianProcessConfig(object):
        def __init__(self, kernel=GaussianProcessesKernels.MATERN, length_scale=1.0, nu=1.5,
                     n_restarts_optimizer=0):
            self.kernel = GaussianProcessesKernels.check(kernel, 'gaussian process kernel')
            self.length_scale = length_scale
            self.nu = nu
            self.n_restarts_optimizer = n_restarts_optimizer

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {'kernel': self.kernel, 'length_scale': self.length_scale, 'nu': self.nu,
                    'n_restarts_optimizer': self.n_restarts_optimizer}

    class UtilityFunctionConfig(object):
        def __init__(self, acquisition_function=AcquisitionFunctions.UCB, gaussian_process=None,
                     kappa=None, eps=None, n_warmup=None, n_iter=None):
            self.acquisition_function = AcquisitionFunctions.check(
                acquisition_function, 'acquisition function')
            if AcquisitionFunctions.is_ucb(acquisition_function) and kappa is None:
                raise ConfigError('the acquisition function `ucb` requires a parameter `kappa`')
            if not AcquisitionFunctions.is_ucb(acquisition_function) and eps is None:
                raise ConfigError('the acquisition function `{}` requires a parameter `eps`'.format(
                    acquisition_function))
            if isinstance(gaussian_process, dict):
                gaussian_process = GaussianProcessConfig.from_dict(gaussian_process)
            self.gaussian_process = gaussian_process or GaussianProcessConfig()
            self.kappa = kappa
            self.eps = eps
            self.n_warmup = n_warmup
            self.n_iter = n_iter

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            out = {'acquisition_function': self.acquisition_function,
                   'gaussian_process': self.gaussian_process.to_dict()}
            for key in ('kappa', 'eps', 'n_warmup', 'n_iter'):
                if getattr(self, key) is not None:
                    out[key] = getattr(self, key)
            return out

    class BOConfig(object):
        def __init__(self, metric, utility_function=None, n_initial_trials=None, n_iterations=None):
            if isinstance(metric, dict):
                metric = SearchMetricConfig.from_dict(metric)
            if isinstance(utility_function, dict):
                utility_function = UtilityFunctionConfig.from_dict(utility_function)
            self.metric = metric
            self.utility_function = utility_function
            self.n_initial_trials = n_initial_trials
            self.n_iterations = n_iterations

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {'metric': self.metric.to_dict(),
                    'utility_function': self.utility_function.to_dict(),
                    'n_initial_trials': self.n_initial_trials,
                    'n_iterations': self.n_iterations}

    class GridSearchConfig(object):
        def __init__(self, n_experiments=None):
            self.n_experiments = n_experiments

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {} if self.n_experiments is None else {'n_experiments': self.n_experiments}

    class RandomSearchConfig(GridSearchConfig):
        def __init__(self, n_experiments):
            super().__init__(n_experiments=n_experiments)

    class ResourceConfig(object):
        """Hyperband's resource: the declaration that receives the budget, as int or float."""
        TYPES = {'int': int, 'float': float}

        def __init__(self, name, type='float'):    # noqa: A002 - the section's key is `type`
            if type not in self.TYPES:
                raise ConfigError('resource type must be `int` or `float`')
            self.name = name
            self.type = type

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {'name': self.name, 'type': self.type}

        def cast_value(self, value):
            return self.TYPES[self.type](value)

    class HyperbandConfig(object):
        def __init__(self, max_iter, eta, resource, metric, resume=False):
            self.max_iter = max_iter
            self.eta = eta
            self.resource = ResourceConfig.from_dict(resource) if isinstance(resource, dict) else resource
            self.metric = SearchMetricConfig.from_dict(metric) if isinstance(metric, dict) else metric
            self.resume = resume

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            return {'max_iter': self.max_iter, 'eta': self.eta, 'resource': self.resource.to_dict(),
                    'metric': self.metric.to_dict(), 'resume': self.resume}

    class HPTuningConfig(object):
        """The ``hptuning`` section: the matrix plus the settings of one search algorithm."""

        def __init__(self, matrix=None, seed=None, concurrency=1, bo=None, early_stopping=None,
                     grid_search=None, random_search=None, hyperband=None):
            self.seed = seed
            self.concurrency = concurrency
            self.early_stopping = early_stopping
            self.matrix = {
                key: value if isinstance(value, MatrixConfig) else MatrixConfig.from_dict(value)
                for key, value in (matrix or {}).items()}
            if isinstance(bo, dict):
                bo = BOConfig.from_dict(bo)
            self.bo = bo
            self.grid_search = (GridSearchConfig.from_dict(grid_search) if isinstance(grid_search, dict)
                                else grid_search)
            self.random_search = (RandomSearchConfig.from_dict(random_search)
                                  if isinstance(random_search, dict) else random_search)
            self.hyperband = HyperbandConfig.from_dict(hyperband) if isinstance(hyperband, dict) else hyperband
            if sum(x is not None for x in (self.bo, self.grid_search, self.random_search, self.hyperband)) > 1:
                raise ConfigError('Only one search algorithm can be configured.')

        @classmethod
        def from_dict(cls, value):
            return cls(**value)

        def to_dict(self):
            out = {'concurrency': self.concurrency,
                   'matrix': {k: v.to_dict() for k, v in self.matrix.items()}}
            if self.seed is not None:
                out['seed'] = self.seed
            for key in ('bo', 'grid_search', 'random_search', 'hyperband'):
                if getattr(self, key) is not None:
                    out[key] = getattr(self, key).to_dict()
            return out

        @property
        def search_algorithm(self):
            if self.bo is not None:
                return SearchAlgorithms.BO
            if self.hyperband is not None:
                return SearchAlgorithms.HYPERBAND
            if self.random_search is not None:
                return SearchAlgorithms.RANDOM
            return SearchAlgorithms.GRID
