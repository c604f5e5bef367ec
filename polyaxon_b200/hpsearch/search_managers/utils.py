"""Random-number contract and the initial random trials (hpsearch/search_managers/utils.py)."""
import copy
import uuid
from functools import reduce
from operator import mul

import numpy as np


class Suggestion(object):
    """One hyper-parameter assignment; equality / hash by sorted ``key:value`` text (utils.py:9-34)."""

    def __init__(self, params):
        self.params = params

    def __repr__(self):
        return ','.join('{}:{}'.format(k, v) for k, v in sorted(self.params.items()))

    def __eq__(self, other):
        return (self.params.keys() == other.params.keys()
                and all(v == other.params[k] for k, v in self.params.items()))

    def __hash__(self):
        return hash(repr(self))

    def uuid(self):
        return uuid.uuid5(uuid.NAMESPACE_DNS, repr(self))


def get_random_generator(seed=None):
    """utils.py:37-38: a falsy seed (None or 0) means the global numpy generator."""
    return np.random.RandomState(seed) if seed else np.random


def get_random_suggestions(matrix, n_suggestions, suggestion_params=None, seed=None):
    """utils.py:41-64: distinct random assignments; capped by the size of an all-discrete space."""
    if not n_suggestions:
        raise ValueError('This search algorithm requires `n_experiments`.')
    base = suggestion_params or {}
    rng = get_random_generator(seed=seed)
    if not any(v.is_continuous for v in matrix.values()):
        space = reduce(mul, [len(v.to_numpy()) for v in matrix.values()])
        n_suggestions = min(n_suggestions, space)
    chosen = []
    while len(chosen) < n_suggestions:
        params = copy.deepcopy(base)
        params.update({k: v.sample(rand_generator=rng) for k, v in matrix.items()})
        candidate = Suggestion(params=params)
        if candidate not in chosen:
            chosen.append(candidate)
    return [s.params for s in chosen]
