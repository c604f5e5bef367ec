"""Random search (hpsearch/search_managers/random.py:6-21): ``random_search.n_experiments`` distinct
draws from the matrix, seeded by ``hptuning.seed``.  Host-only."""
from polyaxon_b200.hpsearch.search_managers.base import BaseSearchAlgorithmManager
from polyaxon_b200.hpsearch.search_managers.utils import get_random_suggestions
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


class RandomSearchManager(BaseSearchAlgorithmManager):
    NAME = SearchAlgorithms.RANDOM

    def get_suggestions(self, iteration_config=None):
        cfg = self.hptuning_config
        return get_random_suggestions(matrix=cfg.matrix, n_suggestions=cfg.random_search.n_experiments,
                                      seed=cfg.seed)
