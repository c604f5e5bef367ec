"""Grid search (hpsearch/search_managers/grid.py:7-31): the Cartesian product of every declared
value list, in matrix key order, optionally cut to ``grid_search.n_experiments``.  Host-only."""
import itertools

from polyaxon_b200.hpsearch.search_managers.base import BaseSearchAlgorithmManager
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


class GridSearchManager(BaseSearchAlgorithmManager):
    NAME = SearchAlgorithms.GRID

    def get_suggestions(self, iteration_config=None):
        matrix = self.hptuning_config.matrix
        names = list(matrix)
        axes = [matrix[name].to_numpy() for name in names]
        grid = [dict(zip(names, point)) for point in itertools.product(*axes)]
        limit = getattr(self.hptuning_config.grid_search, 'n_experiments', None)
        return grid[:limit] if limit else grid
