from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.manager import BOSearchManager
from polyaxon_b200.hpsearch.search_managers.grid import GridSearchManager
from polyaxon_b200.hpsearch.search_managers.hyperband import HyperbandSearchManager
from polyaxon_b200.hpsearch.search_managers.random import RandomSearchManager
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


def get_search_algorithm_manager(hptuning_config):
    """hpsearch/search_managers/__init__.py:8-21.  Only the Bayesian-optimisation manager touches the
    GPU; grid / random / hyperband are host bookkeeping kept here so the entry point is whole."""
    if not hptuning_config:
        return None
    algorithm = hptuning_config.search_algorithm
    for accepts, manager in ((SearchAlgorithms.is_grid, GridSearchManager),
                             (SearchAlgorithms.is_random, RandomSearchManager),
                             (SearchAlgorithms.is_hyperband, HyperbandSearchManager),
                             (SearchAlgorithms.is_bo, BOSearchManager)):
        if accepts(algorithm):
            return manager(hptuning_config=hptuning_config)
    return None
