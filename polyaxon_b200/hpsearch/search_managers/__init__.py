from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.manager import BOSearchManager
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


def get_search_algorithm_manager(hptuning_config):
    """hpsearch/search_managers/__init__.py:8-21.  Only the Bayesian-optimisation branch lives in
    this library; grid / random / hyperband stay with the reference (no numerical kernel there)."""
    if not hptuning_config:
        return None
    if SearchAlgorithms.is_bo(hptuning_config.search_algorithm):
        return BOSearchManager(hptuning_config=hptuning_config)
    return None
