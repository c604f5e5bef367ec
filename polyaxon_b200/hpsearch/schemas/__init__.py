from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


def get_iteration_config(search_algorithm, iteration=None):
    """hpsearch/schemas/__init__.py:6-15, Bayesian-optimisation branch (hyperband is out of scope)."""
    if SearchAlgorithms.is_bo(search_algorithm):
        if not iteration:
            raise ValueError('No iteration was provided')
        return BOIterationConfig.from_dict(iteration)
    return None
