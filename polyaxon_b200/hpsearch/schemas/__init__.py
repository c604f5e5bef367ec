from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
from polyaxon_b200.hpsearch.schemas.hyperband import HyperbandIterationConfig
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


def get_iteration_config(search_algorithm, iteration=None):
    """hpsearch/schemas/__init__.py:6-15: the stored iteration record of the algorithms that have one."""
    if SearchAlgorithms.is_hyperband(search_algorithm) or SearchAlgorithms.is_bo(search_algorithm):
        if not iteration:
            raise ValueError('No iteration was provided')
        cls = HyperbandIterationConfig if SearchAlgorithms.is_hyperband(search_algorithm) else BOIterationConfig
        return cls.from_dict(iteration)
    return None
