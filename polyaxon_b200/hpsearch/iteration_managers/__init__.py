from polyaxon_b200.hpsearch.iteration_managers.bayesian_optimization import BOIterationManager
from polyaxon_b200.hpsearch.iteration_managers.hyperband import HyperbandIterationManager
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


def get_search_iteration_manager(experiment_group, store=None):
    """hpsearch/iteration_managers/__init__.py:6-12: grid and random search have no rounds."""
    if SearchAlgorithms.is_hyperband(experiment_group.search_algorithm):
        return HyperbandIterationManager(experiment_group=experiment_group, store=store)
    if SearchAlgorithms.is_bo(experiment_group.search_algorithm):
        return BOIterationManager(experiment_group=experiment_group, store=store)
    return None
