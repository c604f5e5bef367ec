class BaseSearchAlgorithmManager(object):
    """hpsearch/search_managers/base.py:1-13: refuses a config written for another algorithm."""
    NAME = None

    def __init__(self, hptuning_config):
        self.hptuning_config = hptuning_config
        if self.NAME != hptuning_config.search_algorithm:
            raise ValueError(
                'The current search manger `{}` is not compatible '
                'with the search algorithm `{}` defined in the config.'.format(
                    self.NAME, hptuning_config.search_algorithm))

    def get_suggestions(self, iteration_config=None):
        raise NotImplementedError
