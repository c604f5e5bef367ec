"""Hyperband (hpsearch/search_managers/hyperband.py:9-147).  Host-only bookkeeping, no numerics.

With R = max_iter and eta the elimination rate: s_max = floor(log_eta R), budget per bracket
B = (s_max + 1) R.  Group iteration i runs bracket s = s_max - i: it starts
n(s) = ceil((B / R) eta^s / (s + 1)) random configurations on r(s) = R eta^-s resource units, and
bracket iteration j keeps floor(n eta^-j / eta) of them, each on r(s) eta^j units.
"""
import math

from polyaxon_b200.hpsearch.schemas.hyperband import HyperbandIterationConfig
from polyaxon_b200.hpsearch.search_managers.base import BaseSearchAlgorithmManager
from polyaxon_b200.hpsearch.search_managers.utils import get_random_suggestions
from polyaxon_b200.schemas.hptuning import SearchAlgorithms


class HyperbandSearchManager(BaseSearchAlgorithmManager):
    NAME = SearchAlgorithms.HYPERBAND

    def __init__(self, hptuning_config):
        super().__init__(hptuning_config=hptuning_config)
        settings = hptuning_config.hyperband
        self.max_iter = settings.max_iter
        self.eta = settings.eta
        self.s_max = int(math.log(self.max_iter) / math.log(self.eta))
        self.B = (self.s_max + 1) * self.max_iter

    # --- bracket geometry -------------------------------------------------------------------
    def get_bracket(self, iteration):
        return self.s_max - iteration

    def get_n_configs(self, bracket):
        return int(math.ceil((self.B / self.max_iter) * (self.eta ** bracket) / (bracket + 1)))

    def get_resources(self, bracket):
        return self.max_iter * (self.eta ** (-bracket))

    def get_resources_for_iteration(self, iteration):
        return self.get_resources(bracket=self.get_bracket(iteration=iteration))

    def get_n_config_to_keep(self, n_suggestions, bracket_iteration):
        survivors = n_suggestions * (self.eta ** -bracket_iteration)
        return int(survivors / self.eta)

    def get_n_config_to_keep_for_iteration(self, iteration, bracket_iteration):
        bracket = self.get_bracket(iteration=iteration)
        if bracket_iteration == bracket + 1:       # the bracket's last halving is done
            return 0
        return self.get_n_config_to_keep(n_suggestions=self.get_n_configs(bracket=bracket),
                                         bracket_iteration=bracket_iteration)

    def get_n_resources(self, n_resources, bracket_iteration):
        return n_resources * self.eta ** bracket_iteration

    def get_n_resources_for_iteration(self, iteration, bracket_iteration):
        return self.get_n_resources(n_resources=self.get_resources_for_iteration(iteration=iteration),
                                    bracket_iteration=bracket_iteration)

    # --- scheduling ------------------------------------------------------------------------
    def get_suggestions(self, iteration_config=None):
        if not iteration_config or not isinstance(iteration_config, HyperbandIterationConfig):
            raise ValueError('Hyperband get suggestions requires an iteration.')
        settings = self.hptuning_config.hyperband
        bracket = self.get_bracket(iteration=iteration_config.iteration)
        budget = self.get_n_resources_for_iteration(iteration=iteration_config.iteration,
                                                    bracket_iteration=iteration_config.bracket_iteration)
        return get_random_suggestions(
            matrix=self.hptuning_config.matrix,
            n_suggestions=self.get_n_configs(bracket=bracket),
            suggestion_params={settings.resource.name: settings.resource.cast_value(budget)},
            seed=self.hptuning_config.seed)

    def should_reschedule(self, iteration, bracket_iteration):
        """A new bracket may start once this one has run all its halvings and brackets remain."""
        if bracket_iteration < self.get_bracket(iteration=iteration):
            return False
        return self.get_bracket(iteration=iteration + 1) >= 0

    def should_reduce_configs(self, iteration, bracket_iteration):
        return self.get_n_config_to_keep_for_iteration(iteration=iteration,
                                                       bracket_iteration=bracket_iteration) > 0
