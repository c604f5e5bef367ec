"""cProfile of the drop-in get_suggestion() at config C1 (d=2, n=15, 1e5 candidates, 10 seeds): where the host time goes."""
import cProfile
import pstats
import sys
import time
sys.path.insert(0, '.')
import numpy as np
from oracle.reference_path import synthetic_problem
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
from polyaxon_b200.schemas.hptuning import HPTuningConfig

X, y, _ = synthetic_problem(15, 2)
keys = ['x000', 'x001']
cfg = HPTuningConfig.from_dict({
    'seed': 17, 'concurrency': 1,
    'bo': {'n_iterations': 5, 'n_initial_trials': 5, 'metric': {'name': 'm', 'optimization': 'maximize'},
           'utility_function': {'acquisition_function': 'ucb', 'kappa': 2.576, 'n_warmup': 100000, 'n_iter': 10,
                                'gaussian_process': {'kernel': 'matern', 'length_scale': 1.0, 'nu': 2.5,
                                                     'n_restarts_optimizer': 0}}},
    'matrix': {k: {'uniform': [0, 1]} for k in keys}})
configs = [dict(zip(keys, row.tolist())) for row in X]


def once():
    opt = BOOptimizer(hptuning_config=cfg)
    opt.add_observations(configs=configs, metrics=y.tolist())
    return opt.get_suggestion()


for _ in range(5):
    once()
t0 = time.perf_counter()
for _ in range(200):
    once()
print('p50-ish %.3f ms per call' % ((time.perf_counter() - t0) / 200 * 1e3))
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    once()
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
