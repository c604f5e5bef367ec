"""Config C5 on one GPU: G independent BO groups (d=16, n_obs=512) through suggest_groups, with 1, 2, 4 and 8
worker threads (each with its own C-ABI handle and stream).  Prints groups/s and checks that the suggestions
do not depend on the number of workers."""
import sys
import time

import numpy as np
sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200.groups import suggest_groups
from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
from polyaxon_b200.schemas.hptuning import HPTuningConfig

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n, d = 512, 16
keys = ['x%03d' % i for i in range(d)]
cfgs, raw = [], []
for g in range(G):
    X, y, _ = synthetic_problem(n, d, seed=1234 + g)
    cfgs.append(HPTuningConfig.from_dict({
        'seed': 1234 + g, 'concurrency': 1,
        'bo': {'n_iterations': 5, 'n_initial_trials': 5, 'metric': {'name': 'm', 'optimization': 'maximize'},
               'utility_function': {'acquisition_function': 'ei', 'eps': 0.0, 'n_warmup': 100000, 'n_iter': 10,
                                    'gaussian_process': {'kernel': 'matern', 'length_scale': 1.0, 'nu': 2.5,
                                                         'n_restarts_optimizer': 4}}},
        'matrix': {k: {'uniform': [0, 1]} for k in keys}}))
    raw.append({'iteration': 1, 'experiment_ids': list(range(n)),
                'experiments_configs': [[i, dict(zip(keys, row.tolist()))] for i, row in enumerate(X)],
                'experiments_metrics': [[i, float(v)] for i, v in enumerate(y)]})


def its():
    return [BOIterationConfig.from_dict(r) for r in raw]


base = None
for workers in (1, 2, 4, 8, 1, 4):
    suggest_groups(cfgs, its(), workers=workers)            # first call of every handle: draws, allocations
    t0 = time.perf_counter()
    out = suggest_groups(cfgs, its(), workers=workers)
    dt = time.perf_counter() - t0
    if base is None:
        base = out
    print('workers %d: %.1f ms for %d groups = %.1f groups/s, same suggestions: %s'
          % (workers, dt * 1e3, G, G / dt, out == base))
