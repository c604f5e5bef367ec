"""C5 share of one GPU (8 groups, 8 worker threads) with the host restricted to FEW cores, spinning against sleeping
waits: emulates a rank of an 8-GPU job on a 16-core box (2 cores per rank).
    taskset -c 0,1 python tools/groups_cores.py"""
import os
import sys
import time

sys.path.insert(0, '.')
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import device
from polyaxon_b200.groups import suggest_groups
from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
from polyaxon_b200.schemas.hptuning import HPTuningConfig

G, n, d = 8, 512, 16
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
print('cores available:', len(os.sched_getaffinity(0)))
for mode in ('0', 'auto', '0', 'auto'):
    if mode == '0':
        os.environ['PLX_BO_BLOCKING_SYNC'] = '0'            # present => suggest_groups leaves it alone: spinning waits
    else:
        os.environ.pop('PLX_BO_BLOCKING_SYNC', None)
    for h in device.open_handles():
        h.set_option('blocking_sync', 0)
    times = []
    for _ in range(4):
        its = [BOIterationConfig.from_dict(r) for r in raw]
        t0 = time.perf_counter()
        suggest_groups(cfgs, its, workers=8)
        times.append(time.perf_counter() - t0)
    print('blocking_sync %s: %.1f groups/s' % (mode, G / sorted(times[1:])[1]))
