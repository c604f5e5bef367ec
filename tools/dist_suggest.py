"""Multi-GPU get_suggestion under torchrun: every rank must return the single-process answer.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_suggest.py"""
import os
import sys
import time
import numpy as np
sys.path.insert(0, '.')
import torch
import torch.distributed as dist
from oracle.reference_path import synthetic_problem
from polyaxon_b200 import device
from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
from polyaxon_b200.schemas.hptuning import HPTuningConfig

rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
n, d, N = int(os.environ.get('N_OBS', 512)), 16, int(os.environ.get('N_CAND', 400000))
X, y, _ = synthetic_problem(n, d)
keys = ['x%03d' % i for i in range(d)]
cfg = HPTuningConfig.from_dict({
    'seed': 9, 'concurrency': 1,
    'bo': {'n_iterations': 5, 'n_initial_trials': 5, 'metric': {'name': 'm', 'optimization': 'maximize'},
           'utility_function': {'acquisition_function': 'ei', 'eps': 0.0, 'n_warmup': N, 'n_iter': 7,
                                'gaussian_process': {'kernel': 'matern', 'length_scale': 1.0, 'nu': 2.5,
                                                     'n_restarts_optimizer': 5}}},
    'matrix': {k: {'uniform': [0, 1]} for k in keys}})
configs = [dict(zip(keys, row.tolist())) for row in X]


def run():
    device.get_handle().seed_cache.clear()
    opt = BOOptimizer(hptuning_config=cfg)
    opt.add_observations(configs=configs, metrics=y.tolist())
    t0 = time.perf_counter()
    s = opt.get_suggestion()
    return s, (time.perf_counter() - t0) * 1e3


os.environ['PLX_BO_DISTRIBUTED'] = '0'
single, t_single = run()
os.environ['PLX_BO_DISTRIBUTED'] = '1'
multi, t_multi = run()
assert multi == single, (rank, multi, single)
gathered = [None] * dist.get_world_size()
dist.all_gather_object(gathered, multi)
assert all(g == gathered[0] for g in gathered)
if rank == 0:
    print('world %d: identical suggestion on every rank and to the single-GPU path; single %.1f ms, sharded %.1f ms'
          % (dist.get_world_size(), t_single, t_multi))
dist.destroy_process_group()
