"""Many independent experiment groups at once (BASELINE config C5: 64 concurrent BO groups).

Groups are independent Gaussian processes, so this is "replicas only": under torch.distributed the
groups are dealt to the ranks by contiguous range, every rank works through its share on its own GPU
with the single-GPU path (no collective on the data path), and the suggestion dicts are gathered at
the end.  Without torch.distributed all groups run on this process's GPU.

On one GPU the groups of a rank are worked through by a few threads, each with a C-ABI handle of its
own (own stream and workspace, ``device.use_slot``): a group of C5's size (n_obs = 512) fills neither
the GPU nor the host thread that drives it, so the fit kernels, sweeps and host steps of different
groups overlap.  ``workers=1`` (or ``PLX_BO_GROUP_WORKERS=1``) is the plain loop.  The result of a
group does not depend on the worker that ran it.

Each group is one call of ``BOSearchManager.get_suggestions(iteration_config)`` exactly as
``ExperimentGroup.get_suggestions()`` would make it (db/models/experiment_groups.py:270-274).
"""
import os
from concurrent.futures import ThreadPoolExecutor

from polyaxon_b200 import device, dist as pdist
from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager


def _one_group(hptuning_config, iteration_config, g):
    manager = get_search_algorithm_manager(hptuning_config)
    if manager is None:
        raise ValueError('group {} has no search algorithm manager'.format(g))
    return manager.get_suggestions(iteration_config)


def suggest_groups(hptuning_configs, iteration_configs, gather=True, workers=None):
    """``hptuning_configs[g]`` / ``iteration_configs[g]`` describe group g.  Returns a list with one
    entry per group (the list ``get_suggestions`` returns, or None); with ``gather=False`` under
    torch.distributed only this rank's groups are filled in.  ``workers``: threads (and handles) used
    for this rank's share, default ``PLX_BO_GROUP_WORKERS`` or 4."""
    if len(hptuning_configs) != len(iteration_configs):
        raise ValueError('one iteration config per hptuning config is required')
    rank, size = pdist.world()
    n = len(hptuning_configs)
    lo, hi = pdist.shard_range(n, rank, size)
    out = [None] * n
    previous = os.environ.get('PLX_BO_DISTRIBUTED')
    sleepers = False
    os.environ['PLX_BO_DISTRIBUTED'] = '0'          # inside a group the single-GPU path is used
    try:
        if workers is None:
            workers = int(os.environ.get('PLX_BO_GROUP_WORKERS', '4') or 4)
        workers = max(1, min(int(workers), hi - lo))
        # more spinning waiters than cores would starve the threads that hold the interpreter: let them sleep on an
        # event for the duration of this call (PLX_BO_BLOCKING_SYNC, read by device.get_handle for new handles)
        try:
            cores = len(os.sched_getaffinity(0))
        except AttributeError:
            cores = os.cpu_count() or 1
        local_ranks = int(os.environ.get('LOCAL_WORLD_SIZE', '1') or 1)
        sleepers = workers * local_ranks > cores and 'PLX_BO_BLOCKING_SYNC' not in os.environ
        if sleepers:
            os.environ['PLX_BO_BLOCKING_SYNC'] = '1'
            for h in device.open_handles():
                h.set_option('blocking_sync', 1)
        if workers == 1:
            for g in range(lo, hi):
                out[g] = _one_group(hptuning_configs[g], iteration_configs[g], g)
        else:
            def work(slot):
                device.use_slot(slot + 1)                      # slot 0 stays with the calling thread
                return [(g, _one_group(hptuning_configs[g], iteration_configs[g], g))
                        for g in range(lo + slot, hi, workers)]
            with ThreadPoolExecutor(max_workers=workers) as pool:
                for part in pool.map(work, range(workers)):
                    for g, suggestion in part:
                        out[g] = suggestion
    finally:
        if previous is None:
            del os.environ['PLX_BO_DISTRIBUTED']
        else:
            os.environ['PLX_BO_DISTRIBUTED'] = previous
        if sleepers:
            del os.environ['PLX_BO_BLOCKING_SYNC']
            for h in device.open_handles():
                h.set_option('blocking_sync', 0)
    if size > 1 and gather:
        parts = [None] * size
        pdist.dist.all_gather_object(parts, out[lo:hi])
        out = [s for part in parts for s in part]
    return out
