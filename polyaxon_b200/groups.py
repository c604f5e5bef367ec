"""Many independent experiment groups at once (BASELINE config C5: 64 concurrent BO groups).

Groups are independent Gaussian processes, so this is "replicas only": under torch.distributed the
groups are dealt to the ranks by contiguous range, every rank works through its share on its own GPU
with the single-GPU path (no collective on the data path), and the suggestion dicts are gathered at
the end.  Without torch.distributed all groups run on this process's GPU.

Each group is one call of ``BOSearchManager.get_suggestions(iteration_config)`` exactly as
``ExperimentGroup.get_suggestions()`` would make it (db/models/experiment_groups.py:270-274).
"""
import os

from polyaxon_b200 import dist as pdist
from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager


def suggest_groups(hptuning_configs, iteration_configs, gather=True):
    """``hptuning_configs[g]`` / ``iteration_configs[g]`` describe group g.  Returns a list with one
    entry per group (the list ``get_suggestions`` returns, or None); with ``gather=False`` under
    torch.distributed only this rank's groups are filled in."""
    if len(hptuning_configs) != len(iteration_configs):
        raise ValueError('one iteration config per hptuning config is required')
    rank, size = pdist.world()
    n = len(hptuning_configs)
    lo, hi = pdist.shard_range(n, rank, size)
    out = [None] * n
    previous = os.environ.get('PLX_BO_DISTRIBUTED')
    os.environ['PLX_BO_DISTRIBUTED'] = '0'          # inside a group the single-GPU path is used
    try:
        for g in range(lo, hi):
            manager = get_search_algorithm_manager(hptuning_configs[g])
            if manager is None:
                raise ValueError('group {} is not a Bayesian-optimisation group'.format(g))
            out[g] = manager.get_suggestions(iteration_configs[g])
    finally:
        if previous is None:
            del os.environ['PLX_BO_DISTRIBUTED']
        else:
            os.environ['PLX_BO_DISTRIBUTED'] = previous
    if size > 1 and gather:
        parts = [None] * size
        pdist.dist.all_gather_object(parts, out[lo:hi])
        out = [s for part in parts for s in part]
    return out
