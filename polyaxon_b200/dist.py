"""Multi-GPU plumbing for the suggestion path: one process per GPU, candidates partitioned by
contiguous global index range, one tiny exchange to merge the per-rank bests.

The reference has no distributed code; the only data-path exchange the
algorithm needs is the arg-max merge of ``ys.argmax()`` (acquisition_function.py:91) across shards.
NCCL has no MAXLOC, so every rank all-gathers one {value, index} record (16 bytes) and reduces
locally with numpy.argmax semantics (NaN wins, ties -> lowest global index); every rank therefore
ends with the same answer.  Works on any torch.distributed backend (nccl on GPUs, gloo in tests).
"""
import math

import torch
import torch.distributed as dist


def shard_range(total, rank, world):
    """Contiguous [lo, hi) of `total` items for `rank`; sizes differ by at most one."""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def better(av, ai, bv, bi):
    """True when candidate (av, ai) precedes (bv, bi) in numpy.argmax order."""
    an, bn = math.isnan(av), math.isnan(bv)
    if an != bn:
        return an
    if an:
        return ai < bi
    if av != bv:
        return av > bv
    return ai < bi


def merge_records(values, indices):
    """Reduce a list of per-shard (value, global index) records.  index < 0 marks an empty shard."""
    best = None
    for v, i in zip(values, indices):
        if i < 0:
            continue
        if best is None or better(v, i, best[0], best[1]):
            best = (v, i)
    return best if best is not None else (float('nan'), -1)


def merge_best(value, index, device=None, group=None):
    """All ranks contribute their local best; all ranks return the global (value, index)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value), int(index)
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device('cpu')
    v = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    i = torch.tensor([int(index)], dtype=torch.int64, device=dev)
    vs = torch.empty(world, dtype=torch.float64, device=dev)
    idx = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(vs, v, group=group)
    dist.all_gather_into_tensor(idx, i, group=group)
    return merge_records(vs.tolist(), idx.tolist())


def merge_refined(value, seed_order, device=None, group=None):
    """Merge of the refinement stage (acquisition_function.py:109: `>=`, so among equal values the
    LATER seed wins).  `seed_order` is the global position of the seed; -1 for 'no successful run'."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value), int(seed_order)
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device('cpu')
    v = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    i = torch.tensor([int(seed_order)], dtype=torch.int64, device=dev)
    vs = torch.empty(world, dtype=torch.float64, device=dev)
    idx = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(vs, v, group=group)
    dist.all_gather_into_tensor(idx, i, group=group)
    best = (float('nan'), -1)
    for val, order in sorted(zip(vs.tolist(), idx.tolist()), key=lambda t: t[1]):
        if order < 0:
            continue
        if best[1] < 0 or val >= best[0]:
            best = (val, order)
    return best


def world():
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not in use."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def gather_refined(xs, values, success, n_seeds, device=None, group=None):
    """Each rank refined seeds rank, rank + world, ...; return the three arrays for ALL seeds in seed
    order on every rank (so the acceptance rule of acquisition_function.py:105-111 runs unchanged)."""
    import numpy as np
    rank, size = world()
    if size == 1:
        return xs, values, success
    dev = device if device is not None else torch.device('cpu')
    d = xs.shape[1] if xs.ndim == 2 else 0
    per = (n_seeds + size - 1) // size
    pack = torch.zeros(per, d + 2, dtype=torch.float64, device=dev)
    if len(values):
        pack[:len(values), :d] = torch.from_numpy(np.ascontiguousarray(xs)).to(dev)
        pack[:len(values), d] = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float64)).to(dev)
        pack[:len(values), d + 1] = torch.from_numpy(np.asarray(success, dtype=np.float64)).to(dev)
    out = torch.empty(size * per, d + 2, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(out, pack, group=group)
    out = out.cpu().numpy().reshape(size, per, d + 2)
    all_x = np.empty((n_seeds, d))
    all_v = np.empty(n_seeds)
    all_s = np.zeros(n_seeds, dtype=bool)
    for s in range(n_seeds):
        r, k = s % size, s // size
        all_x[s], all_v[s], all_s[s] = out[r, k, :d], out[r, k, d], out[r, k, d + 1] > 0.5
    return all_x, all_v, all_s
