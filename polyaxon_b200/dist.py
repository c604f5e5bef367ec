"""Multi-GPU plumbing for the suggestion path: one process per GPU, candidates partitioned by
contiguous global index range, one tiny exchange to merge the per-rank bests.

The reference has no distributed code; the only data-path exchange the
algorithm needs is the arg-max merge of ``ys.argmax()`` (acquisition_function.py:91) across shards.
NCCL has no MAXLOC, so every rank all-gathers one {value, index} record (16 bytes) and reduces
locally with numpy.argmax semantics (NaN wins, ties -> lowest global index); every rank therefore
ends with the same answer.  Works on any torch.distributed backend (nccl on GPUs, gloo in tests).
"""
import math

import numpy as np
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


def _active(group=None):
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def _gather_rows(record, device=None, group=None):
    """ONE collective and ONE device->host copy: every rank contributes `record` (a 1-D float64 numpy array, the
    same length on every rank) and gets the (world, len) array of all of them.  Integers travel as the bit pattern
    of a float64 (no arithmetic touches them)."""
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device('cpu')
    mine = torch.from_numpy(np.ascontiguousarray(record, dtype=np.float64)).to(dev)
    out = torch.empty(world * mine.numel(), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(out, mine, group=group)
    return out.cpu().numpy().reshape(world, -1)


def _record(value, index, extra=()):
    rec = np.empty(2 + len(extra), dtype=np.float64)
    rec[0] = value
    rec[1:2].view(np.int64)[0] = int(index)
    rec[2:] = extra
    return rec


def merge_best(value, index, device=None, group=None):
    """All ranks contribute their local best as one 16-byte {value, index} record; all ranks return the global
    (value, index)."""
    if not _active(group):
        return float(value), int(index)
    rows = _gather_rows(_record(value, index), device, group)
    return merge_records(rows[:, 0].tolist(), np.ascontiguousarray(rows[:, 1]).view(np.int64).tolist())


def merge_best_x(value, index, x, device=None, group=None):
    """merge_best for ranks that hold only their own block of candidates: the record also carries the local
    winner's row (16 + 8 d bytes), so every rank returns (value, index, x of the global winner) although only one
    of them has that row in memory."""
    x = np.asarray(x, dtype=np.float64).ravel()
    if not _active(group):
        return float(value), int(index), x
    rows = _gather_rows(_record(value, index, x), device, group)
    indices = np.ascontiguousarray(rows[:, 1]).view(np.int64)
    v, i = merge_records(rows[:, 0].tolist(), indices.tolist())
    if i < 0:
        return v, i, x
    return v, i, rows[int(np.flatnonzero(indices == i)[0]), 2:].copy()


def merge_refined(value, seed_order, device=None, group=None):
    """Merge of the refinement stage (acquisition_function.py:109: `>=`, so among equal values the
    LATER seed wins).  `seed_order` is the global position of the seed; -1 for 'no successful run'."""
    if not _active(group):
        return float(value), int(seed_order)
    rows = _gather_rows(_record(value, seed_order), device, group)
    orders = np.ascontiguousarray(rows[:, 1]).view(np.int64).tolist()
    best = (float('nan'), -1)
    for val, order in sorted(zip(rows[:, 0].tolist(), orders), key=lambda t: t[1]):
        if order < 0:
            continue
        if best[1] < 0 or val >= best[0]:
            best = (val, order)
    return best


def gather_rows(record, device=None, group=None):
    """Public form of the one-collective gather (used by the sharded hyper-parameter fit)."""
    if not _active(group):
        return np.asarray(record, dtype=np.float64).reshape(1, -1)
    return _gather_rows(record, device, group)


def collective_device(cuda_index):
    """Where the records of a collective live: the rank's GPU under NCCL, host memory under gloo (CPU tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_backend() == 'nccl':
        return torch.device('cuda', int(cuda_index))
    return torch.device('cpu')


def world():
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not in use."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def gather_refined(xs, values, success, n_seeds, device=None, group=None):
    """Each rank refined seeds rank, rank + world, ...; return the three arrays for ALL seeds in seed
    order on every rank (so the acceptance rule of acquisition_function.py:105-111 runs unchanged)."""
    rank, size = world()
    if size == 1:
        return xs, values, success
    xs = np.asarray(xs, dtype=np.float64)
    d = xs.shape[1] if xs.ndim == 2 else 0
    per = (n_seeds + size - 1) // size
    pack = np.zeros((per, d + 2))
    k = len(values)
    if k:
        pack[:k, :d] = xs
        pack[:k, d] = values
        pack[:k, d + 1] = np.asarray(success, dtype=np.float64)
    out = _gather_rows(pack.ravel(), device, group).reshape(size, per, d + 2)
    all_x = np.empty((n_seeds, d))
    all_v = np.empty(n_seeds)
    all_s = np.zeros(n_seeds, dtype=bool)
    for s in range(n_seeds):
        r, k = s % size, s // size
        all_x[s], all_v[s], all_s[s] = out[r, k, :d], out[r, k, d], out[r, k, d + 1] > 0.5
    return all_x, all_v, all_s
