"""World-size-2 gloo tests of the shard/merge logic (host side of the multi-GPU path)."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from polyaxon_b200 import dist as pd


def test_shard_ranges_cover_everything():
    for total in (0, 1, 7, 128, 1000003):
        for world in (1, 2, 3, 8):
            spans = [pd.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_merge_records_is_numpy_argmax():
    rs = np.random.RandomState(0)
    for trial in range(200):
        ys = rs.randint(0, 4, size=23).astype(float)
        if trial % 3 == 0:
            ys[rs.randint(0, 23, size=2)] = np.nan
        cuts = sorted(rs.choice(np.arange(1, 23), size=3, replace=False))
        vals, idxs = [], []
        for lo, hi in zip([0] + cuts, cuts + [23]):
            j = int(np.argmax(ys[lo:hi]))
            vals.append(ys[lo + j])
            idxs.append(lo + j)
        v, i = pd.merge_records(vals, idxs)
        assert i == int(np.argmax(ys))
    assert pd.merge_records([1.0, 2.0], [-1, -1]) == pytest.approx((float('nan'), -1), nan_ok=True)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, ys, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    lo, hi = pd.shard_range(len(ys), rank, world)
    local = ys[lo:hi]
    if len(local):
        j = int(np.argmax(local))
        v, i = pd.merge_best(local[j], lo + j)
    else:
        v, i = pd.merge_best(float('nan'), -1)
    rv, ri = pd.merge_refined(float(rank == 0) * 2.0 + 1.0 * (rank == 1) * 2.0, rank)
    # refinement results of 5 seeds dealt round-robin: rank r holds seeds r, r + 2, ...
    seeds = np.arange(5)
    mine = seeds[rank::world]
    gx, gv, gs = pd.gather_refined(np.stack([mine * 1.0, mine * 10.0], axis=1), mine * 0.5, mine % 2 == 0, 5)
    assert np.array_equal(gx, np.stack([seeds * 1.0, seeds * 10.0], axis=1))
    assert np.array_equal(gv, seeds * 0.5) and np.array_equal(gs, seeds % 2 == 0)
    assert pd.world() == (rank, world)
    out[rank] = (v, i, rv, ri)
    dist.destroy_process_group()


@pytest.mark.parametrize('case', ['nan', 'tie', 'short'])
def test_two_rank_merge_over_gloo(case):
    rs = np.random.RandomState(1)
    ys = rs.randn(1001)
    if case == 'nan':
        ys[700] = np.nan
        ys[900] = np.nan
    elif case == 'tie':
        ys[100] = ys[800] = 10.0
    elif case == 'short':
        ys = ys[:1]
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, ys, out), nprocs=2, join=True)
    want = int(np.argmax(ys))
    for rank in range(2):
        v, i, rv, ri = out[rank]
        assert i == want
        assert (np.isnan(v) and np.isnan(ys[want])) or v == ys[want]
        assert (rv, ri) == (2.0, 1)        # equal values: the later seed wins


def _groups_worker(rank, world, port, out):
    """suggest_groups deals groups to ranks by contiguous range and gathers the suggestion lists.  Random-search
    groups need no GPU, so the dealing / gathering logic of config C5 runs here on CPU."""
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from polyaxon_b200.groups import suggest_groups
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    cfgs = [HPTuningConfig.from_dict({'seed': 10 + g, 'concurrency': 1, 'random_search': {'n_experiments': 3},
                                      'matrix': {'lr': {'uniform': [0, 1]}, 'units': {'values': [16, 32, 64]}}})
            for g in range(5)]
    everything = suggest_groups(cfgs, [None] * 5, workers=2)
    mine_only = suggest_groups(cfgs, [None] * 5, gather=False, workers=1)
    out[rank] = (everything, mine_only)
    dist.destroy_process_group()


def test_groups_are_dealt_to_ranks_and_gathered_over_gloo():
    from polyaxon_b200.hpsearch.search_managers import get_search_algorithm_manager
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_groups_worker, args=(2, port, out), nprocs=2, join=True)
    cfgs = [HPTuningConfig.from_dict({'seed': 10 + g, 'concurrency': 1, 'random_search': {'n_experiments': 3},
                                      'matrix': {'lr': {'uniform': [0, 1]}, 'units': {'values': [16, 32, 64]}}})
            for g in range(5)]
    want = [get_search_algorithm_manager(c).get_suggestions() for c in cfgs]
    for rank in range(2):
        everything, mine_only = out[rank]
        assert everything == want                                   # every rank holds every group's suggestions
        lo, hi = pd.shard_range(5, rank, 2)
        assert mine_only[lo:hi] == want[lo:hi]
        assert all(s is None for s in mine_only[:lo] + mine_only[hi:])
    assert os.environ.get('PLX_BO_DISTRIBUTED') is None
