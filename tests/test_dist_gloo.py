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


class _FakeHandle(object):
    """Stands in for the C-ABI handle in the host-side tests of the sharded hyper-parameter fit and the sharded
    warm-up sweep (no GPU here): every run converges where it started with LML = -(theta - 1)^2, except that
    thetas above 5 are 'not positive definite' (LML = -inf), as a very long length scale is on real data."""
    device = 0

    def __init__(self):
        self.set_theta_calls = []
        self.held = None
        self.owner = None
        self.seed_cache = {}
        self.resident_key = None

    def fit(self, theta0, restarts=(), bounds=None, apply=True):
        starts = [theta0] + list(restarts)
        # a rank of a sharded fit must not build the posterior state from its own runs
        assert apply is False or len(starts) == 1
        rows = np.array([[t, t, -np.inf if t > 5 else -(t - 1.0) ** 2, 3.0] for t in starts])
        best = int(np.argmin(-rows[:, 2]))
        if apply:
            self.set_theta(rows[best, 1])
        return float(rows[best, 1]), float(rows[best, 2]), rows

    def set_theta(self, theta):
        if theta > 5:
            raise np.linalg.LinAlgError('kernel matrix is not positive definite')
        self.set_theta_calls.append(theta)


def _fit_worker(rank, world, port, starts, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import (
        DeviceGaussianProcess)
    gp = DeviceGaussianProcess('matern', 1.0, 1.5, len(starts) - 1, np.random.RandomState(1), handle=_FakeHandle())
    try:
        theta, lml, per_run = gp._fit_runs(starts[0], starts[1:])
        out[rank] = ('ok', theta, lml, per_run.tolist(), gp.handle.set_theta_calls)
    except np.linalg.LinAlgError as exc:
        out[rank] = ('linalg', str(exc))
    dist.destroy_process_group()


@pytest.mark.parametrize('starts,want', [
    ([0.0, 9.0, 2.0, 8.0, 1.5], ('ok', 1.5)),      # rank 1 gets 9.0 and 8.0: every one of its runs is non-PD
    ([7.0, 9.0], ('linalg', None)),                # no run anywhere is PD: every rank raises, none hangs
    ([0.5], ('ok', 0.5)),
])
def test_sharded_fit_survives_a_rank_without_a_positive_definite_run(starts, want):
    """a rank whose share of the optimiser runs is all non-positive-definite used to raise before the
    all-gather and leave the other rank blocked in it.  Now the ranks exchange their runs first and build the
    state at the global winner (sklearn: argmin over -LML with -inf LML = +inf, _gpr.py:336-337)."""
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_fit_worker, args=(2, port, starts, out), nprocs=2, join=True)
    for rank in range(2):
        assert out[rank][0] == want[0]
        if want[0] == 'ok':
            _, theta, lml, per_run, calls = out[rank]
            assert theta == want[1] and calls == [want[1]]
            assert [row[0] for row in per_run] == starts        # gathered back in run order
    if want[0] == 'ok' and len(starts) > 1:
        assert out[0][1:4] == out[1][1:4]


def _merge_x_worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    # rank 0 holds rows 0..4, rank 1 rows 5..9; values tie at 3.0 on rows 2 and 7 -> the lower index wins
    local = (3.0, 2 + 5 * rank, np.array([10.0 * rank + 1, 10.0 * rank + 2]))
    res = [pd.merge_best_x(*local)]
    # an empty shard contributes index -1 and is ignored
    res.append(pd.merge_best_x(float('nan'), -1, np.zeros(2)) if rank == 0 else pd.merge_best_x(1.0, 8, np.array([7.0, 8.0])))
    # a huge index survives the trip through the float64 record bit for bit
    res.append(pd.merge_best(float(rank), (1 << 62) + 12345 + rank))
    out[rank] = [(v, i) + ((x.tolist(),) if len(r) == 3 else ()) for r in res for v, i, *rest in [r]
                 for x in [rest[0] if rest else None]]
    dist.destroy_process_group()


def test_single_record_merge_carries_index_bits_and_the_winning_row():
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_merge_x_worker, args=(2, port, out), nprocs=2, join=True)
    for rank in range(2):
        a, b, c = out[rank]
        assert a == (3.0, 2, [1.0, 2.0])
        assert b == (1.0, 8, [7.0, 8.0])
        assert c == (1.0, (1 << 62) + 12346)


class _FakeSweepHandle(_FakeHandle):
    """Host-side stand-in with a toy acquisition (sum of the coordinates) so that the sharded warm-up of
    UtilityFunction can run on CPU: every rank holds the candidates it was given and sweeps the rows it is asked for."""
    d = 3

    def candidates_draw(self, rng, low, high, N, key=None, first=0, count=None):
        return False                                     # "odd stream position": the caller draws on the host

    def candidates_set(self, Xc, key=None):
        self.Xc = np.asarray(Xc)
        self.resident_key = key

    def sweep_resident_range(self, first, count, acq, y_max=0.0, kappa=0.0, eps=0.0):
        if count == 0:
            return float('nan'), -1, np.zeros(self.d)
        ys = self.Xc[first:first + count].sum(axis=1)
        j = int(np.argmax(ys))
        return float(ys[j]), first + j, self.Xc[first + j].copy()

    def last_sweep_clamped(self):
        return 0


def _warmup_worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.acquisition_function import UtilityFunction
    uf = UtilityFunction.__new__(UtilityFunction)
    uf.random_generator = np.random.RandomState(42)
    h = _FakeSweepHandle()
    bounds = np.array([[0.0, 1.0], [-1.0, 2.0], [5.0, 6.0]])
    res = []
    for _ in range(2):                                   # second call: the resident-matrix cache of the group is hit
        uf.random_generator = np.random.RandomState(42)
        max_acq, x_max, x_seeds = uf._warmup(h, 0, dict(y_max=0.0, kappa=0.0, eps=0.0), bounds, 1001, 4)
        res.append((max_acq, x_max.tolist(), x_seeds.tolist(), uf.random_generator.uniform()))
    out[rank] = res
    dist.destroy_process_group()


def test_sharded_warmup_returns_the_single_process_answer_on_every_rank():
    """Each rank sweeps its block of the one candidate matrix and a single all-gather of {value, index, row} records
    gives every rank the global winner's row and leaves every generator where numpy would be."""
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_warmup_worker, args=(2, port, out), nprocs=2, join=True)
    rs = np.random.RandomState(42)
    bounds = np.array([[0.0, 1.0], [-1.0, 2.0], [5.0, 6.0]])
    x_tries = rs.uniform(bounds[:, 0], bounds[:, 1], size=(1001, 3))
    x_seeds = rs.uniform(bounds[:, 0], bounds[:, 1], size=(4, 3))
    nxt = rs.uniform()
    j = int(np.argmax(x_tries.sum(axis=1)))
    for rank in range(2):
        for max_acq, x_max, seeds, after in out[rank]:
            assert max_acq == x_tries[j].sum() and x_max == x_tries[j].tolist()
            assert seeds == x_seeds.tolist() and after == nxt
