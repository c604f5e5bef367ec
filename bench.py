#!/usr/bin/env python
"""Benchmark of the B200-native Bayesian-optimisation suggestion path.

Metric (BASELINE.json): acquisition evaluations per second (FP64) at d=32, n_obs=2048, plus
get_suggestions p50 latency.  A "step" is one warm-up sweep (acquisition_function.py:88-92) over the
whole candidate set of config C3 (16M candidates, Matern nu=1.5, POI eps=0) at a fixed GP state.

ONE global candidate matrix: rows of RandomState(CAND_SEED).uniform(0, 1, (N, 32)), the reference's own draw
(acquisition_function.py:88-89).  Every rank draws only its contiguous block of rows on its GPU (MT19937
jump-ahead, bit-identical to numpy), sweeps it, and one all-gather of a 16-byte {value, index} record per rank
merges the bests (strong scaling: the global candidate count is fixed).  `best_index` is therefore the same
number at every N, and rank 0 checks the first `--cpu-sample` rows against the CPU reference path's arg-max.

  value     whole-job candidates/s with the candidates resident in HBM (device-timed, max over ranks)
  e2e       the same through the C-ABI call with HOST (pinned) candidate buffers: host->device copies and the
            device->host read of the result are inside the timed region
  roofline  the sweep kernel.  From 256 observations the variance contraction runs as exact int8 digit-plane
            products on the tcgen05 tensor cores, so the bound is the int8 tensor pipe: achieved = executed int8
            tensor ops per candidate x candidates / kernel time (CUDA events on the kernel's stream), peak = the
            m128n128k32 kind::i8 issue rate measured in this run.  The FP64-equivalent rate (algorithmic flops of
            SURVEY.md 8d over the same time) and the FP64 DMMA form of the sweep (a few steps of sweep_impl=1 against
            the DGEMM rate measured in this run) are reported beside it as scalars of the same object.
  cpu_baseline  the reference's sklearn/scipy path (oracle/reference_path.py) on the host cores, on a bounded
            sample (the first --cpu-sample rows of the same matrix)

`--impl reference` times that CPU path alone (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, N_OBS, NU = 32, 2048, 1.5
THETA = 1.2                      # log length-scale of the benchmark state (near the fitted optimum)
CAND_SEED = 20240                # RandomState seed of the global candidate matrix
METRIC = 'acquisition evals/sec (FP64) at d=32,n_obs=2048'
UNIT = 'evals/s'
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the kernel over 151 552 candidates, from the
# `ncu --set full` capture of this build kept under profiles/ (a constant from that file, not measured in this run)
TRAFFIC = {
    'sweep_i8c_kernel': ('profiles/r02_ncu_sweep_i8c_full_summary.csv', 2.948484e9 + 2.186265e9, 151552),
    'sweep_tma_kernel': ('profiles/r01_ncu_sweep_tma_v3_summary.csv', 18.701e9 + 2.490e9, 151552),
}


def measured_peaks():
    """The driver-written peaks of this pool (MEASURED_PEAKS.json at the repo root), or None."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def flops_per_eval(n, d):
    """SURVEY.md section 8d / BASELINE.md: n(3d+10) + n^2 + 4n (triangular variance)."""
    return n * (3 * d + 10) + n * n + 4 * n


def i8_ops_per_eval(n):
    """Executed int8 tensor ops per candidate: 28 digit-plane pairs, 64-row half-blocks of L^-1, triangular in
    64-wide chunks (sweep_i8.cuh)."""
    nh = (n + 63) // 64
    return 28 * 2.0 * 64 * 64 * nh * (nh + 1) / 2


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
              'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
              'clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu_index = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.gpu_index), '--query-gpu=' + self.FIELDS,
                 '--format=csv,noheader,nounits', '-lms', '200'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(',')])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, power, reasons, sm_max = [], [], set(), None
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                sm_max = float(r[1])
                power.append(float(r[2]))
                for name, flag in zip(names, r[3:7]):
                    if flag.lower().startswith('active'):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {'sm_mhz': float(np.median(busy)) if busy else None, 'sm_max_mhz': sm_max,
                'reasons': sorted(reasons), 'samples': len(sm), 'power_w_max': max(power) if power else None}


def candidate_rows(count):
    """The first `count` rows of the global candidate matrix, drawn by numpy on the host (row-major fill: the
    first rows of the N x d draw are the draw of a count x d matrix from the same state)."""
    return np.random.RandomState(CAND_SEED).uniform(0.0, 1.0, size=(count, D))


def cpu_reference_step_fn(sample):
    """The reference's CPU implementation of one step on `sample` candidates (sklearn predict + scipy norm)."""
    import warnings
    from oracle import gp_numpy as G
    from oracle import reference_path as R
    X, y, _ = R.synthetic_problem(N_OBS, D)
    gp = R.make_gp(G.KERNEL_MATERN, float(np.exp(THETA)), NU, 0, np.random.RandomState(1))
    gp.optimizer = None                    # fixed GP state, as in our arm
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        gp.fit(X, y)
    y_max = float(y.max())
    xc = candidate_rows(sample)

    def step():
        # chunks of 65 536 rows: BASELINE.md section 2 (K* of one predict call is chunk x n_obs doubles)
        ys = np.concatenate([R.acquisition(gp, xc[s:s + 65536], G.ACQ_POI, y_max, 0.0, 0.0)
                             for s in range(0, sample, 65536)])
        return int(ys.argmax()), float(ys.max())
    return step


def run_cpu_reference(sample, steps, warmup):
    from threadpoolctl import threadpool_limits
    step = cpu_reference_step_fn(sample)
    cores = os.cpu_count()
    # torchrun exports OMP_NUM_THREADS=1; the baseline is entitled to every host core
    with threadpool_limits(limits=cores):
        best = None
        for _ in range(warmup):
            best = step()
        times = []
        for _ in range(steps):
            t0 = time.perf_counter()
            best = step()
            times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    return {'value': sample / dt, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': 'first %d rows of the global candidate matrix per step (predict calls of 65 536 rows), %d steps '
                      'after %d warm-up, p50 %.0f ms; sklearn %s / scipy path of oracle/reference_path.py'
                      % (sample, steps, warmup, 1e3 * float(np.median(times)), __import__('sklearn').__version__)}, dt, best


def reference_arm(args, rank):
    """`--impl reference`: the reference's CPU implementation on the host cores, same config, same K and W."""
    if rank != 0:
        return
    base, dt, best = run_cpu_reference(args.cpu_sample, args.steps, args.warmup)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': base['value'], 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3,
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': workload_config(args, 1),
        'cpu_baseline': base,
        'e2e': {'value': base['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'best_index': best[0], 'best_value': best[1],
    }
    print(json.dumps(line))


def c3_config():
    from oracle.reference_path import synthetic_problem
    X, y, _ = synthetic_problem(2048, 32)
    return ('C3 d=32 n=2048 matern1.5 poi N=16M S=10 R=17',
            dict(kernel='matern', nu=1.5, acq='poi', kappa=None, eps=0.0, n_warmup=1 << 24, n_iter=10,
                 restarts=16, reps=2, cpu=False), X, y)


def latency_configs():
    """C1 and C2 of BASELINE.json as hptuning sections + observations (for get_suggestions latency).  C1's `lr` is
    declared `uniform` over the log range: the reference's SearchSpace gives a `loguniform` parameter a feature
    but no bound (space.py:90 / :117), so a loguniform C1 cannot run through the reference either."""
    from oracle.reference_path import synthetic_problem
    out = []
    X, y, _ = synthetic_problem(15, 2)
    out.append(('C1 d=2 n=15 matern2.5 ucb N=1e5 S=10 (lr declared uniform over the log range, dropout uniform)',
                dict(kernel='matern', nu=2.5, acq='ucb', kappa=2.576, eps=None, n_warmup=100000, n_iter=10,
                     restarts=0), X, y))
    X, y, _ = synthetic_problem(256, 8)
    out.append(('C2 d=8 n=256 rbf ei N=1e6 S=10 (8 uniform dimensions)',
                dict(kernel='rbf', nu=1.5, acq='ei', kappa=None, eps=0.0, n_warmup=1000000, n_iter=10,
                     restarts=0), X, y))
    return out


def measure_latency(configs, reps_gpu=7, with_cpu=True):
    """p50 of BOOptimizer.get_suggestion() through the drop-in Python classes (under torch.distributed every rank
    makes the same call: sharded fit, sharded draw + sweep, sharded refinement), and the same call on the CPU
    reference path (oracle/reference_path.py), same seed and observations."""
    import warnings
    from oracle import gp_numpy as G
    from oracle import reference_path as R
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    from threadpoolctl import threadpool_limits
    results = []
    for label, c, X, y in configs:
        d = X.shape[1]
        keys = ['x%03d' % i for i in range(d)]
        uf = {'acquisition_function': c['acq'], 'n_warmup': c['n_warmup'], 'n_iter': c['n_iter'],
              'gaussian_process': {'kernel': c['kernel'], 'length_scale': 1.0, 'nu': c['nu'],
                                   'n_restarts_optimizer': c['restarts']}}
        uf['kappa' if c['acq'] == 'ucb' else 'eps'] = c['kappa'] if c['acq'] == 'ucb' else c['eps']
        cfg = HPTuningConfig.from_dict({
            'seed': 17, 'concurrency': 1,
            'bo': {'n_iterations': 5, 'n_initial_trials': 5, 'metric': {'name': 'm', 'optimization': 'maximize'},
                   'utility_function': uf},
            'matrix': {k: {'uniform': [0, 1]} for k in keys}})
        configs_ = [dict(zip(keys, row.tolist())) for row in X]
        times, suggestion = [], None
        for _ in range(c.get('reps', reps_gpu) + 1):
            t0 = time.perf_counter()
            opt = BOOptimizer(hptuning_config=cfg)
            opt.add_observations(configs=configs_, metrics=y.tolist())
            suggestion = opt.get_suggestion()
            times.append((time.perf_counter() - t0) * 1e3)
        gpu_ms = float(np.median(times[1:]))
        first_ms = float(times[0])
        bounds = np.tile(np.array([[0.0, 1.0]]), (d, 1))
        cpu = []
        if with_cpu and c.get('cpu', True):
            with threadpool_limits(limits=os.cpu_count()), warnings.catch_warnings():
                warnings.simplefilter('ignore')
                for _ in range(2 if c['n_warmup'] <= 100000 else 1):
                    t0 = time.perf_counter()
                    R.suggest(X, y, bounds, G.KERNEL_RBF if c['kernel'] == 'rbf' else G.KERNEL_MATERN, 1.0, c['nu'],
                              c['restarts'], {'ucb': 0, 'ei': 1, 'poi': 2}[c['acq']], c['kappa'] or 0.0, c['eps'] or 0.0,
                              17, c['n_warmup'], c['n_iter'], chunk=65536)
                    cpu.append((time.perf_counter() - t0) * 1e3)
        results.append({'config': label, 'get_suggestions_p50_ms': gpu_ms, 'first_call_ms': first_ms,
                        'cpu_reference_ms': float(np.median(cpu)) if cpu else None,
                        'suggestion_x000': float(suggestion['x000']),
                        'includes': 'fit, sweep, refinement, decode; the first call of the seeded group also draws the '
                                    'candidates (numpy MT19937 stream regenerated on the device, bit-identical), later '
                                    'calls reuse the device-resident matrix (same generator state => same candidates)'})
    return results


def measure_fit(h):
    """Config C4: hyper-parameter fit (33 L-BFGS-B runs on theta advanced in lock step) at growing n, d=32,
    Matern nu=1.5; the sklearn fit of the CPU reference path beside it while it stays within seconds."""
    import warnings
    from oracle import gp_numpy as G
    from oracle import reference_path as R
    from oracle.reference_path import synthetic_problem
    from threadpoolctl import threadpool_limits
    out = []
    for n in (64, 128, 256, 512, 1024, 2048, 4096):
        X, y, _ = synthetic_problem(n, 32)
        rs = np.random.RandomState(17)
        restarts = rs.uniform(G.THETA_BOUNDS[0], G.THETA_BOUNDS[1], size=32)
        times = []
        for _ in range(3 if n <= 2048 else 2):
            t0 = time.perf_counter()
            h.set_data(X, y, 1, 1.5)
            theta, lml, per_run = h.fit(0.0, restarts)
            times.append((time.perf_counter() - t0) * 1e3)
        evals = int(per_run[:, 3].sum())
        flop = evals * (n * n * (3 * 32 + 10) + float(n) ** 3 + 4 * n * n)
        rec = {'n_obs': n, 'restarts': 33, 'fit_ms': float(np.median(times[1:])), 'lml_evaluations': evals,
               'tflops_model': flop / (float(np.median(times[1:])) * 1e-3) / 1e12, 'theta': theta, 'lml': lml,
               'cpu_reference_ms': None}
        if n <= 256:
            with threadpool_limits(limits=os.cpu_count()), warnings.catch_warnings():
                warnings.simplefilter('ignore')
                gp = R.make_gp(G.KERNEL_MATERN, 1.0, 1.5, 32, R.make_rng(17))
                t0 = time.perf_counter()
                gp.fit(X, y)
                rec['cpu_reference_ms'] = (time.perf_counter() - t0) * 1e3
                rec['cpu_theta'] = float(gp.kernel_.theta[0])
        out.append(rec)
    return out


def measure_groups(n_groups=64, reps=3):
    """Config C5: 64 independent groups (d=16, n_obs=512) through polyaxon_b200.groups.suggest_groups; under
    torch.distributed they are dealt to the ranks by contiguous range (8 per GPU on 8 GPUs), each rank works through
    its share with several worker threads (a C-ABI handle and stream each), suggestions are gathered at the end."""
    from oracle.reference_path import synthetic_problem
    from polyaxon_b200.groups import suggest_groups
    from polyaxon_b200.hpsearch.schemas.bayesian_optimization import BOIterationConfig
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    n, d = 512, 16
    keys = ['x%03d' % i for i in range(d)]
    cfgs, raw = [], []
    for g in range(n_groups):
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
    times, got = [], None
    for _ in range(reps + 1):
        its = [BOIterationConfig.from_dict(r) for r in raw]
        t0 = time.perf_counter()
        got = suggest_groups(cfgs, its, workers=8)
        times.append(time.perf_counter() - t0)
    return n_groups / float(np.median(times[1:])), got


def workload_config(args, world):
    return {'workload': 'C3: BO 32-dim continuous, POI eps=0, matern nu=1.5, n_obs=2048, '
                        'num_warmup=%d candidates (global), fixed GP state theta=%.1f' % (args.candidates, THETA),
            'candidates': args.candidates, 'd': D, 'n_obs': N_OBS, 'parallelism': 'shard%d' % world,
            'candidate_matrix': 'rows of numpy RandomState(%d).uniform(0, 1, (N, 32)); each rank draws its block on '
                                'the device (MT19937 jump-ahead, bit-identical)' % CAND_SEED,
            'l2': 'inputs larger than L2 (candidate block >= 0.5 GB per rank)'}


def measure_general_nu(nu=1.9, N=148 * 128 * 8):
    """Matern nu = 1.9 (the value in the reference's own fixtures) at the C3 shape: one sweep over 151 552 device
    candidates with the kernel value from the per-nu Chebyshev tables (default) and from the Bessel routine per value,
    and a three-run fit the same two ways.  Same arg-max both ways."""
    import torch
    from polyaxon_b200 import _lib
    from oracle.reference_path import synthetic_problem
    X, y, _ = synthetic_problem(N_OBS, D)
    g = _lib.Handle(int(os.environ.get('LOCAL_RANK', '0')))
    out = {'nu': nu, 'candidates': N, 'n_obs': N_OBS, 'd': D}
    try:
        xc = torch.rand(N, D, dtype=torch.float64, device='cuda', generator=torch.Generator(device='cuda').manual_seed(7))
        torch.cuda.synchronize()
        rs = np.random.RandomState(3)
        restarts = rs.uniform(np.log(1e-5), np.log(1e5), 2)
        best = {}
        for tab, key in ((1, 'tables'), (0, 'routine')):
            g.set_option('gen_table', tab)
            g.set_data(X, y, _lib.KERNEL_MATERN, nu)
            t0 = time.perf_counter()
            g.fit(0.0, restarts, apply=False)
            out['fit3_ms_' + key] = (time.perf_counter() - t0) * 1e3
            g.set_theta(THETA)
            for _ in range(2):
                v, i, _ = g.sweep_dev(xc.data_ptr(), N, _lib.ACQ_POI, float(y.max()))
            out['sweep_ms_' + key] = g.last_sweep_stats()[0]
            best[key] = i
        out['same_argmax'] = bool(best['tables'] == best['routine'])
    finally:
        g.close()
    return out


def measure_dgemm_tflops(torch, dev, n=8192, reps=6):
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = float('inf')
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def sweep_kernel_name(impl, n):
    if impl == 0:
        return 'sweep_simple_kernel'
    if impl in (3, 7) and n <= 64:
        return 'sweep_small_kernel'
    if impl == 1 or (impl in (3, 7) and n < 256):
        return 'sweep_tma_kernel'
    if impl == 2:
        return 'kstar_slices_kernel + sweep_i8_kernel'
    if impl == 4 or (impl in (3, 7) and n < (1024 if 16 <= D <= 32 else 1536)):
        return 'sweep_i8f_kernel'
    if impl == 6:
        return 'sweep_i8p_kernel'
    return 'sweep_i8c_kernel'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--candidates', type=int, default=1 << 24)
    ap.add_argument('--cpu-sample', type=int, default=1 << 18)
    ap.add_argument('--sweep-impl', type=int, default=3,
                    help='3 = library default (cooperative int8-sliced sweep from 1536 observations), 5 = the same forced, 6 = its paired form, '
                         '4 = fused int8-sliced (one CTA per tile), '
                         '2 = int8-sliced as two kernels, 1 = FP64 DMMA, 0 = plain-load cross-check')
    ap.add_argument('--i8-operand', type=int, default=0, help='0 = A planes through tensor memory, 1 = from shared memory')
    ap.add_argument('--skip-latency', action='store_true', help='skip the get_suggestions / C4 / C5 legs')
    ap.add_argument('--skip-cpu', action='store_true', help='skip the CPU baseline leg (profiling runs)')
    args = ap.parse_args()

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))

    if args.impl == 'reference':
        reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    from polyaxon_b200 import _lib
    from polyaxon_b200 import dist as pdist
    from oracle.reference_path import synthetic_problem

    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: polyaxon_b200 has no CPU path')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    X, y, _ = synthetic_problem(N_OBS, D)
    y_max = float(y.max())
    h = _lib.Handle(local_rank)
    h.set_option('sweep_impl', args.sweep_impl)
    h.set_option('i8_operand', args.i8_operand)
    stream = torch.cuda.current_stream()
    h.set_stream(stream.cuda_stream)
    h.set_data(X, y, _lib.KERNEL_MATERN, NU)
    h.set_theta(THETA)

    # this rank's block of the one global candidate matrix, drawn on the device
    N = args.candidates
    lo, hi = pdist.shard_range(N, rank, world)
    n_local = hi - lo
    rng = np.random.RandomState(CAND_SEED)
    zeros, ones = np.zeros(D), np.ones(D)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    drawn = h.candidates_draw(rng, zeros, ones, N, first=lo, count=n_local)
    draw_ms = (time.perf_counter() - t0) * 1e3
    assert drawn, 'device draw refused'
    xc_host = torch.empty(n_local, D, dtype=torch.float64, pin_memory=True)
    h.candidates_get(lo, n_local, D, out=xc_host.numpy())
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def step_device():
        v, i, _ = h.sweep_resident_range(lo, n_local, _lib.ACQ_POI, y_max, eps=0.0)
        return pdist.merge_best(v, i if n_local else -1, device=dev)

    def step_host():
        v, i, _ = h.sweep(xc_host.numpy(), _lib.ACQ_POI, y_max, eps=0.0, idx_base=lo)
        return pdist.merge_best(v, i if n_local else -1, device=dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        launches0 = h.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kernel_ms = 0.0
        e0.record(stream)
        result = None
        for _ in range(steps):
            result = fn()
            kernel_ms += h.last_sweep_stats()[0]
        e1.record(stream)
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1) / steps)
        return ms, kernel_ms / steps, h.launch_count() - launches0, result

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_dev, kernel_ms, launches, best_dev = timed(step_device, args.steps, args.warmup)
    clocks = sampler.stop() if sampler else None
    ms_e2e, _, _, best_host = timed(step_host, args.steps, max(1, min(args.warmup, 1)))
    assert best_dev == best_host, 'device-resident and host-buffer sweeps disagree: %r %r' % (best_dev, best_host)

    # the FP64 DMMA form of the same sweep, a few steps (the 0.9-of-DGEMM kernel of round 1, kept as the arithmetic
    # cross-check of the int8 form: same winner required)
    dmma = None
    if args.sweep_impl != 1:
        h.set_option('sweep_impl', 1)
        dmma_sampler = ClockSampler(local_rank) if rank == 0 else None
        if dmma_sampler:
            dmma_sampler.start()
        ms_dmma, kernel_ms_dmma, _, best_dmma = timed(step_device, min(args.steps, 3), 1)
        dmma_clocks = dmma_sampler.stop() if dmma_sampler else None
        h.set_option('sweep_impl', args.sweep_impl)
        assert best_dmma[1] == best_dev[1], 'FP64 DMMA and int8-sliced sweeps pick different candidates: %r %r' % (best_dmma, best_dev)
        dmma = (ms_dmma, kernel_ms_dmma, dmma_clocks)

    line = None
    if rank == 0:
        value = N / (ms_dev * 1e-3)
        e2e = N / (ms_e2e * 1e-3)
        fpe = flops_per_eval(N_OBS, D)
        fp64_equiv = fpe * n_local / (kernel_ms * 1e-3) / 1e12
        dgemm = measure_dgemm_tflops(torch, dev)
        kernel = sweep_kernel_name(args.sweep_impl, N_OBS)
        sliced = 'i8' in kernel
        traffic_file, traffic_bytes, traffic_cands = TRAFFIC.get(kernel, (None, None, 1))
        roofline = {'bound': 'tensor', 'kernel': kernel, 'kernel_ms_per_launch': kernel_ms,
                    'candidates_per_launch': n_local,
                    'traffic': traffic_bytes / traffic_cands * n_local if traffic_bytes else None,
                    'traffic_note': ('dram__bytes_read.sum + dram__bytes_write.sum of one launch over %d candidates (%.3e '
                                     'bytes, a constant from %s: ncu --set full of this kernel), scaled per candidate to '
                                     'this launch; algorithmic bytes are 256 per candidate, the rest is the digit planes '
                                     'of K* spilling from L2' % (traffic_cands, traffic_bytes, traffic_file)) if traffic_bytes else
                                    'no ncu capture of this kernel form under profiles/',
                    'traffic_candidates': traffic_cands if traffic_bytes else None,
                    'fp64_equivalent_tflops': fp64_equiv, 'flops_per_eval': fpe, 'dgemm_tflops': dgemm,
                    'fp64_equivalent_over_dgemm': fp64_equiv / dgemm}
        if sliced:
            sampler = ClockSampler(local_rank)
            sampler.start()
            i8_peak, t_end = 0.0, time.perf_counter() + 1.5        # long enough for the 200 ms clock sampler
            while time.perf_counter() < t_end:
                i8_peak = max(i8_peak, h.measure_i8_peak())
            peak_clocks = sampler.stop()
            ops = i8_ops_per_eval(N_OBS)
            achieved = ops * n_local / (kernel_ms * 1e-3) / 1e12
            roofline.update({
                'pipe': 'int8 tensor (tcgen05.mma kind::i8, exact digit-plane products)',
                'achieved': achieved, 'peak': i8_peak, 'unit': 'TOP/s', 'frac': achieved / i8_peak,
                'int8_ops_per_eval': ops, 'peak_sm_mhz': peak_clocks['sm_mhz'],
                'peak_source': 'tcgen05.mma kind::i8 m128n128k32 issue rate measured in this run with constant operands '
                               '(plx_measure_i8_peak, best of the repeats within 1.5 s); with random operands the same loop draws more power and '
                               'measures ~4070 TOP/s on this pool (tools/i8_umma_gemm.cu), so this is the upper figure',
                'a_operand': 'shared memory' if args.i8_operand else 'tensor memory (tcgen05.cp ring)'})
            mp = measured_peaks()
            if mp and mp.get('bf16_tflops'):
                # the driver's own figures for this pool: dense bf16 by torch.matmul, burst and sustained; the int8 rate of
                # the same tensor pipe is twice the bf16 rate, so these are the denominators a reader of that file would use
                roofline['vs_measured_peaks'] = {
                    'source': 'MEASURED_PEAKS.json: 2 x bf16_tflops (burst) and 2 x bf16_tflops_sustained (a long step)',
                    'int8_peak_burst_tops': 2.0 * mp['bf16_tflops'], 'frac_of_burst': achieved / (2.0 * mp['bf16_tflops']),
                    'int8_peak_sustained_tops': 2.0 * mp['bf16_tflops_sustained'] if mp.get('bf16_tflops_sustained') else None,
                    'frac_of_sustained': achieved / (2.0 * mp['bf16_tflops_sustained']) if mp.get('bf16_tflops_sustained') else None}
        else:
            roofline.update({'pipe': 'FP64 tensor (DMMA.8x8x4)', 'achieved': fp64_equiv, 'peak': dgemm,
                             'unit': 'TFLOP/s', 'frac': fp64_equiv / dgemm,
                             'peak_source': 'cuBLAS DGEMM 8192^3 FP64 measured in this run (best of 6); '
                                            'MEASURED_PEAKS.json has no FP64 figure'})
        if dmma:
            d_tf = fpe * n_local / (dmma[1] * 1e-3) / 1e12
            roofline.update({'dmma_evals_per_s': N / (dmma[0] * 1e-3), 'dmma_tflops': d_tf, 'dmma_frac': d_tf / dgemm,
                             'dmma_sm_mhz': dmma[2]['sm_mhz'] if dmma[2] else None,
                             'dmma_note': 'sweep_tma_kernel (sweep_impl=1) on the same candidates, %d steps: algorithmic '
                                          'FP64 flops over the DGEMM rate of this run; same arg-max index' % min(args.steps, 3)})
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_dev, 'higher_is_better': True, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(args, world),
            'e2e': {'value': e2e, 'unit': UNIT, 'h2d_bytes_per_step': n_local * D * 8 * world,
                    'd2h_bytes_per_step': 16 * world, 'ms_per_step': ms_e2e,
                    'best_index': best_dev[1], 'best_value': best_dev[0]},
            'gpu_launches': launches,
            'clocks': clocks,
            'roofline': roofline,
            'best_index': best_dev[1], 'best_value': best_dev[0],
            'candidate_draw_ms': draw_ms,
        }
        for k in ('frac', 'peak', 'achieved'):
            if sliced:
                line['int8_' + k if k != 'peak' else 'int8_peak_tops'] = roofline[k]
        if dmma:
            line['dmma_evals_per_s'] = roofline['dmma_evals_per_s']
            line['dmma_frac'] = roofline['dmma_frac']

        # CPU baseline beside it (bounded sample, all host cores) and the cross-check of the winner on that sample
        # (timed at N = 1 only: at N > 1 the other ranks' host threads spin in the barrier and take the cores away)
        if not args.skip_cpu and world == 1:
            try:
                sample = min(args.cpu_sample, n_local)
                base, _, cpu_best = run_cpu_reference(sample, 2, 1)
                line['cpu_baseline'] = base
                assert np.array_equal(xc_host.numpy()[:sample], candidate_rows(sample)), \
                    'device-drawn candidates differ from numpy RandomState rows'
                v, i, _ = h.sweep_resident_range(0, sample, _lib.ACQ_POI, y_max, eps=0.0)
                line['cpu_slice_argmax'] = {'rows': sample, 'cpu_index': cpu_best[0], 'gpu_index': i,
                                            'cpu_value': cpu_best[1], 'gpu_value': v, 'match': bool(i == cpu_best[0])}
                line['cpu_slice_argmax_match'] = bool(i == cpu_best[0])
                assert i == cpu_best[0], 'GPU and CPU reference disagree on the arg-max of the first %d candidates: %r' % (
                    sample, line['cpu_slice_argmax'])
            except AssertionError:
                raise
            except Exception as exc:  # the baseline must never hide the GPU numbers
                line['cpu_baseline'] = {'value': None, 'unit': UNIT, 'cores': os.cpu_count(), 'kind': 'port',
                                        'sample': 'failed: %r' % (exc,)}

    h.set_stream(None)
    if not args.skip_latency:
        # get_suggestions through the drop-in classes: every rank takes part (sharded fit / draw / sweep / refinement)
        try:
            barrier()
            c3 = measure_latency([c3_config()], with_cpu=False)[0]
            barrier()
            if line is not None:
                line['c3_p50_ms'] = c3['get_suggestions_p50_ms']
                line['c3_first_call_ms'] = c3['first_call_ms']
                line['e2e']['c3_p50_ms'] = c3['get_suggestions_p50_ms']
                line['e2e']['c3_first_call_ms'] = c3['first_call_ms']
                line['latency'] = [c3]
        except Exception as exc:
            if line is not None:
                line['latency'] = 'failed: %r' % (exc,)
        try:
            barrier()
            gps, _ = measure_groups()
            t = torch.tensor([gps], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if line is not None:
                line['c5_groups_per_s'] = float(t.item())
                line['e2e']['c5_groups_per_s'] = float(t.item())
                line['groups_c5'] = {'config': 'C5: 64 groups, d=16 n=512 matern2.5 ei N=1e5 S=10 R=5, %d per GPU, 8 worker '
                                               'threads per GPU' % (64 // world), 'groups_per_s': float(t.item())}
        except Exception as exc:
            if line is not None:
                line['groups_c5'] = 'failed: %r' % (exc,)
        if world == 1:
            os.environ['PLX_BO_DISTRIBUTED'] = '0'
            try:
                lat = measure_latency(latency_configs())
                line['latency'] = lat + (line['latency'] if isinstance(line.get('latency'), list) else [])
                line['c1_p50_ms'] = lat[0]['get_suggestions_p50_ms']
                line['c2_p50_ms'] = lat[1]['get_suggestions_p50_ms']
                line['e2e']['c1_p50_ms'] = lat[0]['get_suggestions_p50_ms']
                line['e2e']['c2_p50_ms'] = lat[1]['get_suggestions_p50_ms']
                line['e2e']['c2_first_call_ms'] = lat[1]['first_call_ms']
            except Exception as exc:
                line['latency_c1_c2'] = 'failed: %r' % (exc,)
            try:
                line['fit_c4'] = measure_fit(h)
            except Exception as exc:
                line['fit_c4'] = 'failed: %r' % (exc,)
            try:
                line['general_nu'] = measure_general_nu()
                line['general_nu_sweep_ms'] = line['general_nu']['sweep_ms_tables']
            except Exception as exc:
                line['general_nu'] = 'failed: %r' % (exc,)
    if line is not None:
        print(json.dumps(line))
    h.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
