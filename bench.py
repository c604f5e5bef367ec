#!/usr/bin/env python
"""Benchmark of the B200-native Bayesian-optimisation suggestion path.

Metric (BASELINE.json): acquisition evaluations per second (FP64) at d=32, n_obs=2048, plus
get_suggestions p50 latency.  A "step" is one warm-up sweep (acquisition_function.py:88-92) over the
whole candidate set of config C3 (16M candidates, Matern nu=1.5, POI eps=0) at a fixed GP state;
candidates are sharded over the ranks by contiguous index range and the per-rank bests are merged with
one all-gather (strong scaling: the global candidate count is fixed).

  value   whole-job candidates/s with candidates resident in HBM (device-timed, max over ranks)
  e2e     the same through the C-ABI call with HOST (pinned) candidate buffers: host->device copies and
          the device->host read of the result are inside the timed region
  roofline  tensor bound; achieved = algorithmic FP64 flops per candidate x candidates / duration of the sweep
          kernels (CUDA events on their own stream); peak = cuBLAS DGEMM 8192^3 measured in this run
          (MEASURED_PEAKS.json has no FP64 entry).  On the default path the variance contraction runs as exact
          int8 slices on the tcgen05 tensor cores, so the fraction of the FP64 rate can exceed 1; the `int8`
          sub-object gives the executed int8 tensor ops against the instruction's measured issue rate
  cpu_baseline  the reference's sklearn/scipy path (oracle/reference_path.py) on the host cores, on a
          bounded sample of the same workload

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
METRIC = 'acquisition evals/sec (FP64) at d=32,n_obs=2048'
# dram__bytes_read.sum + dram__bytes_write.sum per candidate: sweep_tma_kernel (profiles/r01_ncu_sweep_tma_v3_summary.csv)
# and kstar_slices_kernel + sweep_i8_kernel (profiles/r01_ncu_sweep_i8_summary.csv, one 37 888-candidate chunk)
DRAM_BYTES_PER_CANDIDATE = (18.701e9 + 2.490e9) / 151552
DRAM_BYTES_PER_CANDIDATE_I8 = (9.846e9 + 0.006e9 + 0.011e9 + 0.567e9) / 37888
DRAM_BYTES_PER_CANDIDATE_I8C = (2.622e9 + 2.196e9) / 151552       # sweep_i8c_kernel, same csv
UNIT = 'evals/s'


def flops_per_eval(n, d):
    """SURVEY.md section 8d / BASELINE.md: n(3d+10) + n^2 + 4n (triangular variance)."""
    return n * (3 * d + 10) + n * n + 4 * n


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
        sm, reasons, sm_max = [], set(), None
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                sm_max = float(r[1])
                for name, flag in zip(names, r[3:7]):
                    if flag.lower().startswith('active'):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {'sm_mhz': float(np.median(busy)) if busy else None, 'sm_max_mhz': sm_max,
                'reasons': sorted(reasons), 'samples': len(sm)}


def reference_arm(args, rank):
    """The reference's CPU implementation (sklearn predict + scipy norm) on the host cores."""
    if rank != 0:
        return
    import warnings
    from oracle import gp_numpy as G
    from oracle import reference_path as R
    X, y, bounds = R.synthetic_problem(N_OBS, D)
    gp = R.make_gp(G.KERNEL_MATERN, float(np.exp(THETA)), NU, 0, np.random.RandomState(1))
    gp.optimizer = None                    # fixed GP state, as in our arm
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        gp.fit(X, y)
    y_max = float(y.max())
    sample = args.cpu_sample
    rs = np.random.RandomState(7)
    xc = rs.uniform(0, 1, (sample, D))

    def step():
        ys = R.acquisition(gp, xc, G.ACQ_POI, y_max, 0.0, 0.0)
        return int(ys.argmax())

    # torchrun exports OMP_NUM_THREADS=1; the baseline is entitled to every host core
    from threadpoolctl import threadpool_limits
    cores = os.cpu_count()
    with threadpool_limits(limits=cores):
        for _ in range(min(args.warmup, 1)):
            step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = (time.perf_counter() - t0) / args.steps
    value = sample / dt
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': min(args.warmup, 1), 'ms_per_step': dt * 1e3,
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': workload_config(args, 1),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': '%d candidates per step (one predict call), sklearn %s / scipy path of '
                                   'oracle/reference_path.py' % (sample, __import__('sklearn').__version__)},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


def latency_configs():
    """C1 and C2 of BASELINE.json as hptuning sections + observations (for get_suggestions latency)."""
    from oracle.reference_path import synthetic_problem
    out = []
    X, y, _ = synthetic_problem(15, 2)
    out.append(('C1 d=2 n=15 matern2.5 ucb N=1e5 S=10',
                dict(kernel='matern', nu=2.5, acq='ucb', kappa=2.576, eps=None, n_warmup=100000, n_iter=10,
                     restarts=0), X, y))
    X, y, _ = synthetic_problem(256, 8)
    out.append(('C2 d=8 n=256 rbf ei N=1e6 S=10',
                dict(kernel='rbf', nu=1.5, acq='ei', kappa=None, eps=0.0, n_warmup=1000000, n_iter=10,
                     restarts=0), X, y))
    X, y, _ = synthetic_problem(2048, 32)
    out.append(('C3 d=32 n=2048 matern1.5 poi N=16M S=10 R=17',
                dict(kernel='matern', nu=1.5, acq='poi', kappa=None, eps=0.0, n_warmup=1 << 24, n_iter=10,
                     restarts=16, reps=2, cpu=False), X, y))
    return out


def measure_latency(reps_gpu=7):
    """p50 of BOOptimizer.get_suggestion() through the drop-in Python classes, and the same call on
    the CPU reference path (oracle/reference_path.py), same seed and observations."""
    import warnings
    from oracle import gp_numpy as G
    from oracle import reference_path as R
    from polyaxon_b200.hpsearch.search_managers.bayesian_optimization.optimizer import BOOptimizer
    from polyaxon_b200.schemas.hptuning import HPTuningConfig
    from threadpoolctl import threadpool_limits
    results = []
    for label, c, X, y in latency_configs():
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
        configs = [dict(zip(keys, row.tolist())) for row in X]
        times = []
        for _ in range(c.get('reps', reps_gpu) + 1):
            t0 = time.perf_counter()
            opt = BOOptimizer(hptuning_config=cfg)
            opt.add_observations(configs=configs, metrics=y.tolist())
            suggestion = opt.get_suggestion()
            times.append((time.perf_counter() - t0) * 1e3)
        gpu_ms = float(np.median(times[1:]))
        first_ms = float(times[0])
        bounds = np.tile(np.array([[0.0, 1.0]]), (d, 1))
        cpu = []
        with threadpool_limits(limits=os.cpu_count()), warnings.catch_warnings():
            warnings.simplefilter('ignore')
            for _ in range(0 if not c.get('cpu', True) else (2 if c['n_warmup'] <= 100000 else 1)):
                t0 = time.perf_counter()
                R.suggest(X, y, bounds, G.KERNEL_RBF if c['kernel'] == 'rbf' else G.KERNEL_MATERN, 1.0, c['nu'],
                          c['restarts'], {'ucb': 0, 'ei': 1, 'poi': 2}[c['acq']], c['kappa'] or 0.0, c['eps'] or 0.0,
                          17, c['n_warmup'], c['n_iter'], chunk=65536)
                cpu.append((time.perf_counter() - t0) * 1e3)
        results.append({'config': label, 'get_suggestions_p50_ms': gpu_ms, 'first_call_ms': first_ms,
                        'cpu_reference_ms': float(np.median(cpu)) if cpu else None,
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


def measure_groups(n_groups=8):
    """Config C5, one GPU's share: 8 of the 64 independent groups (d=16, n_obs=512) through
    polyaxon_b200.groups.suggest_groups, one worker thread against eight (a C-ABI handle and stream each)."""
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
    out = {'config': 'C5 share of one GPU: %d groups, d=16 n=512 matern2.5 ei N=1e5 S=10 R=5' % n_groups}
    first = None
    for workers in (1, 8):
        times = []
        for _ in range(4):
            its = [BOIterationConfig.from_dict(r) for r in raw]
            t0 = time.perf_counter()
            got = suggest_groups(cfgs, its, workers=workers)
            times.append(time.perf_counter() - t0)
        first = got if first is None else first
        out['groups_per_s_workers%d' % workers] = n_groups / float(np.median(times[1:]))
        out['same_suggestions'] = bool(got == first)
    return out


def workload_config(args, world):
    return {'workload': 'C3: BO 32-dim continuous, POI eps=0, matern nu=1.5, n_obs=2048, '
                        'num_warmup=%d candidates (global), fixed GP state theta=%.1f' % (args.candidates, THETA),
            'candidates': args.candidates, 'd': D, 'n_obs': N_OBS, 'parallelism': 'shard%d' % world,
            'l2': 'inputs larger than L2 (candidate shard >= 0.5 GB, K* slabs 310 MB)'}


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--candidates', type=int, default=1 << 24)
    ap.add_argument('--cpu-sample', type=int, default=1 << 15)
    ap.add_argument('--sweep-impl', type=int, default=3,
                    help='3 = library default (cooperative int8-sliced sweep from 256 observations), 5 = the same forced, '
                         '4 = fused int8-sliced (one CTA per tile), '
                         '2 = int8-sliced as two kernels, 1 = FP64 DMMA, 0 = plain-load cross-check')
    ap.add_argument('--skip-latency', action='store_true')
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
    stream = torch.cuda.current_stream()
    h.set_stream(stream.cuda_stream)
    h.set_data(X, y, _lib.KERNEL_MATERN, NU)
    h.set_theta(THETA)

    lo, hi = pdist.shard_range(args.candidates, rank, world)
    n_local = hi - lo
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    xc_dev = torch.rand(n_local, D, dtype=torch.float64, device=dev, generator=gen)
    xc_host = torch.empty(n_local, D, dtype=torch.float64, pin_memory=True)
    xc_host.copy_(xc_dev)
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
        v, i, _ = h.sweep_dev(xc_dev.data_ptr(), n_local, _lib.ACQ_POI, y_max, eps=0.0, idx_base=lo)
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

    if rank == 0:
        value = args.candidates / (ms_dev * 1e-3)
        e2e = args.candidates / (ms_e2e * 1e-3)
        fpe = flops_per_eval(N_OBS, D)
        achieved = fpe * n_local / (kernel_ms * 1e-3) / 1e12
        peak = measure_dgemm_tflops(torch, dev)
        sliced = args.sweep_impl in (2, 4, 5) or (args.sweep_impl == 3 and N_OBS >= 256)
        if sliced:
            nh = (N_OBS + 63) // 64                          # 64-row half-blocks of L^-1, triangular in 64-wide chunks
            i8_ops = 28 * 2.0 * 64 * 64 * nh * (nh + 1) / 2   # 28 digit-plane pairs, executed tensor ops per candidate
            i8_peak = h.measure_i8_peak()
            i8_achieved = i8_ops * n_local / (kernel_ms * 1e-3) / 1e12
            roof_kernel = ('kstar_slices_kernel + sweep_i8_kernel (alternating, chunks of %d candidates)' % (148 * 256)
                           if args.sweep_impl == 2 else 'sweep_i8f_kernel (digit-plane builders + int8 contraction fused)'
                           if args.sweep_impl == 4 or (args.sweep_impl == 3 and N_OBS < 1536) else
                           'sweep_i8c_kernel (builders + int8 contraction fused, four CTAs per candidate tile)')
            roof_extra = {'int8': {'executed_tops': i8_achieved, 'peak_tops': i8_peak, 'frac': i8_achieved / i8_peak,
                                   'ops_per_eval': i8_ops,
                                   'peak_source': 'tcgen05.mma kind::i8 m128n128k32 issue rate measured in this run '
                                                  '(plx_measure_i8_peak); the sweep uses n64 accumulators, whose '
                                                  'measured ceiling is 0.69 of it (tools/i8_umma_n64.cu)'},
                          'note': 'frac is algorithmic FP64 flops over the measured DGEMM rate: the contraction runs as '
                                  'exact 7-bit slices on the int8 tensor cores, not on the FP64 pipe'}
            coop = args.sweep_impl == 5 or (args.sweep_impl == 3 and N_OBS >= 1536)
            traffic = (DRAM_BYTES_PER_CANDIDATE_I8C if coop else DRAM_BYTES_PER_CANDIDATE_I8) * n_local
            traffic_note = ('DRAM bytes of the two kernels from profiles/ (ncu, one chunk), scaled per candidate: digit '
                            'planes of K* written once and re-read per 64-row half-block of L^-1')
        else:
            roof_kernel = 'sweep_tma_kernel' if args.sweep_impl in (1, 3) else 'sweep_simple_kernel'
            roof_extra = {}
            traffic = DRAM_BYTES_PER_CANDIDATE * n_local
            traffic_note = ('DRAM bytes of this kernel from profiles/ (ncu --set full at 151552 candidates), scaled per '
                            'candidate: the K* slab round trip, by design')
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_dev, 'higher_is_better': True, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(args, world),
            'e2e': {'value': e2e, 'unit': UNIT, 'h2d_bytes_per_step': n_local * D * 8 * world,
                    'd2h_bytes_per_step': 16 * world, 'ms_per_step': ms_e2e},
            'gpu_launches': launches,
            'clocks': clocks,
            'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak, 'unit': 'TFLOP/s',
                         'frac': achieved / peak, 'traffic': traffic,
                         'kernel': roof_kernel,
                         'traffic_note': traffic_note,
                         'kernel_ms_per_launch': kernel_ms,
                         'flops_per_eval': fpe,
                         'peak_source': 'cuBLAS DGEMM 8192^3 FP64 measured in this run (best of 6); '
                                        'MEASURED_PEAKS.json has no FP64 figure', **roof_extra},
            'best': {'value': best_dev[0], 'index': best_dev[1]},
        }
        # CPU baseline beside it: bounded sample, all host cores
        try:
            import io
            import contextlib
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                ref_args = argparse.Namespace(**vars(args))
                ref_args.steps, ref_args.warmup = 2, 1
                reference_arm(ref_args, 0)
            ref = json.loads(buf.getvalue().strip().splitlines()[-1])
            line['cpu_baseline'] = ref['cpu_baseline']
        except Exception as exc:  # the baseline must never hide the GPU numbers
            line['cpu_baseline'] = {'value': None, 'unit': UNIT, 'cores': os.cpu_count(), 'kind': 'port',
                                    'sample': 'failed: %r' % (exc,)}
        if world == 1 and not args.skip_latency:
            try:
                h.set_stream(None)
                line['latency'] = measure_latency()
            except Exception as exc:
                line['latency'] = 'failed: %r' % (exc,)
            try:
                line['fit_c4'] = measure_fit(h)
            except Exception as exc:
                line['fit_c4'] = 'failed: %r' % (exc,)
            try:
                line['groups_c5'] = measure_groups()
            except Exception as exc:
                line['groups_c5'] = 'failed: %r' % (exc,)
        print(json.dumps(line))
    h.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
