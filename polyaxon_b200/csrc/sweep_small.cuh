// Warm-up sweep for SMALL observation sets (n <= SS_NMAX = 64): one thread per candidate.
//
// Polyaxon's real groups hold tens of observations.  The tiled sweeps pad the observations to 128 and
// run a 128 x 128 tensor-core tile per 128 candidates whatever n is, behind a TMA ring and several warp roles: at
// n = 15 the 1e5 candidates of config C1 took 0.12 ms of kernel time for ~5 us of arithmetic.  Here the posterior
// state lives in shared memory (scaled observations, alpha, the lower triangle of L^-1: at most 32 KB + n d doubles)
// and thread t of a CTA walks candidate t of a 128-candidate tile through the reference's formulas
// (sklearn _gpr.py:446-500, acquisition_function.py:48-68):
//   k*_j = k(|x/l - X_j/l|^2)  (direct differences, the sweep's kernel functions), mean = sum_j k*_j alpha_j,
//   v_i = sum_{j <= i} Z_ij k*_j, var = 1 - sum_i v_i^2 (clamped at 0 and counted), acquisition, running arg-max.
// k* of the thread's candidate is kept in shared memory (kbuf[j][t]: conflict-free), Z_ij is a broadcast read.
#pragma once
#include "sweep.cuh"

namespace plx {

constexpr int SS_NMAX = 64;
constexpr int SS_THREADS = 128;
constexpr int SS_XMAX = 4096;                      // scaled observations in shared memory up to n * d doubles

inline size_t sweep_small_smem(int n, int d) {
  size_t dbl = static_cast<size_t>(n) * SS_THREADS + static_cast<size_t>(n) * n + n + 2 * SS_THREADS;
  if (n * d <= SS_XMAX) dbl += static_cast<size_t>(n) * d;
  return dbl * sizeof(double);
}

template <int KIND>
__device__ __forceinline__ void sweep_small_body(const SweepParams& p, double* smem) {
  const int n = p.n, d = p.d, tid = threadIdx.x;
  double* kbuf = smem;                              // [n][128]
  double* Zs = kbuf + n * SS_THREADS;               // [n][n], lower triangle used
  double* al = Zs + n * n;                          // [n]
  double* sval = al + n;                            // [128]
  long long* sidx = reinterpret_cast<long long*>(sval + SS_THREADS);
  const bool xs_smem = n * d <= SS_XMAX;
  double* Xs = reinterpret_cast<double*>(sidx + SS_THREADS);      // [n][d] = X / l
  for (int e = tid; e < n * n; e += SS_THREADS) {
    const int i = e / n, j = e - i * n;
    Zs[e] = (j <= i) ? p.Z[static_cast<size_t>(i) * p.npad + j] : 0.0;
  }
  for (int j = tid; j < n; j += SS_THREADS) al[j] = p.alpha[j];
  if (xs_smem)
    for (int e = tid; e < n * d; e += SS_THREADS) Xs[e] = p.X[e] / p.ell;
  __syncthreads();
  Best best = {-INFINITY, 0x7fffffffffffffffLL};
  const long long ntiles = (p.N + SS_THREADS - 1) / SS_THREADS;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long ci = tile * SS_THREADS + tid;
    if (ci >= p.N) continue;
    const double* x = p.Xc + ci * d;
    double mean = 0.0;
    for (int j = 0; j < n; ++j) {
      double r2 = 0.0;
      if (xs_smem) {
        const double* xj = Xs + j * d;
        for (int q = 0; q < d; ++q) { const double df = x[q] / p.ell - xj[q]; r2 += df * df; }
      } else {
        const double* xj = p.X + static_cast<size_t>(j) * d;
        for (int q = 0; q < d; ++q) { const double df = x[q] / p.ell - xj[q] / p.ell; r2 += df * df; }
      }
      const double k = kern_value_t<KIND>(r2, p.mg);
      kbuf[j * SS_THREADS + tid] = k;
      mean += k * al[j];
    }
    double sumsq = 0.0;
    for (int i = 0; i < n; ++i) {
      const double* Zi = Zs + i * n;
      double v0 = 0.0, v1 = 0.0;
      int j = 0;
      for (; j + 1 <= i; j += 2) {
        v0 += Zi[j] * kbuf[j * SS_THREADS + tid];
        v1 += Zi[j + 1] * kbuf[(j + 1) * SS_THREADS + tid];
      }
      if (j <= i) v0 += Zi[j] * kbuf[j * SS_THREADS + tid];
      const double v = v0 + v1;
      sumsq = fma(v, v, sumsq);
    }
    double var = 1.0 - sumsq;
    if (var < 0.0) { var = 0.0; if (p.clamped) atomicAdd(p.clamped, 1ULL); }   // sklearn clamps negative variances to zero, and says so
    const double sd = sqrt(var);
    const double a = acq_value(p.acq, mean, sd, p.y_max, p.kappa, p.eps);
    if (p.out_mean) p.out_mean[ci] = mean;
    if (p.out_std) p.out_std[ci] = sd;
    if (p.out_acq) p.out_acq[ci] = a;
    const long long gi = p.idx_base + ci;
    if (better(a, gi, best.v, best.i)) { best.v = a; best.i = gi; }
  }
  sval[tid] = best.v; sidx[tid] = best.i;
  __syncthreads();
  if (tid == 0) {
    Best b = {sval[0], sidx[0]};
    for (int t = 1; t < SS_THREADS; ++t)
      if (better(sval[t], sidx[t], b.v, b.i)) { b.v = sval[t]; b.i = sidx[t]; }
    p.blk_val[blockIdx.x] = b.v;
    p.blk_idx[blockIdx.x] = b.i;
  }
}

__global__ void __launch_bounds__(SS_THREADS) sweep_small_kernel(const SweepParams p) {
  extern __shared__ double ss_smem[];
  switch (p.kind) {
    case K_RBF: sweep_small_body<K_RBF>(p, ss_smem); break;
    case K_M05: sweep_small_body<K_M05>(p, ss_smem); break;
    case K_M15: sweep_small_body<K_M15>(p, ss_smem); break;
    case K_M25: sweep_small_body<K_M25>(p, ss_smem); break;
    default: sweep_small_body<K_GEN>(p, ss_smem); break;
  }
}

}  // namespace plx
