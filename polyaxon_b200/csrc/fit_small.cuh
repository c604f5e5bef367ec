// Hyper-parameter fit for SMALL observation sets (n <= FS_NUSE = 64) in ONE launch: CTA r runs optimiser run r
// (sklearn: the configured length scale first, then the restarts; _gpr.py:299-337) from start to convergence --
// every log-marginal-likelihood + gradient evaluation AND the L-BFGS-B / More'-Thuente state machine
// (lbfgsb1d.h, the same code the host drives for large n) on the device, with no launch and no device->host
// round trip between evaluations.
//
// Why: Polyaxon's real groups hold tens of observations.  There the batched launch sequence of the
// general path (gp_kernels.cuh: ~9 launches and one synchronising copy per evaluation) is pure latency: config C1
// (n = 15) spent 1.1 ms in 9 evaluations of a few microseconds of arithmetic each.
//
// One evaluation at theta, everything in shared memory (one n x (n+1) matrix M):
//   1. Xs = X / exp(theta) (global scratch); K into the lower triangle of M (diagonal 1 + jitter), dK/dtheta
//      into the strict upper triangle (its diagonal is zero for every kernel here).  Same arithmetic as
//      kmat_kernel: direct differences, library exp / sqrt.
//   2. Right-looking factorisation fused with the inversion of the factor: at step k the pivot column updates the
//      trailing matrix (A[i][j] -= A[i][k] A[j][k] / d_k, j > k) and, in the same pass, the rows of the growing inverse
//      (R[i][j] -= Lu[i][k] R[k][j], j <= k).  L is never needed again, so L^-1 overwrites it in place; log det comes
//      from the pivots.
//   3. w = L^-1 y, alpha = L^-T w, LML = -1/2 y.alpha - sum log l_kk - n/2 log 2 pi  (_gpr.py:583-654).
//   4. dLML/dtheta = sum_{i > j} (alpha_i alpha_j - Kinv_ij) dK_ij with Kinv_ij = sum_{m >= i} Z_mi Z_mj  (:631-633).
// A non-positive pivot gives LML = -inf, gradient 0, as sklearn does (:592-593).
#pragma once
#include "gp_kernels.cuh"
#include "lbfgsb1d.h"
#ifdef PLX_PROFILE
#include <cstdio>
#endif

namespace plx {

constexpr int FS_NMAX = 128;                   // what the kernel can hold
constexpr int FS_NUSE = 64;                    // what it is used for: measured against the batched launch sequence
                                               // (33 runs excluded) 0.33 / 0.59 / 1.4 / 3.6 / 5.1 ms at n = 16 / 32 / 64 /
                                               // 96 / 128 against 1.9 / 2.0 / 2.2 / 2.9 / 2.5 ms: the 128 dependent steps of
                                               // the in-CTA factorisation cost more than the launches they save above ~80
constexpr int FS_THREADS = 512;
constexpr int FS_XS = 8192;                    // scaled observations stay in shared memory up to n * d = 8192 doubles
constexpr int FS_SMEM = (FS_NMAX * (FS_NMAX + 1) + 8 * FS_NMAX + 64 + FS_XS) * 8;

struct FitSmallParams {
  const double* X; const double* y;      // n x d, n
  int n, d, kind, R;
  MaternGen mg;
  double jitter, lo, hi;
  const double* starts;                  // [R]
  double* xs;                            // scratch [R][n * d]
  double* per_run;                       // [R][4]: start, theta*, LML*, evaluations
};

__device__ __forceinline__ double fs_block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < FS_THREADS / 32; ++w) t += red[w];
  return t;
}

// LML and its theta-derivative at theta; all threads of the CTA call this and all get the result
// (L_out: when not null, the factor L is also written there, leading dimension ldl, and the gradient is skipped)
__device__ void fs_eval(const FitSmallParams& p, double theta, double* M, double* vec, double* xs, double& lml,
                        double& grad, double* L_out = nullptr, int ldl = 0) {
  const int n = p.n, d = p.d, ld = n + 1, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* colA = vec;                 // [2][NMAX] pivot column of the step (double-buffered)
  double* rowR = vec + 2 * FS_NMAX;   // [2][NMAX] finished row of the unit inverse
  double* dl = vec + 4 * FS_NMAX;     // pivots d_k
  double* sk = vec + 5 * FS_NMAX;     // d_k^-1/2
  double* wv = vec + 6 * FS_NMAX;     // w = L^-1 y
  double* al = vec + 7 * FS_NMAX;     // alpha
  double* red = vec + 8 * FS_NMAX;    // reductions (16 doubles)
  double* pv = red + 32;              // [2] pivot of the step
#ifdef PLX_PROFILE
  long long t0 = clock64(), t1, t2, t3, t4;
#endif
  const double ell = exp(theta);
  for (int e = tid; e < n * d; e += FS_THREADS) xs[e] = p.X[e] / ell;
  __syncthreads();
  // 1. K (lower, with diagonal) and dK/dtheta (strict upper)
  const int npairs = n * (n + 1) / 2;
  for (int q = tid; q < npairs; q += FS_THREADS) {
    int i = static_cast<int>((sqrt(8.0 * q + 1.0) - 1.0) * 0.5);
    while ((i + 1) * (i + 2) / 2 <= q) ++i;
    while (i * (i + 1) / 2 > q) --i;
    const int j = q - i * (i + 1) / 2;
    if (i == j) { M[i * ld + i] = 1.0 + p.jitter; continue; }
    const double* xi = xs + static_cast<size_t>(i) * d;
    const double* xj = xs + static_cast<size_t>(j) * d;
    double r2 = 0.0;
    for (int c = 0; c < d; ++c) { const double df = xi[c] - xj[c]; r2 += df * df; }
    double kv, dk;
    if (p.kind == K_GEN) { double bb; matern_gen_eval(p.mg, r2, kv, dk, bb); }
    else { kv = kern_value(p.kind, r2, p.mg); dk = kern_dtheta(p.kind, r2, p.mg); }
    M[i * ld + j] = kv;
    M[j * ld + i] = dk;
  }
  __syncthreads();
#ifdef PLX_PROFILE
  t1 = clock64();
#endif
  // 2. Factorisation fused with the inversion of the factor, as K = Lu D Lu^T (Lu unit lower, D the pivots) so
  //    that no square root and no division sits on the step-to-step critical path: L = Lu sqrt(D), and
  //    L^-1 = D^-1/2 Lu^-1 is scaled once at the end.  The pivot column, the finished row of Lu^-1 and the pivot
  //    of step k + 1 are published by the threads that produce them during step k (double-buffered), which
  //    leaves one barrier per step.  Scalars that every thread needs cost one reciprocal each, not a log / sqrt /
  //    divide per thread: the FP64 pipe (64 lanes per clock and SM) is what a 512-thread CTA runs out of first.
  for (int i = 1 + tid; i < n; i += FS_THREADS) colA[i] = M[i * ld];
  if (tid == 0) pv[0] = M[0];
  __syncthreads();
  bool bad = false;
  for (int k = 0; k < n; ++k) {
    const int b = k & 1;
    const double piv = pv[b];
    if (!(piv > 0.0)) { bad = true; break; }             // the same value in every thread
    const double rd = __drcp_rn(piv);
    const double* ca = colA + b * FS_NMAX;
    const double* rr = rowR + b * FS_NMAX;
    double* ca_next = colA + (b ^ 1) * FS_NMAX;
    double* rr_next = rowR + (b ^ 1) * FS_NMAX;
    if (tid == 0) dl[k] = piv;
    if (L_out != nullptr) {                              // column k of L = Lu[.][k] sqrt(d_k)
      const double rt = sqrt(piv);
      for (int i = k + tid; i < n; i += FS_THREADS) L_out[static_cast<size_t>(i) * ldl + k] = (i == k) ? rt : ca[i] * rd * rt;
    }
    for (int i = k + 1 + warp; i < n; i += FS_THREADS / 32) {
      const double li = ca[i] * rd;                      // Lu[i][k]
      double* Mi = M + i * ld;
      const bool next_row = (i == k + 1);
      for (int j = lane; j <= i; j += 32) {
        double v;
        if (j > k) {                                     // trailing matrix
          v = Mi[j] - li * ca[j];
          if (j == k + 1) { ca_next[i] = v; if (next_row) pv[b ^ 1] = v; }
        } else if (j == k) {
          v = -li;                                       // Lu^-1 [i][k]: the unit diagonal of row k times -Lu[i][k]
        } else {
          v = Mi[j] - li * rr[j];
        }
        Mi[j] = v;
        if (next_row && j <= k) rr_next[j] = v;          // row k + 1 of Lu^-1 is complete after this step
      }
    }
    __syncthreads();
  }
  if (bad) { lml = -INFINITY; grad = 0.0; __syncthreads(); return; }
  // L^-1 = D^-1/2 Lu^-1 and log det L = 1/2 sum log d_k, all rows at once
  double ld_part = 0.0;
  for (int k = tid; k < n; k += FS_THREADS) { sk[k] = 1.0 / sqrt(dl[k]); ld_part += log(dl[k]); }
  const double logdet = 0.5 * fs_block_sum(ld_part, red);
  for (int k = warp; k < n; k += FS_THREADS / 32) {
    const double sc = sk[k];
    for (int j = lane; j <= k; j += 32) M[k * ld + j] = (j == k) ? sc : M[k * ld + j] * sc;
  }
  __syncthreads();
#ifdef PLX_PROFILE
  t2 = clock64();
#endif
  // 3. w = Z y (a warp per row), alpha = Z^T w (a thread per column)
  for (int i = warp; i < n; i += FS_THREADS / 32) {
    double s = 0.0;
    for (int j = lane; j <= i; j += 32) s += M[i * ld + j] * p.y[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) wv[i] = s;
  }
  __syncthreads();
  for (int j = tid; j < n; j += FS_THREADS) {
    double s = 0.0;
    for (int i = j; i < n; ++i) s += M[i * ld + j] * wv[i];
    al[j] = s;
  }
  __syncthreads();
  double ya = 0.0;
  for (int i = tid; i < n; i += FS_THREADS) ya += p.y[i] * al[i];
  ya = fs_block_sum(ya, red);
  lml = -0.5 * ya - logdet - 0.5 * n * 1.8378770664093453;          // log(2 pi)
#ifdef PLX_PROFILE
  t3 = clock64();
#endif
  if (L_out != nullptr) { grad = 0.0; return; }
  // 4. gradient: a warp per row i, lanes over j < i, Kinv_ij as a dot product of columns i and j of Z
  double gs = 0.0;
  for (int i = 1 + warp; i < n; i += FS_THREADS / 32) {
    const double ai = al[i];
    for (int j = lane; j < i; j += 32) {
      double kinv = 0.0;
      for (int m = i; m < n; ++m) kinv += M[m * ld + i] * M[m * ld + j];
      gs += (ai * al[j] - kinv) * M[j * ld + i];
    }
  }
  grad = fs_block_sum(gs, red);
  __syncthreads();
#ifdef PLX_PROFILE
  t4 = clock64();
  if (tid == 0 && blockIdx.x == 0)
    printf("fs_eval n=%d: build %lld factor %lld solve %lld grad %lld clk\n", n, t1 - t0, t2 - t1, t3 - t2, t4 - t3);
#endif
}

// One CTA per optimiser run.  Thread 0 owns the L-BFGS-B state machine; the trial point travels through shared memory.
__global__ void __launch_bounds__(FS_THREADS, 1) fit_small_kernel(const FitSmallParams p) {
  extern __shared__ double fs_smem[];
  double* M = fs_smem;
  double* vec = fs_smem + FS_NMAX * (FS_NMAX + 1);
  double* ctl = vec + 8 * FS_NMAX + 48;                  // [0] trial theta, [1] running flag
  const int r = blockIdx.x;
  double* xs = (p.n * p.d <= FS_XS) ? vec + 8 * FS_NMAX + 64 : p.xs + static_cast<size_t>(r) * p.n * p.d;
  Lbfgsb1D run;                                          // meaningful in thread 0 only
  if (threadIdx.x == 0) {
    run.lo = p.lo; run.hi = p.hi;
    run.start(p.starts[r]);
    ctl[0] = run.x_eval; ctl[1] = 1.0;
  }
  __syncthreads();
  for (;;) {
    if (ctl[1] == 0.0) break;
    const double theta = ctl[0];
    double lml, grad;
    fs_eval(p, theta, M, vec, xs, lml, grad);
    if (threadIdx.x == 0) {
      run.feed(-lml, -grad);                             // minimise -LML
      ctl[0] = run.x_eval; ctl[1] = run.done ? 0.0 : 1.0;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    p.per_run[4 * r + 0] = p.starts[r]; p.per_run[4 * r + 1] = run.x;
    p.per_run[4 * r + 2] = -run.f; p.per_run[4 * r + 3] = run.nfev;
  }
}

// Posterior state at one theta for a small observation set (what plx_gp_set_theta produces): L, Z = L^-1 and alpha,
// zero-padded to npad (identity on the padded diagonal of L), in one launch of one CTA.  status[0] = 0 / 1 (not PD).
__global__ void __launch_bounds__(FS_THREADS, 1) state_small_kernel(const FitSmallParams p, double theta, int npad,
                                                                    double* __restrict__ L_out, double* __restrict__ Z_out,
                                                                    double* __restrict__ alpha_out, double* __restrict__ status) {
  extern __shared__ double fs_smem[];
  double* M = fs_smem;
  double* vec = fs_smem + FS_NMAX * (FS_NMAX + 1);
  double* xs = (p.n * p.d <= FS_XS) ? vec + 8 * FS_NMAX + 64 : p.xs;
  const int n = p.n, ld = n + 1, tid = threadIdx.x;
  for (int e = tid; e < npad * npad; e += FS_THREADS) {          // padding of L; its computed part is written in fs_eval
    const int i = e / npad, j = e - i * npad;
    if (i >= n || j > i) L_out[e] = (i == j) ? 1.0 : 0.0;
  }
  double lml, grad;
  fs_eval(p, theta, M, vec, xs, lml, grad, L_out, npad);
  const bool bad = !(lml > -INFINITY);
  if (tid == 0) { status[0] = bad ? 1.0 : 0.0; status[1] = lml; }
  if (bad) return;
  const double* al = vec + 7 * FS_NMAX;
  for (int e = tid; e < npad * npad; e += FS_THREADS) {
    const int i = e / npad, j = e - i * npad;
    Z_out[e] = (i < n && j <= i) ? M[i * ld + j] : 0.0;
  }
  for (int j = tid; j < npad; j += FS_THREADS) alpha_out[j] = (j < n) ? al[j] : 0.0;
}

}  // namespace plx
