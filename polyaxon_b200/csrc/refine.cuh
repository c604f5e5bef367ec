// Refinement stage of max_compute: bounded quasi-Newton ascent of the acquisition function from
// every seed at once, entirely on the device.
//
// Reference path being replaced: acquisition_function.py:95-111 -- for each of n_iter seeds,
// scipy.optimize.minimize(lambda x: -acq(x), x0, bounds, method="L-BFGS-B") with finite-difference
// gradients (33 predict() calls per gradient at d = 32).  Here the gradient is analytic and all
// seeds advance together:
//
//   per round   refine_kstar_kernel   k*_j, dk*/d(r^2)_j for every seed           (n d flops / seed)
//               refine_v_kernel       v = Z k*     Z = L^-1 read once for ALL seeds (n^2 / 2 / seed)
//               refine_w_kernel       w = Z^T v    (K^-1 k*, for the variance gradient)
//               refine_step_kernel    mean, variance, acquisition and gradient, then one transition of
//                                     the seed's optimiser state machine (warp-cooperative vectors)
//
// The optimiser is a projected limited-memory BFGS in the L-BFGS-B family: active bounds are those
// whose gradient points outward, the two-loop recursion (m = 10) gives the direction on the free
// variables, the step is capped at the first bound it would cross (stpmx, as L-BFGS-B's lnsrlb does)
// and chosen by the same More'-Thuente search (dcsrch: ftol 1e-3, gtol 0.9, xtol 0.1, at most 20
// evaluations) with scipy's stopping rules (projected gradient <= 1e-5, relative reduction
// <= 1e7 eps, 15000 iterations).  It is not scipy's trajectory -- analytic gradients alone rule
// that out -- the contract (BASELINE.json) is the value reached: >= scipy's - 1e-6.
#pragma once
#include "sweep.cuh"
#include "lbfgsb1d.h"
#ifdef PLX_PROFILE
#include <cstdio>
#endif

namespace plx {

constexpr int RF_M = 10;          // stored correction pairs
constexpr int RF_DMAX = 2048;     // largest encoded dimension the refinement supports (per-seed gradient in shared memory: 16 d bytes)
constexpr int RF_SB = 8;          // seeds per register block in the matrix-vector kernels

// d k / d (r^2) for the closed-form kernels (r^2 = squared distance in length-scale units)
__device__ __forceinline__ double kern_dr2(int kind, double r2, const MaternGen& mg) {
  if (kind == K_GEN) { double v, a, b; matern_gen_eval(mg, r2, v, a, b); return b; }
  if (kind == K_RBF) return -0.5 * exp(-0.5 * r2);
  const double r = sqrt(r2);
  if (kind == K_M05) return (r > 0.0) ? -exp(-r) / (2.0 * r) : 0.0;
  if (kind == K_M15) return -1.5 * exp(-r * 1.7320508075688772);
  const double t = r * 2.23606797749979;
  return -(5.0 / 6.0) * (1.0 + t) * exp(-t);
}

struct RefineScalars {            // per-seed optimiser state (all lanes of the warp hold a copy)
  double f, ft, stp, stpmx, gd0, gamma;
  Dcsrch ls;
  int iter, nfev, ifun, hist_n, hist_head, phase, status;   // phase 0 first eval, 1 line search, 2 done
};

struct RefineParams {
  int S, d, n, npad, kind, acq, max_iter;
  MaternGen mg;
  double ell, kappa, eps, y_max;
  const double* X; const double* alpha; const double* Z; const double* bounds;   // bounds [d][2]
  double* kv; double* kp; double* v; double* w;     // [S][npad]
  double* wpart;                                     // nullable: [row chunks][S][npad] partial sums of w (round-by-round path)
  double* vec;                                       // per-seed vectors, stride vec_stride
  size_t vec_stride;
  RefineScalars* sc;
  double* val_out; double* grad_out;                 // evaluation-only mode: acq [S], grad [S][d]
  int* active;                                       // number of seeds still running
};

// vector slots inside a seed's block (each d doubles unless noted)
__device__ __forceinline__ double* rf_x(const RefineParams& p, int s) { return p.vec + s * p.vec_stride; }
__device__ __forceinline__ double* rf_xt(const RefineParams& p, int s) { return rf_x(p, s) + p.d; }
__device__ __forceinline__ double* rf_g(const RefineParams& p, int s) { return rf_x(p, s) + 2 * p.d; }
__device__ __forceinline__ double* rf_gt(const RefineParams& p, int s) { return rf_x(p, s) + 3 * p.d; }
__device__ __forceinline__ double* rf_dir(const RefineParams& p, int s) { return rf_x(p, s) + 4 * p.d; }
__device__ __forceinline__ double* rf_S(const RefineParams& p, int s, int i) { return rf_x(p, s) + (5 + i) * p.d; }
__device__ __forceinline__ double* rf_Y(const RefineParams& p, int s, int i) { return rf_x(p, s) + (5 + RF_M + i) * p.d; }
__device__ __forceinline__ double* rf_rho(const RefineParams& p, int s) { return rf_x(p, s) + (5 + 2 * RF_M) * p.d; }
inline size_t rf_vec_stride(int d) { return static_cast<size_t>(5 + 2 * RF_M) * d + RF_M; }

// k*_j and its r^2-derivative for seed blockIdx.y at its current trial point
__global__ void __launch_bounds__(128) refine_kstar_kernel(RefineParams p) {
  const int s = blockIdx.y, j = blockIdx.x * 128 + threadIdx.x;
  if (j >= p.npad) return;
  double kv = 0.0, kp = 0.0;
  if (j < p.n) {
    const double* x = rf_x(p, s);
    const double* xj = p.X + static_cast<size_t>(j) * p.d;
    double r2 = 0.0;
    for (int q = 0; q < p.d; ++q) { const double df = x[q] / p.ell - xj[q] / p.ell; r2 += df * df; }
    kv = kern_value(p.kind, r2, p.mg);
    kp = kern_dr2(p.kind, r2, p.mg);
  }
  p.kv[static_cast<size_t>(s) * p.npad + j] = kv;
  p.kp[static_cast<size_t>(s) * p.npad + j] = kp;
}

// v[s][i] = sum_{j <= i} Z[i][j] k[s][j]: one warp per row i, the row of Z stays in L1 while the
// seeds are walked in blocks of RF_SB
__global__ void __launch_bounds__(256) refine_v_kernel(RefineParams p) {
  const int lane = threadIdx.x & 31, i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= p.n) return;
  const double* Zi = p.Z + static_cast<size_t>(i) * p.npad;
  for (int s0 = 0; s0 < p.S; s0 += RF_SB) {
    double acc[RF_SB];
#pragma unroll
    for (int b = 0; b < RF_SB; ++b) acc[b] = 0.0;
    // four strides of the row at a time: their loads are issued together (the products are added in the same order as
    // one stride at a time would add them)
    int j = lane;
    for (; j + 96 <= i; j += 128) {
      double z[4], kk[4][RF_SB];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        z[u] = Zi[j + 32 * u];
#pragma unroll
        for (int b = 0; b < RF_SB; ++b) kk[u][b] = (s0 + b < p.S) ? p.kv[static_cast<size_t>(s0 + b) * p.npad + j + 32 * u] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int b = 0; b < RF_SB; ++b)
          if (s0 + b < p.S) acc[b] += z[u] * kk[u][b];
    }
    for (; j <= i; j += 32) {
      const double z = Zi[j];
#pragma unroll
      for (int b = 0; b < RF_SB; ++b)
        if (s0 + b < p.S) acc[b] += z * p.kv[static_cast<size_t>(s0 + b) * p.npad + j];
    }
#pragma unroll
    for (int b = 0; b < RF_SB; ++b) {
      double a = acc[b];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      if (lane == 0 && s0 + b < p.S) p.v[static_cast<size_t>(s0 + b) * p.npad + i] = a;
    }
  }
}

// w[s][j] = sum_{i >= j} Z[i][j] v[s][i] in two steps.  This kernel: CTA (rc, cb) takes rows 64 rc .. 64 rc + 63 and
// columns 128 cb .. 128 cb + 127 (skipped when the block lies above the diagonal), a thread per column, and leaves the
// partial sums of its rows in wpart[rc][s][j]; refine_step_kernel adds the chunks in order.  (One CTA per 32 columns
// walking ALL rows below them, as in round 1, was the slowest kernel of a round at n = 2048: 64 CTAs, 257 us.)
constexpr int RF_WROWS = 64;
__global__ void __launch_bounds__(128) refine_w_kernel(RefineParams p) {
  const int rc = blockIdx.x, cb = blockIdx.y, r0 = rc * RF_WROWS, j = cb * 128 + threadIdx.x;
  const int r1 = min(r0 + RF_WROWS, p.n);
  if (cb * 128 > r1 - 1) return;                                       // above the diagonal: nothing stored, nothing read
  __shared__ double vsm[RF_SB][RF_WROWS];                              // v of this CTA's rows for the seeds of a pass
  for (int s0 = 0; s0 < p.S; s0 += RF_SB) {
    __syncthreads();
    for (int e = threadIdx.x; e < RF_SB * RF_WROWS; e += 128) {
      const int b = e / RF_WROWS, i = r0 + e % RF_WROWS;
      vsm[b][e % RF_WROWS] = (s0 + b < p.S && i < r1) ? p.v[static_cast<size_t>(s0 + b) * p.npad + i] : 0.0;
    }
    __syncthreads();
    double acc[RF_SB];
#pragma unroll
    for (int b = 0; b < RF_SB; ++b) acc[b] = 0.0;
    if (j < p.n) {
      // eight rows of the column at a time: their loads are in flight together, the products are added row by row
      int i = max(r0, j);
      for (; i + 8 <= r1; i += 8) {
        double z[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) z[u] = p.Z[static_cast<size_t>(i + u) * p.npad + j];
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
          for (int b = 0; b < RF_SB; ++b)
            if (s0 + b < p.S) acc[b] += z[u] * vsm[b][i + u - r0];
      }
      for (; i < r1; ++i) {
        const double z = p.Z[static_cast<size_t>(i) * p.npad + j];
#pragma unroll
        for (int b = 0; b < RF_SB; ++b)
          if (s0 + b < p.S) acc[b] += z * vsm[b][i - r0];
      }
#pragma unroll
      for (int b = 0; b < RF_SB; ++b)
        if (s0 + b < p.S) p.wpart[(static_cast<size_t>(rc) * p.S + s0 + b) * p.npad + j] = acc[b];
    }
  }
}

// ---- warp-cooperative vector helpers: lane owns elements q = lane, lane + 32, ... ----
__device__ __forceinline__ double wdot(const double* a, const double* b, int d, int lane) {
  double s = 0.0;
  for (int q = lane; q < d; q += 32) s += a[q] * b[q];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}
__device__ __forceinline__ double wmax(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double wmin(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ bool rf_is_active(double x, double g, double lo, double hi) {
  return (x <= lo && g > 0.0) || (x >= hi && g < 0.0);
}

// projected-gradient infinity norm (L-BFGS-B's projgr)
__device__ __forceinline__ double rf_projgr(const RefineParams& p, const double* x, const double* g, int lane) {
  double m = 0.0;
  for (int q = lane; q < p.d; q += 32) {
    double gi = g[q];
    if (gi < 0.0) gi = fmax(x[q] - p.bounds[2 * q + 1], gi); else gi = fmin(x[q] - p.bounds[2 * q], gi);
    m = fmax(m, fabs(gi));
  }
  return wmax(m);
}

// search direction on the free variables: two-loop recursion, active components zeroed
__device__ void rf_direction(const RefineParams& p, int s, RefineScalars& sc, int lane, bool use_hist) {
  double* dir = rf_dir(p, s);
  const double* x = rf_x(p, s);
  const double* g = rf_g(p, s);
  double* rho = rf_rho(p, s);
  for (int q = lane; q < p.d; q += 32)
    dir[q] = rf_is_active(x[q], g[q], p.bounds[2 * q], p.bounds[2 * q + 1]) ? 0.0 : g[q];
  __syncwarp();
  double a[RF_M];
  const int nh = use_hist ? sc.hist_n : 0;
  for (int t = 0; t < nh; ++t) {                         // newest -> oldest
    const int i = (sc.hist_head - 1 - t + 2 * RF_M) % RF_M;
    a[t] = rho[i] * wdot(rf_S(p, s, i), dir, p.d, lane);
    const double* Y = rf_Y(p, s, i);
    for (int q = lane; q < p.d; q += 32) dir[q] -= a[t] * Y[q];
    __syncwarp();
  }
  const double gamma = nh > 0 ? sc.gamma : 1.0;
  for (int q = lane; q < p.d; q += 32) dir[q] *= gamma;
  __syncwarp();
  for (int t = nh - 1; t >= 0; --t) {                    // oldest -> newest
    const int i = (sc.hist_head - 1 - t + 2 * RF_M) % RF_M;
    const double b = rho[i] * wdot(rf_Y(p, s, i), dir, p.d, lane);
    const double* Sv = rf_S(p, s, i);
    for (int q = lane; q < p.d; q += 32) dir[q] += Sv[q] * (a[t] - b);
    __syncwarp();
  }
  for (int q = lane; q < p.d; q += 32) {
    const double lo = p.bounds[2 * q], hi = p.bounds[2 * q + 1];
    double v = rf_is_active(x[q], g[q], lo, hi) ? 0.0 : -dir[q];
    if ((x[q] >= hi && v > 0.0) || (x[q] <= lo && v < 0.0)) v = 0.0;   // never step out of the box
    dir[q] = v;
  }
  __syncwarp();
}

__device__ __forceinline__ void rf_set_trial(const RefineParams& p, int s, const RefineScalars& sc, int lane) {
  double* x = rf_x(p, s);
  const double* xt = rf_xt(p, s);
  const double* dir = rf_dir(p, s);
  for (int q = lane; q < p.d; q += 32) {
    const double lo = p.bounds[2 * q], hi = p.bounds[2 * q + 1];
    double v = xt[q] + sc.stp * dir[q];
    if (sc.stp == sc.stpmx && dir[q] != 0.0) {
      // the capped step lands exactly on the bound it was capped by
      const double lim = dir[q] > 0.0 ? (hi - xt[q]) / dir[q] : (lo - xt[q]) / dir[q];
      if (lim <= sc.stpmx) v = dir[q] > 0.0 ? hi : lo;
    }
    x[q] = fmin(fmax(v, lo), hi);
  }
  __syncwarp();
}

__device__ void rf_finish(const RefineParams& p, RefineScalars& sc, int status, int lane) {
  sc.phase = 2; sc.status = status;
  if (lane == 0) atomicSub(p.active, 1);
}

// start an iteration at the current iterate (x, f, g); returns with a trial point set or the seed done
__device__ void rf_begin_iteration(const RefineParams& p, int s, RefineScalars& sc, int lane) {
  double* x = rf_x(p, s); double* g = rf_g(p, s);
  for (int attempt = 0; attempt < 2; ++attempt) {
    if (rf_projgr(p, x, g, lane) <= 1e-5) { rf_finish(p, sc, 0, lane); return; }
    rf_direction(p, s, sc, lane, true);
    double* dir = rf_dir(p, s);
    double gd = wdot(g, dir, p.d, lane);
    if (!(gd < 0.0) && sc.hist_n > 0) {                   // not a descent direction: drop the memory
      sc.hist_n = 0;
      rf_direction(p, s, sc, lane, false);
      gd = wdot(g, dir, p.d, lane);
    }
    if (!(gd < 0.0)) { rf_finish(p, sc, 2, lane); return; }
    double lim = 1e10;
    for (int q = lane; q < p.d; q += 32) {
      if (dir[q] > 0.0) lim = fmin(lim, (p.bounds[2 * q + 1] - x[q]) / dir[q]);
      else if (dir[q] < 0.0) lim = fmin(lim, (p.bounds[2 * q] - x[q]) / dir[q]);
    }
    sc.stpmx = wmin(lim);
    const double dnorm = sqrt(wdot(dir, dir, p.d, lane));
    sc.stp = (sc.iter == 0 && sc.hist_n == 0) ? fmin(1.0 / dnorm, sc.stpmx) : fmin(1.0, sc.stpmx);
    double* xt = rf_xt(p, s); double* gt = rf_gt(p, s);
    for (int q = lane; q < p.d; q += 32) { xt[q] = x[q]; gt[q] = g[q]; }
    __syncwarp();
    sc.ft = sc.f; sc.ifun = 0; sc.gd0 = gd;
    sc.ls.step(true, sc.f, gd, sc.stp, 1e-3, 0.9, 0.1, 0.0, sc.stpmx);
    if (sc.ls.task == Dcsrch::ERROR) {
      if (sc.hist_n > 0) { sc.hist_n = 0; continue; }
      rf_finish(p, sc, 2, lane); return;
    }
    sc.ifun = 1; ++sc.nfev; sc.phase = 1;
    rf_set_trial(p, s, sc, lane);
    return;
  }
  rf_finish(p, sc, 2, lane);
}

// consume f, g at the trial point (already stored in sc.f / rf_g) and move the machine on
__device__ void rf_advance(const RefineParams& p, int s, RefineScalars& sc, int lane) {
  double* x = rf_x(p, s); double* g = rf_g(p, s);
  double* xt = rf_xt(p, s); double* gt = rf_gt(p, s);
  if (sc.phase == 0) {
    sc.nfev = 1; sc.iter = 0; sc.hist_n = 0; sc.hist_head = 0; sc.gamma = 1.0;
    rf_begin_iteration(p, s, sc, lane);
    return;
  }
  const double* dir = rf_dir(p, s);
  const double gd = wdot(g, dir, p.d, lane);
  sc.ls.step(false, sc.f, gd, sc.stp, 1e-3, 0.9, 0.1, 0.0, sc.stpmx);
  if (sc.ls.task != Dcsrch::CONVERGENCE && sc.ls.task != Dcsrch::WARNING) {
    ++sc.ifun; ++sc.nfev;
    if (sc.ifun - 1 >= 20 || !(sc.stp == sc.stp)) {       // line search exhausted (or NaN step)
      for (int q = lane; q < p.d; q += 32) { x[q] = xt[q]; g[q] = gt[q]; }
      __syncwarp();
      sc.f = sc.ft;
      if (sc.hist_n > 0) { sc.hist_n = 0; rf_begin_iteration(p, s, sc, lane); }
      else rf_finish(p, sc, 2, lane);
      return;
    }
    rf_set_trial(p, s, sc, lane);
    return;
  }
  // step accepted
  ++sc.iter;
  double* Sv = rf_S(p, s, sc.hist_head); double* Yv = rf_Y(p, s, sc.hist_head);
  for (int q = lane; q < p.d; q += 32) { Sv[q] = x[q] - xt[q]; Yv[q] = g[q] - gt[q]; }
  __syncwarp();
  const double sy = wdot(Sv, Yv, p.d, lane), yy = wdot(Yv, Yv, p.d, lane);
  if (sy > 2.220446049250313e-16 * yy) {
    if (lane == 0) rf_rho(p, s)[sc.hist_head] = 1.0 / sy;
    __syncwarp();
    sc.gamma = sy / yy;
    sc.hist_head = (sc.hist_head + 1) % RF_M;
    if (sc.hist_n < RF_M) ++sc.hist_n;
  }
  if (sc.iter >= p.max_iter) { rf_finish(p, sc, 1, lane); return; }
  if (rf_projgr(p, x, g, lane) <= 1e-5) { rf_finish(p, sc, 0, lane); return; }
  const double scale = fmax(fabs(sc.ft), fmax(fabs(sc.f), 1.0));
  if (sc.ft - sc.f <= 2.220446049250313e-16 * 1e7 * scale) { rf_finish(p, sc, 0, lane); return; }
  rf_begin_iteration(p, s, sc, lane);
}

// Reductions over the observations for seed s, the acquisition and its gradient, then (unless in evaluation-only
// mode) one optimiser transition by warp 0.  Called by all 128 threads of a CTA; grad_smem holds 2 d doubles.
__device__ __forceinline__ void refine_step_body(const RefineParams& p, int s, double* grad_smem) {
  __shared__ double red_mu[16], red_s2[16];
  __shared__ double gsum[16][2][32];
  __shared__ double sh_mu, sh_s2;
  double* gm = grad_smem;
  double* gv = grad_smem + p.d;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nw = nthr >> 5;   // 4 .. 16 warps
  const bool eval_only = p.val_out != nullptr;
  const double* kv = p.kv + static_cast<size_t>(s) * p.npad;
  const double* kp = p.kp + static_cast<size_t>(s) * p.npad;
  const double* v = p.v + static_cast<size_t>(s) * p.npad;
  const double* w = p.w + static_cast<size_t>(s) * p.npad;
  const double* x = rf_x(p, s);
  double mu = 0.0, s2 = 0.0;
  for (int j = tid; j < p.n; j += nthr) { mu += kv[j] * p.alpha[j]; s2 += v[j] * v[j]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { mu += __shfl_xor_sync(0xffffffffu, mu, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  if (lane == 0) { red_mu[warp] = mu; red_s2[warp] = s2; }
  // gradient sums: lane <-> dimension q (blocks of 32), warp <-> interleaved observations
  const double inv_l2 = 1.0 / (p.ell * p.ell);
  for (int q0 = 0; q0 < p.d; q0 += 32) {
    const int q = q0 + lane;
    double a1 = 0.0, a2 = 0.0;
    if (q < p.d) {
      const double xq = x[q];
      int j = warp;
      for (; j + 3 * nw < p.n; j += 4 * nw) {              // four observations' loads in flight, added in the same order
        double kpj[4], xj[4], aj[4], wj[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int ju = j + u * nw;
          kpj[u] = kp[ju]; xj[u] = p.X[static_cast<size_t>(ju) * p.d + q]; aj[u] = p.alpha[ju]; wj[u] = w[ju];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const double c = 2.0 * kpj[u] * inv_l2 * (xq - xj[u]);
          a1 += c * aj[u];
          a2 += c * wj[u];
        }
      }
      for (; j < p.n; j += nw) {
        const double c = 2.0 * kp[j] * inv_l2 * (xq - p.X[static_cast<size_t>(j) * p.d + q]);
        a1 += c * p.alpha[j];
        a2 += c * w[j];
      }
    }
    gsum[warp][0][lane] = a1; gsum[warp][1][lane] = a2;
    __syncthreads();
    if (warp == 0 && q < p.d) {
      // groups of four warps first, then the groups: with four warps this is the order (w0 + w1 + w2 + w3) as before
      double t1 = 0.0, t2 = 0.0;
      for (int w4 = 0; w4 < nw; w4 += 4) {
        t1 += gsum[w4][0][lane] + gsum[w4 + 1][0][lane] + gsum[w4 + 2][0][lane] + gsum[w4 + 3][0][lane];
        t2 += gsum[w4][1][lane] + gsum[w4 + 1][1][lane] + gsum[w4 + 2][1][lane] + gsum[w4 + 3][1][lane];
      }
      gm[q] = t1;
      gv[q] = -2.0 * t2;
    }
    __syncthreads();
  }
  if (tid == 0) {
    double a = 0.0, b = 0.0;
    for (int w4 = 0; w4 < nw; w4 += 4) {
      a += red_mu[w4] + red_mu[w4 + 1] + red_mu[w4 + 2] + red_mu[w4 + 3];
      b += red_s2[w4] + red_s2[w4 + 1] + red_s2[w4 + 2] + red_s2[w4 + 3];
    }
    sh_mu = a; sh_s2 = b;
  }
  __syncthreads();
  if (warp != 0) return;
  const double mean = sh_mu;
  double var = 1.0 - sh_s2;
  if (var < 0.0) var = 0.0;
  const double sd = sqrt(var);
  const double acq = acq_value(p.acq, mean, sd, p.y_max, p.kappa, p.eps);
  // d acq / dx = c_mean * grad(mean) + c_var * grad(var)
  double c_mean, c_var;
  if (p.acq == ACQ_UCB) { c_mean = 1.0; c_var = sd > 0.0 ? p.kappa / (2.0 * sd) : 0.0; }
  else {
    const double imp = mean - p.y_max - p.eps, z = imp / sd;
    const double cdf = norm_cdf(z), pdf = norm_pdf(z);
    if (p.acq == ACQ_EI) { c_mean = cdf; c_var = sd > 0.0 ? pdf / (2.0 * sd) : 0.0; }
    else if (sd > 0.0) { c_mean = pdf / sd; c_var = -pdf * z / (2.0 * sd * sd); }
    else { c_mean = 0.0; c_var = 0.0; }
  }
  if (eval_only) {
    if (lane == 0) p.val_out[s] = acq;
    for (int q = lane; q < p.d; q += 32) p.grad_out[static_cast<size_t>(s) * p.d + q] = c_mean * gm[q] + c_var * gv[q];
    return;
  }
  // minimise f = -acq
  double* g = rf_g(p, s);
  for (int q = lane; q < p.d; q += 32) g[q] = -(c_mean * gm[q] + c_var * gv[q]);
  __syncwarp();
  RefineScalars sc = p.sc[s];
  sc.f = -acq;
  rf_advance(p, s, sc, lane);
  if (lane == 0) p.sc[s] = sc;
}

// One CTA per seed, one round: used with refine_kstar / refine_v / refine_w kernels for large n.
__global__ void __launch_bounds__(512) refine_step_kernel(RefineParams p) {
  extern __shared__ double rf_grad_smem[];        // gm[d] | gv[d]
  const int s = blockIdx.x;
  if (p.val_out == nullptr && p.sc[s].phase == 2) return;
  if (p.wpart != nullptr) {
    // w[s][j] = sum over the row chunks at or below j of the partial sums refine_w_kernel left, in chunk order
    const int nrc = (p.n + RF_WROWS - 1) / RF_WROWS;
    for (int j = threadIdx.x; j < p.n; j += blockDim.x) {
      double t = 0.0;
      int rc = j / RF_WROWS;
      for (; rc + 8 <= nrc; rc += 8) {                                   // eight chunks' loads in flight, added in chunk order
        double u[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) u[q] = p.wpart[(static_cast<size_t>(rc + q) * p.S + s) * p.npad + j];
#pragma unroll
        for (int q = 0; q < 8; ++q) t += u[q];
      }
      for (; rc < nrc; ++rc) t += p.wpart[(static_cast<size_t>(rc) * p.S + s) * p.npad + j];
      p.w[static_cast<size_t>(s) * p.npad + j] = t;
    }
    __syncthreads();
  }
  refine_step_body(p, s, rf_grad_smem);
}

// The whole refinement of a seed in ONE launch (n <= RF_PERSIST_NMAX): the CTA of seed s loops over rounds -- k* and
// its derivative, v = L^-1 k*, w = L^-T v, acquisition, gradient and the optimiser transition -- until the seed has
// converged.  Same arithmetic and the same state machine as the round-by-round kernels above, which need four
// launches per round and a device->host check every few rounds: for small n that latency was the whole cost (config
// C1: 0.7 ms for microseconds of arithmetic).  The seed's optimiser vectors and k*, v, w live in shared memory for
// the duration (when they fit: d <= RF_PERSIST_DSMEM), and so does L^-1 up to 128 observations (128 KB); above that
// it is streamed from L2, several rows in flight per warp.
constexpr int RF_PERSIST_NMAX = 256;
constexpr int RF_PERSIST_DSMEM = 256;             // per-seed vectors in shared memory up to this many dimensions
constexpr int RF_PERSIST_ZSMEM = 128;             // L^-1 in shared memory up to this many observations
constexpr int RF_PERSIST_CHUNK = 32;              // above that: rows of L^-1 staged per pass through shared memory

// dynamic shared memory of refine_persist_kernel, in doubles (the host passes the same number)
inline size_t rf_persist_smem_doubles(int n, int npad, int d) {
  size_t t = static_cast<size_t>(2) * d + static_cast<size_t>(4) * npad;
  if (d <= RF_PERSIST_DSMEM) t += rf_vec_stride(d);
  if (n <= RF_PERSIST_ZSMEM) t += static_cast<size_t>(n) * (n | 1);
  else t += static_cast<size_t>(RF_PERSIST_CHUNK) * (n | 1);          // staging of RF_PERSIST_CHUNK rows of L^-1 at a time
  return t;
}

__global__ void __launch_bounds__(128) refine_persist_kernel(RefineParams p, int max_rounds) {
  extern __shared__ double rf_grad_smem[];        // gm[d] | gv[d] | k* | k*' | v | w | [seed vectors] | [L^-1]
  __shared__ int sh_phase;
  const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* kvs = rf_grad_smem + 2 * p.d;
  double* kps = kvs + p.npad;
  double* vs = kps + p.npad;
  double* ws = vs + p.npad;
  double* next = ws + p.npad;
  // q: the view of this seed with its working vectors redirected to shared memory (index s still selects them)
  RefineParams q = p;
  q.kv = kvs - static_cast<size_t>(s) * p.npad; q.kp = kps - static_cast<size_t>(s) * p.npad;
  q.v = vs - static_cast<size_t>(s) * p.npad; q.w = ws - static_cast<size_t>(s) * p.npad;
  const bool vec_smem = p.d <= RF_PERSIST_DSMEM;
  double* gvec = p.vec + static_cast<size_t>(s) * p.vec_stride;
  if (vec_smem) {
    for (size_t e = tid; e < p.vec_stride; e += 128) next[e] = gvec[e];
    q.vec = next - static_cast<size_t>(s) * p.vec_stride;
    next += p.vec_stride;
  }
  const double* Zp = p.Z;
  int zld = p.npad;
  if (p.n <= RF_PERSIST_ZSMEM) {
    zld = p.n | 1;                                   // odd leading dimension: a thread per ROW reads without bank conflicts
    for (int e = tid; e < p.n * p.n; e += 128) {
      const int i = e / p.n, j = e - i * p.n;
      next[i * zld + j] = (j <= i) ? p.Z[static_cast<size_t>(i) * p.npad + j] : 0.0;
    }
    Zp = next;
  }
  double* chunk = next;                              // n > RF_PERSIST_ZSMEM: staging of RF_PERSIST_CHUNK rows
  const bool staged = p.n > RF_PERSIST_ZSMEM;
  if (staged) zld = p.n | 1;
  __syncthreads();
#ifdef PLX_PROFILE
  long long tk = 0, tv = 0, tw = 0, ts = 0; int nr = 0;
#endif
  for (int round = 0; round < max_rounds; ++round) {
    if (tid == 0) sh_phase = q.sc[s].phase;
    __syncthreads();
    if (sh_phase == 2) break;
#ifdef PLX_PROFILE
    long long c0 = clock64(); ++nr;
#endif
    const double* x = rf_x(q, s);
    for (int j = tid; j < p.npad; j += 128) {
      double kv = 0.0, kp = 0.0;
      if (j < p.n) {
        const double* xj = p.X + static_cast<size_t>(j) * p.d;
        double r2 = 0.0;
        for (int c = 0; c < p.d; ++c) { const double df = x[c] / p.ell - xj[c] / p.ell; r2 += df * df; }
        kv = kern_value(p.kind, r2, p.mg);
        kp = kern_dr2(p.kind, r2, p.mg);
      }
      kvs[j] = kv; kps[j] = kp;
    }
    __syncthreads();
#ifdef PLX_PROFILE
    long long c1 = clock64(); tk += c1 - c0;
#endif
    if (!staged) {
      // v_i = sum_{j <= i} Z[i][j] k*_j: a thread per row, eight products in flight (row-wise reads are free of bank
      // conflicts thanks to the odd leading dimension)
      for (int i = tid; i < p.n; i += 128) {
        const double* Zi = Zp + static_cast<size_t>(i) * zld;
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
        int j = 0;
        for (; j + 8 <= i + 1; j += 8) {
          double z[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) z[u] = Zi[j + u];
#pragma unroll
          for (int u = 0; u < 8; ++u) acc[u & 3] += z[u] * kvs[j + u];
        }
        for (; j <= i; ++j) acc[0] += Zi[j] * kvs[j];
        vs[i] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
      }
      __syncthreads();
#ifdef PLX_PROFILE
      tv += clock64() - c1;
#endif
      // w_j = sum_{i >= j} Z[i][j] v_i, a thread per column
      for (int j = tid; j < p.n; j += 128) {
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
        int i = j;
        for (; i + 8 <= p.n; i += 8) {
          double z[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) z[u] = Zp[static_cast<size_t>(i + u) * zld + j];
#pragma unroll
          for (int u = 0; u < 8; ++u) acc[u & 3] += z[u] * vs[i + u];
        }
        for (; i < p.n; ++i) acc[0] += Zp[static_cast<size_t>(i) * zld + j] * vs[i];
        ws[j] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
      }
      __syncthreads();
    } else {
      // ONE pass over L^-1 from L2: RF_PERSIST_CHUNK rows at a time are staged in shared memory with all threads'
      // loads in flight; v of those rows needs only k*, and w accumulates the rows' contributions right away
      for (int j = tid; j < p.n; j += 128) ws[j] = 0.0;
      for (int r0 = 0; r0 < p.n; r0 += RF_PERSIST_CHUNK) {
        const int rows = min(RF_PERSIST_CHUNK, p.n - r0), width = r0 + rows;      // columns that can be non-zero
        // warp w stages rows w, w + 4, ... two at a time: 16 independent loads per thread in flight (n <= 256)
        for (int rp = warp; rp < rows; rp += 8) {
          double z[2][8];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int ri = rp + 4 * u;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const int j = lane + 32 * c;
              z[u][c] = (ri < rows && j <= r0 + ri) ? p.Z[static_cast<size_t>(r0 + ri) * p.npad + j] : 0.0;
            }
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int ri = rp + 4 * u;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const int j = lane + 32 * c;
              if (ri < rows && j < width) chunk[ri * zld + j] = z[u][c];
            }
          }
        }
        __syncthreads();
        {                                                    // four threads per row
          const int ri = tid >> 2, seg = tid & 3;
          double a = 0.0;
          if (ri < rows)
            for (int j = seg; j <= r0 + ri; j += 4) a += chunk[ri * zld + j] * kvs[j];
          a += __shfl_xor_sync(0xffffffffu, a, 1);
          a += __shfl_xor_sync(0xffffffffu, a, 2);
          if (seg == 0 && ri < rows) vs[r0 + ri] = a;
        }
        __syncthreads();
        for (int j = tid; j < width; j += 128) {
          double a = 0.0;
          for (int ri = 0; ri < rows; ++ri) a += chunk[ri * zld + j] * vs[r0 + ri];
          ws[j] += a;
        }
        __syncthreads();
      }
#ifdef PLX_PROFILE
      tv += clock64() - c1;
#endif
    }
#ifdef PLX_PROFILE
    long long c2 = clock64();
#endif
#ifdef PLX_PROFILE
    long long c3 = clock64(); tw += c3 - c2;   // (v and w are timed together under tv)
#endif
    refine_step_body(q, s, rf_grad_smem);
    __syncthreads();
#ifdef PLX_PROFILE
    ts += clock64() - c3;
#endif
  }
#ifdef PLX_PROFILE
  if (tid == 0 && nr > 0)
    printf("refine seed %d n=%d d=%d rounds %d: kstar %lld v %lld w %lld step %lld clk per round\n", s, p.n, p.d, nr, tk / nr, tv / nr,
           tw / nr, ts / nr);
#endif
  if (vec_smem) {
    // the host reads the final point (first d doubles of the block); the rest is kept consistent too
    const double* mine = rf_x(q, s);
    for (size_t e = tid; e < p.vec_stride; e += 128) gvec[e] = mine[e];
  }
}

}  // namespace plx
