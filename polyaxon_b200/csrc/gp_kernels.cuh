// Gaussian-process state kernels: kernel-matrix build, blocked FP64 Cholesky, triangular inverse,
// alpha = K^-1 y, explicit K^-1 for the marginal-likelihood gradient.  All batched over the
// length-scale restarts (grid.y / grid.z = batch index), all float64.
//
// Reference arithmetic being replaced (third-party, reached from optimizer.py:21):
//   sklearn _gpr.py:583-654  log_marginal_likelihood(theta, eval_gradient=True)
//   sklearn _gpr.py:349-367  L = cholesky(K + alpha I), alpha_ = cho_solve(L, y)
//   sklearn kernels.py:1558-1587 (RBF), :1713-1786 (Matern nu = 0.5 / 1.5 / 2.5)
//
// HBM layout per batch member b (npad = n rounded up to 128, leading dimension ld = npad):
//   A[b]  npad x ld  K + jitter*I (symmetric, identity in the padding); trailing updates in place;
//                    then Z = L^-1 (lower) during the triangular inverse, finally K^-1
//   L[b]  npad x ld  lower Cholesky factor (row-major)
//   U[b]  npad x ld  U = L^-T, upper triangular (row-major), i.e. U[c][r] = (L^-1)[r][c]
//   W[b]  nb x 128 x 128  inverses of the diagonal blocks of L (lower)
//   T[b]  npad x ld  scratch of the triangular inverse
#pragma once
#include "tma_ring.cuh"
#include <math.h>

namespace plx {

enum KernelKind { K_RBF = 0, K_M05 = 1, K_M15 = 2, K_M25 = 3, K_GEN = 4 };

__device__ __forceinline__ double exp_neg(double z);      // defined with the sweep's branch-free sqrt / exp below

// Matern with a general smoothness nu (sklearn kernels.py:1732-1738: 2^(1-nu)/Gamma(nu) * t^nu * K_nu(t),
// t = sqrt(2 nu) r, through scipy.special.kv).  Everything that depends on nu alone is prepared once on
// the host (api.cu: make_matern_gen); the device evaluates the modified Bessel function K by Temme's
// series for t < 1 and Steed's continued fraction CF2 above, then recurs upward in the order --
// the published algorithm (Temme 1975; Press et al., `bessik`), checked against mpmath to 3e-15.
struct MaternGen {
  double nu, mu, coef, s;                // order, fractional order, 2^(1-nu)/Gamma(nu), sqrt(2 nu)
  double gam1, gam2, gampl, gammi;       // Temme's Gamma_1, Gamma_2, 1/Gamma(1+mu), 1/Gamma(1-mu)
  int nl;                                // upward recurrence steps: nu = mu + nl (nl = 0: mu = -nu)
  const double* tab;                     // nullable, device: Chebyshev tables of the kernel value and its derivative (below)
};

// Tables of the general-nu kernel.  The Bessel routine below costs ~2 000 FP64 operations per value (series / continued
// fraction to convergence, pow, log, sinh): at the C3 shape 358 ms per 151 552 candidates of a sweep against 20 ms for
// the closed forms, and ten times a closed form's time for the kernel matrices of a fit.  For ONE nu the value
// k(t) = coef t^nu K_nu(t) and g(t) = coef t^(nu+1) K_(nu-1)(t), which carries both derivatives (dk/dtheta = g,
// dk/d(r^2) = -nu g / t^2), are fixed smooth functions of t > 0.  h1 = k e^t and h2 = g e^t are tabulated once per handle
// as Chebyshev interpolants of degree 15 on the intervals [t0 2^(i/2), t0 2^((i+1)/2)), i < MG_NINT (t0 = 2^-10 ... 512;
// the only singularity, t^(2 nu) at 0, is 4.8 half-widths away from every interval: interpolation error ~1e-16 of the
// interval's maximum), filled from the routine at the Chebyshev nodes.  Outside [t0, 512) the routine itself is used.
// Row i of a table: centre, 1 / half-width, c_0 / 2, c_1 .. c_15; the table of h2 follows that of h1.
constexpr int MG_E0 = 10, MG_NINT = 2 * (MG_E0 + 9), MG_DEG = 16, MG_ROW = MG_DEG + 2;
constexpr double MG_SQRT2 = 1.4142135623730951;
__device__ __forceinline__ double mg_edge(int i) {          // lower edge of interval i: exact products of powers of two and fl(sqrt 2)
  return __hiloint2double((1023 - MG_E0 + (i >> 1)) << 20, 0) * ((i & 1) ? MG_SQRT2 : 1.0);
}

// K_nu(x) and K_{nu-1}(x) for x > 0
__device__ __forceinline__ void bessel_k_pair(const MaternGen& m, double x, double& k, double& km1) {
  const double mu = m.mu, mu2 = mu * mu, xi = 1.0 / x, xi2 = 2.0 * xi;
  double rkmu, rk1;
  if (x < 1.0) {
    const double x2 = 0.5 * x, pimu = 3.141592653589793 * mu;
    double d = -log(x2), e = mu * d;
    const double fact2 = (fabs(e) < 1e-16) ? 1.0 : sinh(e) / e;
    const double fact = (fabs(pimu) < 1e-16) ? 1.0 : pimu / sin(pimu);
    double ff = fact * (m.gam1 * cosh(e) + m.gam2 * fact2 * d);
    double sum = ff;
    e = exp(e);
    double p = 0.5 * e / m.gampl, q = 0.5 / (e * m.gammi), c = 1.0, sum1 = p;
    d = x2 * x2;
    for (int i = 1; i < 500; ++i) {
      ff = (i * ff + p + q) / (i * i - mu2);
      c *= d / i;
      p /= (i - mu);
      q /= (i + mu);
      const double del = c * ff;
      sum += del;
      sum1 += c * (p - i * ff);
      if (fabs(del) < fabs(sum) * 1e-16) break;
    }
    rkmu = sum; rk1 = sum1 * xi2;
  } else {
    double b = 2.0 * (1.0 + x), d = 1.0 / b, h = d, delh = d, q1 = 0.0, q2 = 1.0;
    const double a1 = 0.25 - mu2;
    double q = a1, c = a1, a = -a1, s = 1.0 + q * delh;
    for (int i = 2; i < 10000; ++i) {
      a -= 2 * (i - 1);
      c = -a * c / i;
      const double qnew = (q1 - b * q2) / a;
      q1 = q2; q2 = qnew;
      q += c * qnew;
      b += 2.0;
      d = 1.0 / (b + a * d);
      delh = (b * d - 1.0) * delh;
      h += delh;
      const double dels = q * delh;
      s += dels;
      if (fabs(dels / s) < 1e-16) break;
    }
    h = a1 * h;
    rkmu = sqrt(3.141592653589793 / (2.0 * x)) * exp(-x) / s;
    rk1 = rkmu * (mu + x + 0.5 - h) * xi;
  }
  if (m.nl == 0) { k = rkmu; km1 = rk1; return; }      // mu = -nu: K_nu = K_mu, K_{nu-1} = K_{1-nu} = K_{mu+1}
  double prev = rkmu;
  for (int i = 1; i <= m.nl; ++i) {
    const double t = (mu + i) * xi2 * rk1 + rkmu;
    prev = rkmu; rkmu = rk1; rk1 = t;
  }
  k = rkmu; km1 = prev;
}

// k and g at t = sqrt(2 nu) r from the Bessel routine (the tables are filled from this; not inlined: it is the rare path of
// the inner loops, which would otherwise carry a copy of the routine per unrolled value)
__device__ __noinline__ void matern_gen_routine(const MaternGen& m, double t, double& val, double& g) {
  double k, km1;
  bessel_k_pair(m, t, k, km1);
  const double tp = m.coef * pow(t, m.nu);
  val = tp * k;
  g = tp * t * km1;
}

// One CTA per interval, one thread per Chebyshev node / coefficient.
__global__ void __launch_bounds__(MG_DEG) matern_tab_kernel(MaternGen m, double* __restrict__ tab) {
  __shared__ double hs[2][MG_DEG];
  const int i = blockIdx.x, k = threadIdx.x;
  const double a = mg_edge(i), b = mg_edge(i + 1), cen = 0.5 * (a + b), hw = 0.5 * (b - a);
  const double t = cen + hw * cospi((k + 0.5) / MG_DEG);
  double val, g;
  matern_gen_routine(m, t, val, g);
  const double et = exp(t);
  hs[0][k] = val * et; hs[1][k] = g * et;
  __syncthreads();
#pragma unroll
  for (int f = 0; f < 2; ++f) {
    double c = 0.0;
    for (int q = 0; q < MG_DEG; ++q) c += hs[f][q] * cospi(k * (q + 0.5) / MG_DEG);
    c *= 2.0 / MG_DEG;
    double* row = tab + (f * MG_NINT + i) * MG_ROW;
    row[2 + k] = (k == 0) ? 0.5 * c : c;
    if (k == 0) { row[0] = cen; row[1] = 1.0 / hw; }
  }
}

// interval of t in the tables, or -1 (no table, out of range, NaN): 2 floor(log2(t / t0)) + (mantissa >= sqrt 2)
__device__ __forceinline__ int mg_interval(const MaternGen& m, double t) {
  if (m.tab == nullptr || !(t >= mg_edge(0)) || !(t < 512.0)) return -1;
  const int hi = __double2hiint(t);
  const int e = ((hi >> 20) & 0x7ff) - 1023 + MG_E0;
  const double mant = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(t));
  return 2 * e + (mant >= MG_SQRT2 ? 1 : 0);
}
__device__ __forceinline__ double mg_cheb(const double* row, double x) {
  const double x2 = 2.0 * x;
  double b1 = row[2 + MG_DEG - 1], b2 = 0.0;
#pragma unroll
  for (int j = MG_DEG - 2; j >= 1; --j) { const double b0 = fma(x2, b1, row[2 + j] - b2); b2 = b1; b1 = b0; }
  return fma(x, b1, row[2] - b2);
}

// value, d/dtheta and d/d(r^2) of the general-nu Matern kernel at squared scaled distance r2
__device__ __forceinline__ void matern_gen_eval(const MaternGen& m, double r2, double& val, double& dtheta, double& dr2) {
  double r = sqrt(r2);
  if (r == 0.0) r = 2.220446049250313e-16;             // sklearn: K[K == 0.0] += eps
  const double t = m.s * r;
  double g;
  const int idx = mg_interval(m, t);
  if (idx < 0) {
    matern_gen_routine(m, t, val, g);
  } else {
    const double* row = m.tab + idx * MG_ROW;
    const double x = (t - row[0]) * row[1], et = exp_neg(-t);
    val = mg_cheb(row, x) * et;
    g = mg_cheb(row + MG_NINT * MG_ROW, x) * et;
  }
  dtheta = g;                                           // d/dtheta, theta = log(length_scale): dt/dtheta = -t
  dr2 = -m.nu * g / (t * t);                            // d/d(r^2)
}

// The kernel value alone (the sweeps).
__device__ __forceinline__ double matern_gen_value(const MaternGen& m, double r2) {
  double r = sqrt(r2);
  if (r == 0.0) r = 2.220446049250313e-16;             // sklearn: K[K == 0.0] += eps
  const double t = m.s * r;
  const int idx = mg_interval(m, t);
  if (idx < 0) { double val, g; matern_gen_routine(m, t, val, g); return val; }
  const double* row = m.tab + idx * MG_ROW;
  return mg_cheb(row, (t - row[0]) * row[1]) * exp_neg(-t);
}

__device__ __forceinline__ double kern_value(int kind, double r2, const MaternGen& mg) {
  if (kind == K_GEN) { double v, a, b; matern_gen_eval(mg, r2, v, a, b); return v; }
  if (kind == K_RBF) return exp(-0.5 * r2);
  const double r = sqrt(r2);
  if (kind == K_M05) return exp(-r);
  if (kind == K_M15) { const double t = r * 1.7320508075688772; return (1.0 + t) * exp(-t); }
  const double t = r * 2.23606797749979;
  return (1.0 + t + t * t / 3.0) * exp(-t);
}

// Branch-free square root and exponential for the sweep's inner loop.  The library versions carry
// a slow-path call each (denormals, overflow), which fences the scheduler: independent evaluations
// cannot be interleaved and the loop becomes latency-bound (ncu: ~210 cycles per evaluation against
// ~60 cycles of FP64 pipe work).  The arguments here are squared distances (>= 0) and non-positive
// exponents, so the special cases reduce to two selects.

// The range checks below compare the HIGH WORD of the argument as an integer: DSETP runs on the FP64 pipe, which is
// what the int8 sweep's builders are short of, the integer pipe is not.
constexpr int HI_1E_290 = 0x03b8f2b1;                    // high word just above that of 1e-290
constexpr int HI_708 = 0x40862000;                       // high word of 708.0 (low word 0)

// k * 2^54 for a kernel value 0 <= k <= 1 by an integer add on the exponent field (exact, like the multiplication;
// zero and denormal k turn into something below 2^-900, which the digit extraction rounds to zero all the same)
__device__ __forceinline__ double scale_2p54(double k) {
  return __hiloint2double(__double2hiint(k) + (54 << 20), __double2loint(k));
}

// sqrt(x) for x >= 0: hardware reciprocal-square-root seed (~2^-22), one coupled Goldschmidt step for the root and
// half the reciprocal root, one Heron correction with an exact FMA residual (<= 1 ulp); x below 1e-290 (and -0) -> 0,
// NaN -> NaN
__device__ __forceinline__ double sqrt_pos(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double r0 = x * y, h0 = 0.5 * y;
  const double e = fma(-r0, h0, 0.5);
  const double r1 = fma(r0, e, r0), h1 = fma(h0, e, h0);
  const double res = fma(-r1, r1, x);
  const double out = fma(res, h1, r1);
  return (__double2hiint(x) >= HI_1E_290) ? out : 0.0;
}

// exp(z) for z <= 0: z = k ln2 + r, |r| <= ln2/2, degree-13 Taylor polynomial in two interleaved
// Horner chains, exponent patched in by integer add (1-2 ulp); z below -708 -> 0, NaN -> NaN
__device__ __forceinline__ double exp_neg(double z) {
  const double kf = fma(z, 1.4426950408889634, 6755399441055744.0);
  const int k = __double2loint(kf);
  const double kd = kf - 6755399441055744.0;
  double r = fma(kd, -6.93147180369123816490e-01, z);
  r = fma(kd, -1.90821492927058770002e-10, r);
  const double r2 = r * r;
  // even and odd halves of sum r^n / n!
  double pe = 1.0 / 479001600.0;                         // 1/12!
  pe = fma(pe, r2, 1.0 / 3628800.0);
  pe = fma(pe, r2, 1.0 / 40320.0);
  pe = fma(pe, r2, 1.0 / 720.0);
  pe = fma(pe, r2, 1.0 / 24.0);
  pe = fma(pe, r2, 0.5);
  pe = fma(pe, r2, 1.0);
  double po = 1.0 / 6227020800.0;                        // 1/13!
  po = fma(po, r2, 1.0 / 39916800.0);
  po = fma(po, r2, 1.0 / 362880.0);
  po = fma(po, r2, 1.0 / 5040.0);
  po = fma(po, r2, 1.0 / 120.0);
  po = fma(po, r2, 1.0 / 6.0);
  po = fma(po, r2, 1.0);
  const double p = fma(po, r, pe);
  const double scaled = __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
  const int az = __double2hiint(z) & 0x7fffffff;         // |z| by its high word
  const double out = (az <= HI_708) ? scaled : 0.0;
  return (az > 0x7ff00000) ? z : out;                    // NaN
}

// N independent evaluations in lock step: every stage is written as a loop over the N values so the
// N dependency chains are interleaved instruction by instruction (the FP64 pipe needs ~10 independent
// operations in flight per scheduler to stay busy; one evaluation is a chain of ~22).
template <int N>
__device__ __forceinline__ void sqrt_pos_n(double (&x)[N]) {
  double y[N], e[N], r[N];
#pragma unroll
  for (int i = 0; i < N; ++i) asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y[i]) : "d"(x[i]));
#pragma unroll
  for (int i = 0; i < N; ++i) { r[i] = x[i] * y[i]; y[i] = 0.5 * y[i]; }
#pragma unroll
  for (int i = 0; i < N; ++i) e[i] = fma(-r[i], y[i], 0.5);
#pragma unroll
  for (int i = 0; i < N; ++i) { r[i] = fma(r[i], e[i], r[i]); y[i] = fma(y[i], e[i], y[i]); }
#pragma unroll
  for (int i = 0; i < N; ++i) e[i] = fma(-r[i], r[i], x[i]);
#pragma unroll
  for (int i = 0; i < N; ++i) { const double o = fma(e[i], y[i], r[i]); x[i] = (__double2hiint(x[i]) >= HI_1E_290) ? o : 0.0; }
}

template <int N>
__device__ __forceinline__ void exp_neg_n(double (&z)[N]) {
  double kf[N], r[N], r2[N], pe[N], po[N];
#pragma unroll
  for (int i = 0; i < N; ++i) kf[i] = fma(z[i], 1.4426950408889634, 6755399441055744.0);
#pragma unroll
  for (int i = 0; i < N; ++i) { const double kd = kf[i] - 6755399441055744.0; r[i] = fma(kd, -6.93147180369123816490e-01, z[i]); r[i] = fma(kd, -1.90821492927058770002e-10, r[i]); }
#pragma unroll
  for (int i = 0; i < N; ++i) { r2[i] = r[i] * r[i]; pe[i] = 1.0 / 479001600.0; po[i] = 1.0 / 6227020800.0; }
  const double ce[6] = {1.0 / 3628800.0, 1.0 / 40320.0, 1.0 / 720.0, 1.0 / 24.0, 0.5, 1.0};
  const double co[6] = {1.0 / 39916800.0, 1.0 / 362880.0, 1.0 / 5040.0, 1.0 / 120.0, 1.0 / 6.0, 1.0};
#pragma unroll
  for (int s = 0; s < 6; ++s)
#pragma unroll
    for (int i = 0; i < N; ++i) { pe[i] = fma(pe[i], r2[i], ce[s]); po[i] = fma(po[i], r2[i], co[s]); }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const double p = fma(po[i], r[i], pe[i]);
    const double scaled = __hiloint2double(__double2hiint(p) + (__double2loint(kf[i]) << 20), __double2loint(p));
    const int az = __double2hiint(z[i]) & 0x7fffffff;
    const double out = (az <= HI_708) ? scaled : 0.0;
    z[i] = (az > 0x7ff00000) ? z[i] : out;
  }
}

// k(r^2) for N squared scaled distances at once (closed-form kernels), in place
template <int KIND, int N>
__device__ __forceinline__ void kern_values_n(double (&v)[N]) {
  if (KIND == K_RBF) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = -0.5 * v[i];
    exp_neg_n<N>(v);
    return;
  }
  sqrt_pos_n<N>(v);
  double t[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    t[i] = (KIND == K_M05) ? v[i] : (KIND == K_M15 ? v[i] * 1.7320508075688772 : v[i] * 2.23606797749979);
    v[i] = -t[i];
  }
  exp_neg_n<N>(v);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    if (KIND == K_M15) v[i] = (1.0 + t[i]) * v[i];
    else if (KIND == K_M25) v[i] = (1.0 + t[i] + t[i] * t[i] * 0.3333333333333333) * v[i];   // no FP64 division in the sweep's inner loop
  }
}

// the same for any kernel kind: the general-nu Matern value by value (table, gp_kernels.cuh above)
template <int KIND, int N>
__device__ __forceinline__ void kern_values_any(double (&v)[N], const MaternGen& mg) {
  if constexpr (KIND == K_GEN) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = matern_gen_value(mg, v[i]);
  } else {
    kern_values_n<KIND, N>(v);
  }
}

template <int KIND>
__device__ __forceinline__ double kern_value_t(double r2, const MaternGen& mg) {
  if (KIND == K_GEN) return matern_gen_value(mg, r2);
  if (KIND == K_RBF) return exp_neg(-0.5 * r2);
  const double r = sqrt_pos(r2);
  if (KIND == K_M05) return exp_neg(-r);
  if (KIND == K_M15) { const double t = r * 1.7320508075688772; return (1.0 + t) * exp_neg(-t); }
  const double t = r * 2.23606797749979;
  return (1.0 + t + t * t * 0.3333333333333333) * exp_neg(-t);
}

// test hook: out[i] = kern_value_t<kind>(r2[i])
__global__ void kern_value_probe_kernel(const double* __restrict__ r2, int n, int kind, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v[1] = {r2[i]};
  switch (kind) {
    case K_RBF: kern_values_n<K_RBF, 1>(v); break;
    case K_M05: kern_values_n<K_M05, 1>(v); break;
    case K_M15: kern_values_n<K_M15, 1>(v); break;
    default: kern_values_n<K_M25, 1>(v); break;
  }
  out[i] = v[0];
}

// dK/dtheta for theta = log(length_scale) as sklearn forms it (kernels.py:1575, :1752-1773)
__device__ __forceinline__ double kern_dtheta(int kind, double r2, const MaternGen& mg) {
  if (kind == K_GEN) { double v, a, b; matern_gen_eval(mg, r2, v, a, b); return a; }
  if (kind == K_RBF) return exp(-0.5 * r2) * r2;
  const double r = sqrt(r2);
  const double d2 = r * r;                       // sklearn squares the rooted distance again
  if (kind == K_M05) return (r != 0.0) ? exp(-r) * d2 / r : 0.0;
  if (kind == K_M15) return 3.0 * d2 * exp(-sqrt(3.0 * d2));
  const double t = sqrt(5.0 * d2);
  return 5.0 / 3.0 * d2 * (t + 1.0) * exp(-t);
}

struct BatchPtrs {
  double* A; double* L; double* U; double* W; double* T;
  size_t mat_stride;     // npad * ld
  size_t blk_stride;     // nb * 128 * 128
};

// ---------------------------------------------------------------------------------------------
// K(X, X) + jitter*I for every theta of the batch.  32x32 output tile per CTA, X rows staged in
// shared memory in dimension chunks of 32.  HBM-bound: writes npad^2 doubles per batch member.
// mode 0: write K into A (padding = identity).
// mode 1: gradient reduction  sum_{a,b<n} (alpha_a alpha_b - Kinv[a][b]) * dK/dtheta[a][b]
//         over the lower 32-tiles (off-diagonal tiles weighted 2), one partial per CTA.
// ---------------------------------------------------------------------------------------------
constexpr int DCH = 32;

// X / length_scale for every theta of the batch: Xs[b][n][d] (the division sklearn does, done once)
// (also clears the batch member's log-determinant and status cells: the first kernel of every evaluation)
__global__ void scale_x_batch_kernel(const double* __restrict__ X, int nd, const double* __restrict__ thetas,
                                     double* __restrict__ Xs, double* __restrict__ logdet, int* __restrict__ info) {
  const int b = blockIdx.y, e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nd) Xs[static_cast<size_t>(b) * nd + e] = X[e] / exp(thetas[b]);
  if (e == 0) { logdet[b] = 0.0; info[b] = 0; }
}

template <int MODE>
__global__ void __launch_bounds__(256) kmat_kernel(const double* __restrict__ Xs, int n, int d, int npad,
                                                   int kind, MaternGen mg, double jitter, BatchPtrs bp,
                                                   const double* __restrict__ alpha,
                                                   double* __restrict__ partial, int ntri) {
  __shared__ double Xi[32][DCH + 1];
  __shared__ double Xj[32][DCH + 1];
  __shared__ double red[8];
  const int b = blockIdx.z;
  // linear index over the lower triangle of 32-tiles: p = ti*(ti+1)/2 + tj.  Only the lower triangle
  // of K is ever read (factorisation, panels and trailing updates all work on i >= j).
  int ti, tj;
  {
    const int p = blockIdx.x;
    ti = static_cast<int>((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= p) ++ti;
    while (ti * (ti + 1) / 2 > p) --ti;
    tj = p - ti * (ti + 1) / 2;
  }
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const double* Xb = Xs + static_cast<size_t>(b) * n * d;
  const int i0 = ti * 32, j0 = tj * 32;
  double r2[4] = {0.0, 0.0, 0.0, 0.0};
  for (int dc = 0; dc < d; dc += DCH) {
    const int dn = min(DCH, d - dc);
    __syncthreads();
    for (int e = threadIdx.x; e < 32 * DCH; e += 256) {
      const int row = e / DCH, col = e % DCH;
      double vi = 0.0, vj = 0.0;
      if (col < dn) {
        if (i0 + row < n) vi = Xb[static_cast<size_t>(i0 + row) * d + dc + col];
        if (j0 + row < n) vj = Xb[static_cast<size_t>(j0 + row) * d + dc + col];
      }
      Xi[row][col] = vi; Xj[row][col] = vj;
    }
    __syncthreads();
    for (int q = 0; q < dn; ++q) {
      const double xj = Xj[tx][q];
#pragma unroll
      for (int r = 0; r < 4; ++r) { const double df = Xi[ty + 8 * r][q] - xj; r2[r] += df * df; }
    }
  }
  const size_t ld = npad;
  double kv[4] = {r2[0], r2[1], r2[2], r2[3]};
  if (MODE == 0) {
    // correctly rounded library sqrt / exp here, not the sweep's 4-ulp versions: the optimiser's line
    // search works on LML differences at the 1e-12 level near convergence, and extra noise in K costs
    // it evaluations
#pragma unroll
    for (int r = 0; r < 4; ++r) kv[r] = kern_value(kind, r2[r], mg);
    double* A = bp.A + b * bp.mat_stride;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = i0 + ty + 8 * r, j = j0 + tx;
      double v;
      if (i < n && j < n) v = (i == j) ? 1.0 + jitter : kv[r];
      else v = (i == j) ? 1.0 : 0.0;
      A[i * ld + j] = v;
    }
  } else {
#pragma unroll
    for (int r = 0; r < 4; ++r) kv[r] = kern_dtheta(kind, r2[r], mg);
    const double* Kinv = bp.A + b * bp.mat_stride;
    const double* al = alpha + static_cast<size_t>(b) * npad;
    double s = 0.0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = i0 + ty + 8 * r, j = j0 + tx;
      if (i < n && j < n) {
        const double g = (i == j) ? 0.0 : kv[r];
        s += (al[i] * al[j] - Kinv[i * ld + j]) * g;
      }
    }
    if (ti != tj) s *= 2.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (tx == 0) red[ty] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < 8; ++w) t += red[w];
      partial[static_cast<size_t>(b) * ntri + blockIdx.x] = t;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Diagonal block k: unblocked Cholesky of the 128x128 block in shared memory, then its inverse by
// column-parallel forward substitution.  Writes L_kk, W_k = L_kk^-1, U_kk = W_k^T, adds the
// block's sum(log diag) to logdet[b] and records the first non-positive pivot in info[b].
// One CTA per batch member; latency-bound (128 dependent column steps).
// ---------------------------------------------------------------------------------------------
constexpr int POTF_LD = 129;
constexpr int POTF_SMEM = (128 * POTF_LD + 256) * 8;

// Columns 16*BJ .. 16*BJ+15 of the register-tiled factorisation.  Thread (ti, tc) holds
// A[ti + 16 a][tc + 16 q]; in this range only a, q >= BJ are live, so that is all that is updated.
// Column j arrives in colbuf[j & 1] already updated through step j - 1: the threads that own column
// j + 1 apply step j to it first and publish it BEFORE the barrier, everybody applies step j to the
// rest of the tile AFTER it -- that bulk update overlaps the pivot chain of step j + 1.
template <int BJ>
__device__ __forceinline__ bool potf_factor_cols(double (&acc)[8][8], double* S, double (*colbuf)[128], double* dl,
                                                 double* wdiag, int* bad, int k, int m, int ti, int tc, int tid) {
#pragma unroll 1
  for (int jj = 0; jj < 16; ++jj) {
    const int j = 16 * BJ + jj;
    if (j >= m) return false;
    const double* cb = colbuf[j & 1];
    double* nx = colbuf[(j + 1) & 1];
    double ajj = cb[j];
    if (!(ajj > 0.0)) {                 // non-positive or NaN pivot: not positive definite
      if (tid == 0 && *bad == 0) *bad = k * 128 + j + 1;
      ajj = 1.0;
    }
    // root and reciprocal root from one hardware seed and a coupled Newton step (<= 2 ulp): the
    // pivot -> sqrt -> divide chain is on the critical path of every column
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(ajj));
    const double e0 = fma(ajj, -(y0 * y0), 1.0);
    double inv = fma(fma(e0, 0.375, 0.5), y0 * e0, y0);
    double ljj = ajj * inv;
    ljj = fma(fma(ljj, -ljj, ajj), 0.5 * inv, ljj);
    inv = fma(fma(-ljj, inv, 1.0), inv, inv);            // 1 / ljj (LAPACK dpotf2 scales by the reciprocal too)
    double li[8], lc[8];
#pragma unroll
    for (int a = BJ; a < 8; ++a) { li[a] = cb[ti + 16 * a] * inv; lc[a] = cb[tc + 16 * a] * inv; }
    if (jj < 15) {
      if (tc == jj + 1) {
#pragma unroll
        for (int a = BJ; a < 8; ++a) nx[ti + 16 * a] = fma(-li[a], lc[BJ], acc[a][BJ]);
      }
    } else if constexpr (BJ < 7) {
      if (tc == 0) {
#pragma unroll
        for (int a = BJ + 1; a < 8; ++a) nx[ti + 16 * a] = fma(-li[a], lc[BJ + 1], acc[a][BJ + 1]);
      }
    }
    if (tc == jj) {
#pragma unroll
      for (int a = BJ; a < 8; ++a) { const int i = ti + 16 * a; if (i > j) S[i * POTF_LD + j] = li[a]; }
    }
    if (tid == 0) { dl[j] = ljj; wdiag[j] = inv; }
    __syncthreads();
#pragma unroll
    for (int a = BJ; a < 8; ++a)
#pragma unroll
      for (int q = BJ; q < 8; ++q) acc[a][q] = fma(-li[a], lc[q], acc[a][q]);
  }
  return true;
}

// Rows 16*BJ .. 16*BJ+15 of W = L^-1 by right-looking forward substitution on the identity, same
// tiling: acc[a][q] = W[ti + 16 a][tc + 16 q], live for a >= BJ (rows still open) and q <= BJ
// (W is lower triangular).  Row j arrives finished in colbuf[j & 1]; the owners of row j + 1 apply
// step j to it, scale it and publish it before the barrier, the bulk update follows the barrier.
// W[j][c] (c < j) is parked at S[c][j], the diagonal stays in wdiag.
template <int BJ>
__device__ __forceinline__ bool potf_inverse_rows(double (&acc)[8][8], double* S, double (*colbuf)[128],
                                                  const double* wdiag, int m, int ti, int tc, int tid) {
#pragma unroll 1
  for (int jj = 0; jj < 16; ++jj) {
    const int j = 16 * BJ + jj;
    if (j >= m) return false;
    const double* rb = colbuf[j & 1];
    double* nx = colbuf[(j + 1) & 1];
    double wr[8], lcol[8];
#pragma unroll
    for (int q = 0; q <= BJ; ++q) wr[q] = rb[tc + 16 * q];
#pragma unroll
    for (int a = BJ; a < 8; ++a) {
      const int i = ti + 16 * a;
      lcol[a] = (i > j && i < m) ? S[i * POTF_LD + j] : 0.0;
    }
    if (j + 1 < m) {
      if (jj < 15) {
        if (ti == jj + 1) {
          const double w1 = wdiag[j + 1];
#pragma unroll
          for (int q = 0; q <= BJ; ++q) nx[tc + 16 * q] = fma(-lcol[BJ], wr[q], acc[BJ][q]) * w1;
        }
      } else if constexpr (BJ < 7) {
        if (ti == 0) {
          const double w1 = wdiag[j + 1];
#pragma unroll
          for (int q = 0; q <= BJ; ++q) nx[tc + 16 * q] = fma(-lcol[BJ + 1], wr[q], acc[BJ + 1][q]) * w1;
          nx[tc + 16 * (BJ + 1)] = acc[BJ + 1][BJ + 1] * w1;
        }
      }
    }
    if (tid < j) S[tid * POTF_LD + j] = rb[tid];
    __syncthreads();
#pragma unroll
    for (int a = BJ; a < 8; ++a)
#pragma unroll
      for (int q = 0; q <= BJ; ++q) acc[a][q] = fma(-lcol[a], wr[q], acc[a][q]);
  }
  return true;
}

__global__ void __launch_bounds__(256) potf2_inv_kernel(BatchPtrs bp, int k, int n, int npad,
                                                        double* __restrict__ logdet, int* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* S = reinterpret_cast<double*>(smem_raw);        // S[i][j] at S[i*129 + j]
  double* dl = S + 128 * POTF_LD;                         // diagonal of L
  double* wdiag = dl + 128;                               // diagonal of W
  const int b = blockIdx.x, tid = threadIdx.x;
  const size_t ld = npad;
  const double* A = bp.A + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  for (int e = tid; e < 128 * 128; e += 256) {
    const int i = e >> 7, j = e & 127;
    S[i * POTF_LD + j] = A[i * ld + j];
  }
  __shared__ int bad;
  __shared__ double colbuf[2][128];
  if (tid == 0) bad = 0;
  // rows / columns beyond the observations are identity padding: the factor and its inverse are the
  // identity there, so every loop stops at the last real row of this block
  const int m = min(128, max(0, n - k * 128));
  // Register-tiled right-looking factorisation: thread (ti, tc) keeps the 8 x 8 strided sub-block
  // A[ti + 16 a][tc + 16 b] in registers for the whole factorisation; per column only the column
  // itself crosses shared memory (double-buffered, one barrier per column).  Entries left of / above
  // the current column are dead and receive harmless garbage updates.
  const int tc = tid & 15, ti = tid >> 4;
  double acc[8][8];
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[a][q] = S[(ti + 16 * a) * POTF_LD + tc + 16 * q];
  if (tc == 0) {
#pragma unroll
    for (int a = 0; a < 8; ++a) colbuf[0][ti + 16 * a] = acc[a][0];
  }
  __syncthreads();
  potf_factor_cols<0>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<1>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<2>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<3>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<4>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<5>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<6>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid) &&
      potf_factor_cols<7>(acc, S, colbuf, dl, wdiag, &bad, k, m, ti, tc, tid);
  __syncthreads();
  if (tid >= m && tid < 128) { dl[tid] = 1.0; wdiag[tid] = 1.0; }
  // strict upper triangle will receive W^T; padding rows of L are zero
  for (int e = tid; e < 128 * 128; e += 256) {
    const int r = e >> 7, c = e & 127;
    if (c > r) S[r * POTF_LD + c] = 0.0;
    else if (c < r && r >= m) S[r * POTF_LD + c] = 0.0;
  }
  __syncthreads();
  if (tid < 128) S[tid * POTF_LD + tid] = dl[tid];
  // W = L^-1 by right-looking forward substitution on all columns of the identity at once, same
  // register tiling: acc[a][q] = W[ti + 16 a][tc + 16 q]; per step only the finished row crosses
  // shared memory.  W[j][c] (c < j) is parked at S[c][j], the diagonal in wdiag.
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[a][q] = (ti + 16 * a == tc + 16 * q) ? 1.0 : 0.0;
  if (ti == 0) colbuf[0][tc] = (tc == 0) ? wdiag[0] : 0.0;       // row 0 of W
  __syncthreads();
  potf_inverse_rows<0>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<1>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<2>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<3>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<4>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<5>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<6>(acc, S, colbuf, wdiag, m, ti, tc, tid) &&
      potf_inverse_rows<7>(acc, S, colbuf, wdiag, m, ti, tc, tid);
  __syncthreads();
  double* Lo = bp.L + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  double* Uo = bp.U + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  double* Wo = bp.W + b * bp.blk_stride + static_cast<size_t>(k) * 128 * 128;
  // the block's inverse also goes where its input came from: A doubles as Z = L^-1 (lower) from here on
  double* Zo = bp.A + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  for (int e = tid; e < 128 * 128; e += 256) {
    const int i = e >> 7, j = e & 127;
    double l, w;
    if (i > j) { l = S[i * POTF_LD + j]; w = S[j * POTF_LD + i]; }
    else if (i == j) { l = S[i * POTF_LD + i]; w = wdiag[i]; }
    else { l = 0.0; w = 0.0; }
    Lo[i * ld + j] = l;
    Wo[i * 128 + j] = w;
    Zo[i * ld + j] = w;
    // U_kk = W^T: element (i, j) of U is W[j][i]
    double u;
    if (j > i) u = S[i * POTF_LD + j]; else if (i == j) u = wdiag[i]; else u = 0.0;
    Uo[i * ld + j] = u;
  }
  if (tid < 32) {
    double s = 0.0;
    for (int i = tid; i < 128; i += 32)
      if (k * 128 + i < n) s += log(S[i * POTF_LD + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (tid == 0) {
      logdet[b] += s;
      if (bad != 0 && info[b] == 0) info[b] = bad;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Fused form of the diagonal-block kernel: the factorisation and the inversion of the factor advance TOGETHER, one
// column step each per barrier, instead of 128 steps of one after 128 steps of the other.  Both updates of step j
// need the same thing, column j of L:
//   trailing matrix   A[i][c] -= L[i][j] L[c][j]                (i, c > j)
//   rows of L^-1      R[i][c] -= L[i][j] W[j][c]                (i > j, c <= j),  W[j][.] = R[j][.] / l_jj
// (R starts as the identity; row j of R is complete after step j - 1 and is published UNSCALED, because l_jj is only
// known at step j; it is scaled where it is consumed).  The serial chain per step is the factor's alone -- pivot,
// reciprocal root, next column -- so the 2 x 128 dependent steps of potf2_inv_kernel become 128.
// Same register tiling: thread (ti, tc) holds A[ti + 16 a][tc + 16 q] and R[ti + 16 a][tc + 16 q]; in block column
// BJ the live parts are a, q >= BJ of A and a >= BJ, q <= BJ of R (column block BJ of R is created when it is
// first needed, which keeps the live registers at 9 (8 - BJ) values).
// ---------------------------------------------------------------------------------------------
// Pivot of column step j from the published column: l_jj and 1 / l_jj (reciprocal-square-root seed, one coupled
// Newton step, one Heron correction, one Newton step for the reciprocal).  A non-positive or NaN pivot is recorded
// and replaced by 1 so that the remaining steps stay finite.
__device__ __forceinline__ void potf_pivot(const double* cb, int j, int k, int* bad, int tid, double& ljj, double& inv) {
  double ajj = cb[j];
  if (!(ajj > 0.0)) {                   // non-positive or NaN pivot: not positive definite
    if (tid == 0 && *bad == 0) *bad = k * 128 + j + 1;
    ajj = 1.0;
  }
  double y0;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(ajj));
  const double e0 = fma(ajj, -(y0 * y0), 1.0);
  inv = fma(fma(e0, 0.375, 0.5), y0 * e0, y0);
  ljj = ajj * inv;
  ljj = fma(fma(ljj, -ljj, ajj), 0.5 * inv, ljj);
  inv = fma(fma(-ljj, inv, 1.0), inv, inv);              // 1 / ljj
}

// (ljj, inv) enter with the pivot of this block's first column and leave with that of the next block's: the pivot of
// step j + 1 is computed right after the barrier of step j, in the same basic block as the rank-1 update, so that its
// dependent chain of ~12 FP64 operations is interleaved with the update's independent ones instead of standing at the
// head of the next step.
template <int BJ>
__device__ __forceinline__ bool potf_fused_cols(double (&accA)[8][8], double (&accW)[8][8], double* S,
                                                double (*colbuf)[128], double (*rowbuf)[128], double* dl, double* wdiag,
                                                int* bad, int k, int m, int ti, int tc, int tid, double& ljj, double& inv) {
  // column block BJ of R: the identity (rows of earlier blocks never reach it)
#pragma unroll
  for (int a = BJ; a < 8; ++a) accW[a][BJ] = (a == BJ && ti == tc) ? 1.0 : 0.0;
#pragma unroll 4
  for (int jj = 0; jj < 16; ++jj) {
    const int j = 16 * BJ + jj;
    if (j >= m) return false;
    const double* cb = colbuf[j & 1];
    double* cnx = colbuf[(j + 1) & 1];
    const double* rb = rowbuf[j & 1];
    double* rnx = rowbuf[(j + 1) & 1];
    double li[8], lc[8], wr[8];
#pragma unroll
    for (int a = BJ; a < 8; ++a) { li[a] = cb[ti + 16 * a] * inv; lc[a] = cb[tc + 16 * a] * inv; }
#pragma unroll
    for (int q = 0; q <= BJ; ++q) wr[q] = rb[tc + 16 * q] * inv;          // W[j][tc + 16 q] (zero beyond column j)
    if (jj < 15) {
      if (tc == jj + 1) {
#pragma unroll
        for (int a = BJ; a < 8; ++a) cnx[ti + 16 * a] = fma(-li[a], lc[BJ], accA[a][BJ]);
      }
      if (ti == jj + 1) {
#pragma unroll
        for (int q = 0; q <= BJ; ++q) rnx[tc + 16 * q] = fma(-li[BJ], wr[q], accW[BJ][q]);
      }
    } else if constexpr (BJ < 7) {
      if (tc == 0) {
#pragma unroll
        for (int a = BJ + 1; a < 8; ++a) cnx[ti + 16 * a] = fma(-li[a], lc[BJ + 1], accA[a][BJ + 1]);
      }
      if (ti == 0) {
#pragma unroll
        for (int q = 0; q <= BJ; ++q) rnx[tc + 16 * q] = fma(-li[BJ + 1], wr[q], accW[BJ + 1][q]);
        rnx[tc + 16 * (BJ + 1)] = (tc == 0) ? 1.0 : 0.0;                  // the identity part of row j + 1
      }
    }
    if (tc == jj) {
#pragma unroll
      for (int a = BJ; a < 8; ++a) { const int i = ti + 16 * a; if (i > j) S[i * POTF_LD + j] = li[a]; }
    }
    if (tid < j) S[tid * POTF_LD + j] = rb[tid] * inv;                    // W[j][c], c < j, parked at S[c][j]
    if (tid == 0) { dl[j] = ljj; wdiag[j] = inv; }
    __syncthreads();
    if (j + 1 < m && j + 1 < 128) potf_pivot(cnx, j + 1, k, bad, tid, ljj, inv);      // column j + 1 is in cnx now
#pragma unroll
    for (int a = BJ; a < 8; ++a) {
#pragma unroll
      for (int q = BJ; q < 8; ++q) accA[a][q] = fma(-li[a], lc[q], accA[a][q]);
#pragma unroll
      for (int q = 0; q <= BJ; ++q) accW[a][q] = fma(-li[a], wr[q], accW[a][q]);
    }
  }
  return true;
}

__global__ void __launch_bounds__(256) potf2_fused_kernel(BatchPtrs bp, int k, int n, int npad,
                                                          double* __restrict__ logdet, int* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* S = reinterpret_cast<double*>(smem_raw);        // S[i][j] at S[i*129 + j]
  double* dl = S + 128 * POTF_LD;                         // diagonal of L
  double* wdiag = dl + 128;                               // diagonal of W
  const int b = blockIdx.x, tid = threadIdx.x;
  const size_t ld = npad;
  const double* A = bp.A + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  for (int e = tid; e < 128 * 128; e += 256) {
    const int i = e >> 7, j = e & 127;
    S[i * POTF_LD + j] = A[i * ld + j];
  }
  __shared__ int bad;
  __shared__ double colbuf[2][128];
  __shared__ double rowbuf[2][128];
  if (tid == 0) bad = 0;
  const int m = min(128, max(0, n - k * 128));
  const int tc = tid & 15, ti = tid >> 4;
  double accA[8][8], accW[8][8];
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int q = 0; q < 8; ++q) accA[a][q] = S[(ti + 16 * a) * POTF_LD + tc + 16 * q];
  if (tc == 0) {
#pragma unroll
    for (int a = 0; a < 8; ++a) colbuf[0][ti + 16 * a] = accA[a][0];
  }
  if (tid < 128) rowbuf[0][tid] = (tid == 0) ? 1.0 : 0.0;        // row 0 of R = e_0
  __syncthreads();
  // everything the column steps do not write must read as zero afterwards (strict upper triangle beyond the
  // parked rows, rows and columns of the identity padding)
  for (int e = tid; e < 128 * POTF_LD; e += 256) S[e] = 0.0;
  if (tid < 128) { dl[tid] = 1.0; wdiag[tid] = 1.0; }
  __syncthreads();
  double ljj = 1.0, inv = 1.0;
  if (m > 0) potf_pivot(colbuf[0], 0, k, &bad, tid, ljj, inv);
  potf_fused_cols<0>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<1>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<2>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<3>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<4>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<5>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<6>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv) &&
      potf_fused_cols<7>(accA, accW, S, colbuf, rowbuf, dl, wdiag, &bad, k, m, ti, tc, tid, ljj, inv);
  __syncthreads();
  if (tid < 128) S[tid * POTF_LD + tid] = dl[tid];
  __syncthreads();
  double* Lo = bp.L + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  double* Uo = bp.U + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  double* Wo = bp.W + b * bp.blk_stride + static_cast<size_t>(k) * 128 * 128;
  double* Zo = bp.A + b * bp.mat_stride + (static_cast<size_t>(k) * 128) * ld + k * 128;
  for (int e = tid; e < 128 * 128; e += 256) {
    const int i = e >> 7, j = e & 127;
    double l, w;
    if (i > j) { l = S[i * POTF_LD + j]; w = S[j * POTF_LD + i]; }
    else if (i == j) { l = S[i * POTF_LD + i]; w = wdiag[i]; }
    else { l = 0.0; w = 0.0; }
    Lo[i * ld + j] = l;
    Wo[i * 128 + j] = w;
    Zo[i * ld + j] = w;
    double u;
    if (j > i) u = S[i * POTF_LD + j]; else if (i == j) u = wdiag[i]; else u = 0.0;
    Uo[i * ld + j] = u;
  }
  if (tid < 32) {
    double s = 0.0;
    for (int i = tid; i < 128; i += 32)
      if (k * 128 + i < n) s += log(S[i * POTF_LD + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (tid == 0) {
      logdet[b] += s;
      if (bad != 0 && info[b] == 0) info[b] = bad;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// 128x128-tile DMMA kernels of the blocked algorithms.  Each CTA: one output tile, 384 threads:
// warps 0-7 run the tensor pipe (232 registers), warps 8-11 are the producer warpgroup (40
// registers), whose elected lane streams 16-wide K-chunks of both operands through a ring of
// shared-memory stages with 3-D bulk tensor loads (column, row, batch member; 128-byte swizzle).
// ---------------------------------------------------------------------------------------------
constexpr int GEMM_SMEM = 4 * TILE_BYTES + 1024;          // plain-load main loop (mma_f64.cuh)
constexpr int FIT_STAGES = 4;
constexpr int FIT_SMEM = FIT_STAGES * 2 * TILE_BYTES + 1024 + 128;

__device__ __forceinline__ unsigned char* align1k(unsigned char* p) {
  return reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~static_cast<uintptr_t>(1023));
}

// Tensor maps of the batched fit workspace (built by the host when the workspace is allocated).
struct FitMaps { CUtensorMap A, L, U, T, W; };

// C[128][128] = sum over nch chunks of A[rowA0.., kA0 + 16c ..] * B[rowB0.., kB0 + 16c ..]^T for batch member b.
// The consumer threads then run epilogue(acc) (acc in the layout of acc_row / acc_col); the producer
// threads have nothing left to do.
template <class Epilogue>
__device__ __forceinline__ void tile_product_tma(const CUtensorMap* tmA, int rowA0, int kA0, const CUtensorMap* tmB,
                                                 int rowB0, int kB0, int nch, int b, unsigned char* smem_raw,
                                                 Epilogue epilogue) {
  unsigned char* smem = align1k(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = s0 + FIT_STAGES * 2 * TILE_BYTES;      // full[ST], empty[ST]
  if (tid == 0) {
    for (int s = 0; s < FIT_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (FIT_STAGES + s), 8); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
    if (warp == 8 && lane == 0) {
      for (int c = 0; c < nch; ++c) {
        const uint32_t s = c % FIT_STAGES, ph = (c / FIT_STAGES) & 1;
        mbar_wait(bar0 + 8 * (FIT_STAGES + s), ph ^ 1);
        const uint32_t full = bar0 + 8 * s;
        mbar_expect_tx(full, 2 * TILE_BYTES);
        tma_load_3d(s0 + s * 2 * TILE_BYTES, tmA, kA0 + c * KCH, rowA0, b, full);
        tma_load_3d(s0 + s * 2 * TILE_BYTES + TILE_BYTES, tmB, kB0 + c * KCH, rowB0, b, full);
      }
    }
    return;
  }
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
  WarpAcc acc;
  acc.zero();
  for (int c = 0; c < nch; ++c) {
    const uint32_t s = c % FIT_STAGES, ph = (c / FIT_STAGES) & 1;
    mbar_wait(bar0 + 8 * s, ph);
    mma_chunk_pf(acc, s0 + s * 2 * TILE_BYTES, s0 + s * 2 * TILE_BYTES + TILE_BYTES, warp, lane, 0);
    __syncwarp();
    if (lane == 0) mbar_arrive(bar0 + 8 * (FIT_STAGES + s));
  }
  epilogue(acc);
}

template <bool NEG, bool SUB>
__device__ __forceinline__ void store_tile(const WarpAcc& acc, double* C, size_t ldc) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int mf = 0; mf < 4; ++mf)
#pragma unroll
    for (int nf = 0; nf < 8; ++nf) {
      double2* p = reinterpret_cast<double2*>(C + static_cast<size_t>(acc_row(warp, lane, mf)) * ldc +
                                              acc_col(warp, lane, nf, 0));
      double2 v = make_double2(acc.c[mf][nf][0], acc.c[mf][nf][1]);
      if (NEG) { v.x = -v.x; v.y = -v.y; }
      if (SUB) { const double2 o = *p; v.x = o.x - v.x; v.y = o.y - v.y; }
      *p = v;
    }
}

// panel: L[i][k] = A[i][k] * W_k^T for i = k+1+blockIdx.x
__global__ void __launch_bounds__(384, 1) chol_panel_kernel(const __grid_constant__ CUtensorMap tmA,
                                                            const __grid_constant__ CUtensorMap tmW, BatchPtrs bp,
                                                            int k, int npad) {
  extern __shared__ unsigned char smem_raw[];
  const int b = blockIdx.y, i = k + 1 + blockIdx.x;
  const size_t ld = npad;
  double* out = bp.L + b * bp.mat_stride + (static_cast<size_t>(i) * 128) * ld + k * 128;
  tile_product_tma(&tmA, i * 128, k * 128, &tmW, k * 128, 0, 128 / KCH, b, smem_raw,
                   [&](const WarpAcc& acc) { store_tile<false, false>(acc, out, ld); });
}

__device__ __forceinline__ void tri_index(int p, int& ti, int& tj) {
  ti = static_cast<int>((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
  while ((ti + 1) * (ti + 2) / 2 <= p) ++ti;
  while (ti * (ti + 1) / 2 > p) --ti;
  tj = p - ti * (ti + 1) / 2;
}

// trailing update: A[i][j] -= sum over block columns kcol .. kcol+kw-1 of L[i][.] * L[j][.]^T for the tiles
// with jlo <= j < jhi and j <= i < nb.  The factorisation works in panels of a few block columns:
// inside a panel only the panel's own columns are updated after every block column (kw = 1), the rest
// of the trailing matrix once per panel with the whole panel as the contraction (kw = panel width) --
// fewer read-modify-write passes over A and longer runs of the tensor pipe per tile.
// Tiles are numbered column by column from (jlo, jlo); the launch covers tiles t0 .. t0 + gridDim.x - 1, which
// lets the host issue tile 0 -- the next diagonal block -- ahead of the rest.
__global__ void __launch_bounds__(384, 1) chol_syrk_kernel(const __grid_constant__ CUtensorMap tmL, BatchPtrs bp,
                                                           int kcol, int kw, int jlo, int jhi, int nb, int npad,
                                                           int t0) {
  extern __shared__ unsigned char smem_raw[];
  const int b = blockIdx.y;
  int j = jlo, t = blockIdx.x + t0;
  while (j < jhi && t >= nb - j) { t -= nb - j; ++j; }
  if (j >= jhi) return;
  const int i = j + t;
  const size_t ld = npad;
  double* out = bp.A + b * bp.mat_stride + (static_cast<size_t>(i) * 128) * ld + j * 128;
  tile_product_tma(&tmL, i * 128, kcol * 128, &tmL, j * 128, kcol * 128, kw * 128 / KCH, b, smem_raw,
                   [&](const WarpAcc& acc) { store_tile<false, true>(acc, out, ld); });
}

// Triangular inverse by recursive doubling.  With L = [[L11, 0], [L21, L22]] and both diagonal
// inverses known, the off-diagonal block of the inverse is Z21 = -Z22 L21 Z11, i.e. for U = L^-T
//   U12 = -U11 L21^T U22.
// Level s (in 128-blocks) joins every aligned pair of s-block inverses into a 2s-block inverse:
// two launches per level (T = U11 L21^T, then U12 = -T Z22^T), log2(nb) levels, every tile of a
// level in parallel.  Both U (upper) and Z = U^T (lower, kept in A's storage, which the
// factorisation no longer needs) are maintained so that every product is K-major on both sides.
// T is a full npad x ld scratch matrix per batch member.
__device__ __forceinline__ bool trtri_tile(int t, int s, int nb, int& left0, int& right0, int& a, int& c) {
  const int p = t / (s * s), rem = t - p * s * s;
  left0 = p * 2 * s; right0 = left0 + s;
  if (right0 >= nb) return false;
  a = rem / s; c = rem - a * s;
  return c < min(s, nb - right0);
}

// part 1: T[left0+a][right0+c] = sum_{m in left block, m >= a} U[left0+a][m] * L[right0+c][m]^T
__global__ void __launch_bounds__(384, 1) trtri_t_kernel(const __grid_constant__ CUtensorMap tmU,
                                                         const __grid_constant__ CUtensorMap tmL, BatchPtrs bp, int s,
                                                         int nb, int npad) {
  extern __shared__ unsigned char smem_raw[];
  const int b = blockIdx.y;
  int left0, right0, a, c;
  if (!trtri_tile(blockIdx.x, s, nb, left0, right0, a, c)) return;
  const size_t ld = npad;
  double* out = bp.T + b * bp.mat_stride + (static_cast<size_t>(left0 + a) * 128) * ld + (right0 + c) * 128;
  tile_product_tma(&tmU, (left0 + a) * 128, (left0 + a) * 128, &tmL, (right0 + c) * 128, (left0 + a) * 128,
                   (s - a) * 128 / KCH, b, smem_raw, [&](const WarpAcc& acc) { store_tile<false, false>(acc, out, ld); });
}

// part 2: U[left0+a][right0+c] = -sum_{q in right block, q <= c} T[left0+a][q] * Z[right0+c][q]^T, and Z = U^T
__global__ void __launch_bounds__(384, 1) trtri_u_kernel(const __grid_constant__ CUtensorMap tmT,
                                                         const __grid_constant__ CUtensorMap tmA, BatchPtrs bp, int s,
                                                         int nb, int npad) {
  extern __shared__ unsigned char smem_raw[];
  const int b = blockIdx.y;
  int left0, right0, a, c;
  if (!trtri_tile(blockIdx.x, s, nb, left0, right0, a, c)) return;
  const size_t ld = npad;
  double* Uo = bp.U + b * bp.mat_stride + (static_cast<size_t>(left0 + a) * 128) * ld + (right0 + c) * 128;
  // transposed copy into Z[right0+c][left0+a]
  double* Zt = bp.A + b * bp.mat_stride + (static_cast<size_t>(right0 + c) * 128) * ld + (left0 + a) * 128;
  tile_product_tma(&tmT, (left0 + a) * 128, right0 * 128, &tmA, (right0 + c) * 128, right0 * 128, (c + 1) * 128 / KCH, b,
                   smem_raw, [&](const WarpAcc& acc) {
    store_tile<true, false>(acc, Uo, ld);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int mf = 0; mf < 4; ++mf)
#pragma unroll
      for (int nf = 0; nf < 8; ++nf)
#pragma unroll
        for (int e = 0; e < 2; ++e)
          Zt[static_cast<size_t>(acc_col(warp, lane, nf, e)) * ld + acc_row(warp, lane, mf)] = -acc.c[mf][nf][e];
  });
}

// K^-1 = U U^T on the lower 128-tiles (diagonal tiles in full), written over A.
__global__ void __launch_bounds__(384, 1) kinv_kernel(const __grid_constant__ CUtensorMap tmU, BatchPtrs bp, int npad) {
  extern __shared__ unsigned char smem_raw[];
  const int b = blockIdx.y;
  int ta, tb; tri_index(blockIdx.x, ta, tb);
  const size_t ld = npad;
  double* out = bp.A + b * bp.mat_stride + (static_cast<size_t>(ta) * 128) * ld + tb * 128;
  tile_product_tma(&tmU, ta * 128, ta * 128, &tmU, tb * 128, ta * 128, (npad - ta * 128) / KCH, b, smem_raw,
                   [&](const WarpAcc& acc) { store_tile<false, false>(acc, out, ld); });
}

// w[m] = sum_{a <= m} U[a][m] y[a]   (w = L^-1 y).  block (32, 8): 32 columns per CTA.
__global__ void __launch_bounds__(256) uty_kernel(BatchPtrs bp, const double* __restrict__ y, int npad,
                                                  double* __restrict__ w) {
  __shared__ double red[8][33];
  const int b = blockIdx.y, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int m = blockIdx.x * 32 + tx;
  const double* U = bp.U + b * bp.mat_stride;
  double s = 0.0;
  for (int a = ty; a <= m; a += 8) s += U[static_cast<size_t>(a) * npad + m] * y[a];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0) {
    double t = 0.0;
    for (int r = 0; r < 8; ++r) t += red[r][tx];
    w[static_cast<size_t>(b) * npad + m] = t;
  }
}

// alpha[a] = sum_{m >= a} U[a][m] w[m]   (alpha = L^-T w).  One warp per row.
__global__ void __launch_bounds__(256) uw_kernel(BatchPtrs bp, const double* __restrict__ w, int npad,
                                                 double* __restrict__ alpha) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int a = blockIdx.x * 8 + (threadIdx.x >> 5);
  const double* U = bp.U + b * bp.mat_stride + static_cast<size_t>(a) * npad;
  const double* wb = w + static_cast<size_t>(b) * npad;
  double s = 0.0;
  for (int m = (a & ~31) + lane; m < npad; m += 32)
    if (m >= a) s += U[m] * wb[m];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) alpha[static_cast<size_t>(b) * npad + a] = s;
}

// lml[b] = -1/2 y^T alpha - logdet - n/2 log(2 pi);  grad[b] = 1/2 * sum(partials)
__global__ void lml_finalize_kernel(const double* __restrict__ y, const double* __restrict__ alpha,
                                    const double* __restrict__ logdet, const int* __restrict__ info,
                                    const double* __restrict__ partial, int ntri, int n, int npad,
                                    double* __restrict__ lml, double* __restrict__ grad) {
  __shared__ double red[2][8];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double ya = 0.0, gs = 0.0;
  for (int i = tid; i < n; i += 256) ya += y[i] * alpha[static_cast<size_t>(b) * npad + i];
  if (partial != nullptr)
    for (int i = tid; i < ntri; i += 256) gs += partial[static_cast<size_t>(b) * ntri + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ya += __shfl_xor_sync(0xffffffffu, ya, o);
    gs += __shfl_xor_sync(0xffffffffu, gs, o);
  }
  if (lane == 0) { red[0][warp] = ya; red[1][warp] = gs; }
  __syncthreads();
  if (tid == 0) {
    double a = 0.0, g = 0.0;
    for (int w = 0; w < 8; ++w) { a += red[0][w]; g += red[1][w]; }
    if (info[b] != 0) { lml[b] = -INFINITY; if (grad) grad[b] = 0.0; }
    else {
      lml[b] = -0.5 * a - logdet[b] - 0.5 * n * 1.8378770664093453;   // log(2 pi)
      if (grad) grad[b] = 0.5 * g;
    }
  }
}

// Z = U^T (row-major L^-1) for the sweep; 32x32 tiles through shared memory.
__global__ void __launch_bounds__(256) transpose_kernel(const double* __restrict__ U, double* __restrict__ Z, int npad) {
  __shared__ double tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int bi = blockIdx.y * 32, bj = blockIdx.x * 32;
  for (int r = ty; r < 32; r += 8) tile[r][tx] = U[static_cast<size_t>(bi + r) * npad + bj + tx];
  __syncthreads();
  for (int r = ty; r < 32; r += 8) Z[static_cast<size_t>(bj + r) * npad + bi + tx] = tile[tx][r];
}

}  // namespace plx
