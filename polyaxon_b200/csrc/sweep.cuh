// The acquisition sweep: for every candidate x, the Gaussian-process posterior mean and variance
// at a fixed GP state, the UCB / EI / POI value, and the arg-max over all candidates.
//
// Reference path being replaced: UtilityFunction.max_compute warm-up
// (acquisition_function.py:88-92) -> compute (:48-68) -> sklearn GaussianProcessRegressor.predict
// (_gpr.py:446-500): K* = k(Xc, X); mean = K* alpha; V = L^-1 K*^T; var = 1 - colsum(V^2),
// negatives clamped to 0; std = sqrt(var).
//
// One persistent CTA per SM walks candidate tiles of 128:
//   phase A  cross-kernel tile K*[128][n16] -> this CTA's private scratch slab in HBM/L2
//            (K-major, the layout the tensor pipe wants), posterior mean accumulated on the way;
//   phase B  V^T = K* . Z^T with Z = L^-1 (lower triangular, so block row ib only needs
//            k < (ib+1)*128): FP64 DMMA, Z tiles and K* tiles staged in shared memory by TMA
//            (128-byte swizzle) through an mbarrier ring; squares of V summed per candidate in
//            registers -- V is never written anywhere;
//   epilogue variance -> std -> acquisition -> running arg-max (numpy semantics: first maximum,
//            NaN wins).
// Algorithmic work per candidate: n*(3d+10) + n^2 + 4n flops (triangular variance), 8d bytes read.
#pragma once
#include "gp_kernels.cuh"

namespace plx {

enum AcqKind { ACQ_UCB = 0, ACQ_EI = 1, ACQ_POI = 2 };

// Work-skipping switches for phase timing exist only in builds made with -DPLX_PROFILE (python -m polyaxon_b200.build
// --profile); in the shipped library the tests below are constant false and the option is refused.
#ifdef PLX_PROFILE
#define PLX_DBG(flags, bit) (((flags) & (bit)) != 0)
#else
#define PLX_DBG(flags, bit) false
#endif

struct SweepParams {
  const double* Xc;        // candidates [N][d], device
  long long N;             // candidates in this launch
  long long idx_base;      // global index of candidate 0 of this launch
  int d, n, npad, n16;     // observations, padded to 128, padded to 16
  const double* X;         // observations [n][d]
  const double* Xt;        // X / length_scale transposed to [d][npad] (TMA-ring kernel only)
  const double* xrow;      // nullable: centred scaled observations, row-major [npad][40] (32 dimensions + padding); when set
                           // (d <= 32) the int8 builders form the squared distances on the FP64 tensor path (sweep_i8f.cuh)
  const double* xnc;       //   |row|^2 of xrow [npad]
  const double* xmean;     //   the centre: mean of X / length_scale per dimension [32]
  const double* alpha;     // [npad]
  const double* Z;         // L^-1, row-major [npad][npad]
  double ell;              // length scale
  int kind, acq;
  MaternGen mg;            // constants of the general-nu Matern kernel (kind == K_GEN)
  double kappa, eps, y_max;
  double* Ks;              // scratch [gridDim.x * 128][npad]
  signed char* planes;     // variant 2: digit planes of K*, [I8_S][plane_rows][npad] int8 (see sweep_i8.cuh)
  size_t plane_stride;     //            bytes between planes
  double* out_mean; double* out_std; double* out_acq;   // optional per-candidate dumps [N]
  double* blk_val; long long* blk_idx;                  // per-CTA best
  unsigned long long* clamped;                          // nullable: counts negative variances set to zero (sklearn _gpr.py:485-491 warns)
  int dbg;                 // read only in PLX_PROFILE builds (tools/*_phases.py): bits skip parts of the work to time the rest
};

struct Best { double v; long long i; };

// numpy.argmax order: NaN beats everything, then larger value, then lower index
__device__ __forceinline__ bool better(double av, long long ai, double bv, long long bi) {
  const bool an = isnan(av), bn = isnan(bv);
  if (an != bn) return an;
  if (an) return ai < bi;
  if (av != bv) return av > bv;
  return ai < bi;
}

// scipy.special.ndtr (Cephes): erf branch for |z|/sqrt(2) < 1, erfc branch otherwise
__device__ __forceinline__ double norm_cdf(double z) {
  if (isnan(z)) return z;
  const double x = z * 0.7071067811865476;
  const double ax = fabs(x);
  if (ax < 1.0) return 0.5 + 0.5 * erf(x);
  const double y = 0.5 * erfc(ax);
  return (x > 0.0) ? 1.0 - y : y;
}
__device__ __forceinline__ double norm_pdf(double z) { return exp(-z * z / 2.0) / 2.5066282746310002; }

// acquisition_function.py:48-61 -- no guard for std == 0 (inf / NaN propagate as in numpy)
__device__ __forceinline__ double acq_value(int acq, double mean, double sd, double y_max, double kappa, double eps) {
  if (acq == ACQ_UCB) return mean + kappa * sd;
  const double imp = mean - y_max - eps;
  const double z = imp / sd;
  if (acq == ACQ_EI) return imp * norm_cdf(z) + sd * norm_pdf(z);
  return norm_cdf(z);
}

// ------------------------------------------------------------------------------------------
// Phase A: K*[c][j] for the 128 candidates of a tile against all observations, into the CTA's
// scratch slab (row pitch npad), plus the posterior mean.  256 threads; thread (tx = tid & 15,
// ty = tid >> 4) owns candidates ty*8 .. ty*8+7 and observations tx + 16*jj of each 64-block.
// Shared staging (aliases the pipeline stages): Cs[32 dims][130], Xs[32 dims][65].
// ------------------------------------------------------------------------------------------
constexpr int CS_LD = 130, XS_LD = 65;

__device__ __forceinline__ void phase_a(const SweepParams& p, long long tile, double* Kslab, double* smem_d,
                                        double* mean_s, int tid) {
  double* Cs = smem_d;
  double* Xs = smem_d + 32 * CS_LD;
  const int tx = tid & 15, ty = tid >> 4;
  const long long c_base = tile * 128;
  const int nchunk_d = (p.d + 31) / 32;
  double mu[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) mu[i] = 0.0;

  for (int jb = 0; jb * 64 < p.n16; ++jb) {
    double r2[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r2[i][j] = 0.0;
    for (int dc = 0; dc < nchunk_d; ++dc) {
      const int d0 = dc * 32, dn = min(32, p.d - d0);
      named_bar_sync(1, 256);
      if (jb == 0 || nchunk_d > 1) {
        for (int e = tid; e < 128 * 32; e += 256) {
          const int c = e >> 5, q = e & 31;
          double v = 0.0;
          if (q < dn && c_base + c < p.N) v = p.Xc[(c_base + c) * p.d + d0 + q] / p.ell;
          Cs[q * CS_LD + c] = v;
        }
      }
      for (int e = tid; e < 64 * 32; e += 256) {
        const int j = e >> 5, q = e & 31;
        const int jo = jb * 64 + j;
        double v = 0.0;
        if (q < dn && jo < p.n) v = p.X[static_cast<size_t>(jo) * p.d + d0 + q] / p.ell;
        Xs[q * XS_LD + j] = v;
      }
      named_bar_sync(1, 256);
      for (int q = 0; q < dn; ++q) {
        double cv[8], xv[4];
        const double2* cp = reinterpret_cast<const double2*>(Cs + q * CS_LD + ty * 8);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const double2 t2 = cp[i]; cv[2 * i] = t2.x; cv[2 * i + 1] = t2.y; }
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[j] = Xs[q * XS_LD + tx + 16 * j];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) { const double df = cv[i] - xv[j]; r2[i][j] += df * df; }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int jo = jb * 64 + tx + 16 * j;
      if (jo < p.n16) {
        const bool real = jo < p.n;
        const double aj = real ? p.alpha[jo] : 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const double kv = real ? kern_value(p.kind, r2[i][j], p.mg) : 0.0;
          mu[i] += kv * aj;
          Kslab[static_cast<size_t>(ty * 8 + i) * p.npad + jo] = kv;
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double m = mu[i];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
    if (tx == 0) mean_s[ty * 8 + i] = m;
  }
}

// After phase B: fold each thread's per-row sums of squares into part[wn][row].
__device__ __forceinline__ void fold_rowsums(const double (&sv)[4], double* part, int warp, int lane) {
#pragma unroll
  for (int mf = 0; mf < 4; ++mf) {
    double s = sv[mf];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if ((lane & 3) == 0) part[(warp >> 2) * 128 + acc_row(warp, lane, mf)] = s;
  }
}

__device__ __forceinline__ void add_squares(double (&sv)[4], const WarpAcc& acc) {
#pragma unroll
  for (int mf = 0; mf < 4; ++mf)
#pragma unroll
    for (int nf = 0; nf < 8; ++nf)
      sv[mf] += acc.c[mf][nf][0] * acc.c[mf][nf][0] + acc.c[mf][nf][1] * acc.c[mf][nf][1];
}

// Epilogue for candidate c = tid (< 128) of the tile.
__device__ __forceinline__ void tile_epilogue(const SweepParams& p, long long tile, const double* part,
                                              const double* mean_s, int c, Best& best) {
  const long long ci = tile * 128 + c;
  if (ci >= p.N) return;
  double var = 1.0 - (part[c] + part[128 + c]);
  if (var < 0.0) { var = 0.0; if (p.clamped) atomicAdd(p.clamped, 1ULL); }   // sklearn clamps negative variances to zero, and says so
  const double sd = sqrt(var), mean = mean_s[c];
  const double a = acq_value(p.acq, mean, sd, p.y_max, p.kappa, p.eps);
  if (p.out_mean) p.out_mean[ci] = mean;
  if (p.out_std) p.out_std[ci] = sd;
  if (p.out_acq) p.out_acq[ci] = a;
  const long long gi = p.idx_base + ci;
  if (better(a, gi, best.v, best.i)) { best.v = a; best.i = gi; }
}

__device__ __forceinline__ void block_best(const SweepParams& p, Best best, double* sval, long long* sidx, int tid) {
  // 128 candidates' running bests -> one per CTA (thread 0 scans in index order: deterministic)
  if (tid < 128) { sval[tid] = best.v; sidx[tid] = best.i; }
  named_bar_sync(1, 256);
  if (tid == 0) {
    Best b = {sval[0], sidx[0]};
    for (int t = 1; t < 128; ++t)
      if (better(sval[t], sidx[t], b.v, b.i)) { b.v = sval[t]; b.i = sidx[t]; }
    p.blk_val[blockIdx.x] = b.v;
    p.blk_idx[blockIdx.x] = b.i;
  }
}

// number of leading N fragments of warp slab `wn` that are structurally zero for K-chunk j0 in
// block row ib (Z is lower triangular: Z[i][j] = 0 for j > i)
__device__ __forceinline__ int tri_skip(int j0, int ib, int wn) {
  const int q = (j0 - ib * 128 - wn) >> 3;
  return q < 0 ? 0 : (q > 8 ? 8 : q);
}

// ------------------------------------------------------------------------------------------
// Phase A for the TMA-ring kernel.  Same arithmetic as phase_a, organised so that the code stays
// small (the straight-line version is ~11k instructions and thrashes the instruction cache):
//   1. distances: thread (tx = tid & 15, ty = tid >> 4) accumulates r^2 for 8 candidates x 4
//      observations of a 64-observation block from shared tiles (observations come pre-scaled and
//      transposed, Xt[q][j], the next block is prefetched into registers meanwhile) and parks the 32
//      values in a shared tile R2s[128][65];
//   2. kernel values: warp w walks candidates 16w .. 16w+15 in a compact loop, lane = observation
//      (two per lane): k(r^2), coalesced store into the K* slab, and the posterior-mean partial into
//      the thread's own slot of Macc[128][32] (deterministic, no atomics).
// Shared staging (aliases the pipeline stages): Cs[32][130] | Xs[32][65] | R2s[128][65] | Macc[128][32].
// ------------------------------------------------------------------------------------------
constexpr int R2_LD = 65;
constexpr int PA_CS = 0, PA_XS = 32 * CS_LD, PA_R2 = PA_XS + 32 * XS_LD, PA_MACC = PA_R2 + 128 * R2_LD;
constexpr int PA_DOUBLES = PA_MACC + 128 * 32;
static_assert(PA_DOUBLES * 8 <= 5 * 2 * TILE_BYTES, "phase A staging must fit in the pipeline stages");

// Digit planes of the int8-sliced sweep (sweep_i8.cuh): I8_S balanced base-256 digits in [-128, 127] per value.
constexpr int I8_S = 7;

// Balanced base-256 digits of a 56-bit fixed-point integer q (|q| < 2^54), least significant first; the leading
// digit takes what is left (|.| <= 65).  Integer arithmetic only: exact.
__device__ __forceinline__ void store_digits_q(signed char* dst, size_t plane_stride, long long q) {
#pragma unroll
  for (int s = I8_S - 1; s >= 1; --s) {
    const int dgt = static_cast<int>((q + 128) & 255) - 128;
    dst[s * plane_stride] = static_cast<signed char>(dgt);
    q = (q - dgt) >> 8;
  }
  dst[0] = static_cast<signed char>(q);
}

// Kernel value k in [0, 1]: k = 4 * sum_s a_s 256^-(s+1) with the digits of round(k * 2^54) (|residual| <= 2^-55).
__device__ __forceinline__ void store_digits(signed char* dst, size_t plane_stride, double k, int dbg = 0) {
  if (PLX_DBG(dbg, 16)) return;
  store_digits_q(dst, plane_stride, __double2ll_rn(scale_2p54(k)));
}

// SL = false: FP64 kernel values into the K* slab; SL = true: their digit planes (Kslab then points at the
// tile's first row of plane 0, reinterpreted).
template <int KIND, bool SL>
__device__ __forceinline__ void kern_rows(const SweepParams& p, uint32_t sm, double* Kslab, int jb, int warp, int lane) {
  const int j0 = jb * 64 + lane, j1 = j0 + 32;
  const bool real0 = j0 < p.n, real1 = j1 < p.n;
  const double a0 = real0 ? p.alpha[j0] : 0.0, a1 = real1 ? p.alpha[j1] : 0.0;
  const bool st0 = j0 < p.n16, st1 = j1 < p.n16;
  if constexpr (KIND == K_GEN) {
    for (int ci = 0; ci < 16; ++ci) {
      const int c = warp * 16 + ci;
      const uint32_t r2a = sm + static_cast<uint32_t>((PA_R2 + c * R2_LD + lane) * 8);
      double k0 = kern_value_t<KIND>(lds64(r2a), p.mg);
      double k1 = kern_value_t<KIND>(lds64(r2a + 32 * 8), p.mg);
      k0 = real0 ? k0 : 0.0;
      k1 = real1 ? k1 : 0.0;
      if constexpr (SL) {
        signed char* row = reinterpret_cast<signed char*>(Kslab) + static_cast<size_t>(c) * p.npad;
        if (st0) store_digits(row + j0, p.plane_stride, k0, p.dbg);
        if (st1) store_digits(row + j1, p.plane_stride, k1, p.dbg);
      } else {
        double* row = Kslab + static_cast<size_t>(c) * p.npad;
        if (st0) row[j0] = k0;
        if (st1) row[j1] = k1;
      }
      const uint32_t ma = sm + static_cast<uint32_t>((PA_MACC + c * 32 + lane) * 8);
      sts64(ma, lds64(ma) + (k0 * a0 + k1 * a1));
    }
  } else {
  // closed forms: four candidates (eight values) per step, evaluated in lock step (see kern_values_n)
#pragma unroll 1
  for (int ci = 0; ci < 16; ci += 4) {
    double v[8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t r2a = sm + static_cast<uint32_t>((PA_R2 + (warp * 16 + ci + u) * R2_LD + lane) * 8);
      v[2 * u] = lds64(r2a);
      v[2 * u + 1] = lds64(r2a + 32 * 8);
    }
    kern_values_n<KIND, 8>(v);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = warp * 16 + ci + u;
      const double k0 = real0 ? v[2 * u] : 0.0, k1 = real1 ? v[2 * u + 1] : 0.0;
      if constexpr (SL) {
        signed char* row = reinterpret_cast<signed char*>(Kslab) + static_cast<size_t>(c) * p.npad;
        if (st0) store_digits(row + j0, p.plane_stride, k0, p.dbg);
        if (st1) store_digits(row + j1, p.plane_stride, k1, p.dbg);
      } else {
        double* row = Kslab + static_cast<size_t>(c) * p.npad;
        if (st0) row[j0] = k0;
        if (st1) row[j1] = k1;
      }
      const uint32_t ma = sm + static_cast<uint32_t>((PA_MACC + c * 32 + lane) * 8);
      sts64(ma, lds64(ma) + (k0 * a0 + k1 * a1));
    }
  }
  }
}

template <bool SL>
__device__ __forceinline__ void phase_a3(const SweepParams& p, long long tile, double* Kslab, uint32_t sm,
                                         double* mean_s, int tid) {
  const int tx = tid & 15, ty = tid >> 4, warp = tid >> 5, lane = tid & 31;
  const long long c_base = tile * 128;
  const int nchunk_d = (p.d + 31) / 32;
  const int nblk = (p.n16 + 63) / 64;

  for (int e = tid; e < 128 * 32; e += 256) sts64(sm + static_cast<uint32_t>((PA_MACC + e) * 8), 0.0);

  double pre[8];                      // element e = tid + 256 r of a 32 x 64 block: q = e >> 6, j = e & 63
  auto fetch = [&](int jb, int dc) {
    const int d0 = dc * 32, dn = min(32, p.d - d0);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int e = tid + 256 * r, q = e >> 6, j = e & 63;
      pre[r] = (q < dn) ? p.Xt[static_cast<size_t>(d0 + q) * p.npad + jb * 64 + j] : 0.0;
    }
  };

  fetch(0, 0);
  for (int jb = 0; jb < nblk; ++jb) {
    double r2[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r2[i][j] = 0.0;
    for (int dc = 0; dc < nchunk_d; ++dc) {
      const int d0 = dc * 32, dn = min(32, p.d - d0);
      if (dc > 0) named_bar_sync(1, 256);                   // readers of the previous dimension chunk are done
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int e = tid + 256 * r;
        sts64(sm + static_cast<uint32_t>((PA_XS + (e >> 6) * XS_LD + (e & 63)) * 8), pre[r]);
      }
      if (jb == 0 || nchunk_d > 1) {
        for (int e = tid; e < 128 * 32; e += 256) {
          const int c = e >> 5, q = e & 31;
          double v = 0.0;
          if (q < dn && c_base + c < p.N) v = p.Xc[(c_base + c) * p.d + d0 + q] / p.ell;
          sts64(sm + static_cast<uint32_t>((PA_CS + q * CS_LD + c) * 8), v);
        }
      }
      named_bar_sync(1, 256);
      {                                                       // prefetch the next (jb, dc)
        int njb = jb, ndc = dc + 1;
        if (ndc == nchunk_d) { ndc = 0; ++njb; }
        if (njb < nblk) fetch(njb, ndc);
      }
      const uint32_t cs = sm + static_cast<uint32_t>((PA_CS + ty * 8) * 8);
      const uint32_t xs = sm + static_cast<uint32_t>((PA_XS + tx) * 8);
#pragma unroll 2
      for (int q = 0; q < (PLX_DBG(p.dbg, 8) ? 1 : dn); ++q) {
        double cv[8], xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { const double2 t2 = lds128(cs + (q * CS_LD + 2 * i) * 8); cv[2 * i] = t2.x; cv[2 * i + 1] = t2.y; }
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[j] = lds64(xs + (q * XS_LD + 16 * j) * 8);
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) { const double df = cv[i] - xv[j]; r2[i][j] += df * df; }
      }
    }
    // park r^2 (previous block's readers passed the barrier above before anyone got here)
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        sts64(sm + static_cast<uint32_t>((PA_R2 + (ty * 8 + i) * R2_LD + tx + 16 * j) * 8), r2[i][j]);
    named_bar_sync(1, 256);
    if (!PLX_DBG(p.dbg, 4)) switch (p.kind) {
      case K_RBF: kern_rows<K_RBF, SL>(p, sm, Kslab, jb, warp, lane); break;
      case K_M05: kern_rows<K_M05, SL>(p, sm, Kslab, jb, warp, lane); break;
      case K_M15: kern_rows<K_M15, SL>(p, sm, Kslab, jb, warp, lane); break;
      case K_GEN: kern_rows<K_GEN, SL>(p, sm, Kslab, jb, warp, lane); break;
      default: kern_rows<K_M25, SL>(p, sm, Kslab, jb, warp, lane); break;
    }
    // no barrier here: the next block re-writes Xs (free since the barrier above) and parks into R2s only
    // after its own barrier, which every thread reaches after finishing this kern_rows
  }
  named_bar_sync(1, 256);
  if (tid < 128) {
    double m = 0.0;
    for (int l = 0; l < 32; ++l) m += lds64(sm + static_cast<uint32_t>((PA_MACC + tid * 32 + l) * 8));
    mean_s[tid] = m;
  }
}

// ------------------------------------------------------------------------------------------
// Variant 1: TMA + mbarrier ring.  384 threads: warps 0-7 (two warpgroups) compute with 232
// registers each, warps 8-11 form the producer warpgroup (40 registers; one elected lane issues
// every TMA load).  No CTA-wide barrier inside the tile loop: the roles meet only through
// mbarriers (slab ready -> producer; stage full -> consumers; stage empty -> producer).
// ------------------------------------------------------------------------------------------
constexpr int SWEEP_STAGES = 5;

struct SweepTail {                 // small per-CTA arrays behind the pipeline stages
  double mean_s[128];              // posterior mean of the tile's candidates
  double part[256];                // [2][128] sums of squares from the two N slabs
  double sval[128];                // block arg-max scratch
  long long sidx[128];
  unsigned long long bars[2 * SWEEP_STAGES + 1];   // full[ST], empty[ST], slab
};
constexpr int SWEEP_SMEM = SWEEP_STAGES * 2 * TILE_BYTES + 1024 /*align*/ + 8192 /*SweepTail*/;
static_assert(sizeof(SweepTail) <= 8192, "tail budget");

__global__ void __launch_bounds__(384, 1) sweep_tma_kernel(const __grid_constant__ CUtensorMap tmZ,
                                                           const __grid_constant__ CUtensorMap tmK,
                                                           const SweepParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  SweepTail* tail = reinterpret_cast<SweepTail*>(smem + SWEEP_STAGES * 2 * TILE_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = smem_u32(tail->bars);
  const uint32_t bar_slab = bar0 + 8 * (2 * SWEEP_STAGES);
  if (tid == 0) {
    for (int s = 0; s < SWEEP_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (SWEEP_STAGES + s), 8); }
    mbar_init(bar_slab, 8);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  const long long ntiles = (p.N + 127) / 128;
  const int nib = (p.n + 127) / 128;
  uint32_t it = 0;                         // ring position, continues across tiles
  uint32_t tcount = 0;                     // tiles processed by this CTA (slab barrier phase)

  if (warp >= 8) {
    // ===== producer warpgroup =====
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
    if (warp == 8 && lane == 0) {
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
        mbar_wait(bar_slab, tcount & 1);   // K* slab of this tile is complete and visible to TMA
        for (int ib = 0; ib < nib; ++ib) {
          const int kend = min((ib + 1) * 128, p.n16);
          for (int j0 = 0; j0 < kend; j0 += KCH, ++it) {
            const uint32_t s = it % SWEEP_STAGES, ph = (it / SWEEP_STAGES) & 1;
            mbar_wait(bar0 + 8 * (SWEEP_STAGES + s), ph ^ 1);
            const uint32_t full = bar0 + 8 * s;
            mbar_expect_tx(full, 2 * TILE_BYTES);
            tma_load_2d(s0 + s * 2 * TILE_BYTES, &tmK, j0, blockIdx.x * 128, full);
            tma_load_2d(s0 + s * 2 * TILE_BYTES + TILE_BYTES, &tmZ, j0, ib * 128, full);
          }
        }
      }
    }
    return;
  }

  // ===== consumer warpgroups =====
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
  double* Kslab = p.Ks + static_cast<size_t>(blockIdx.x) * 128 * p.npad;
  Best best = {-INFINITY, 0x7fffffffffffffffLL};
  const int wn = (warp >> 2) * 64;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    // every TMA load of the previous tile has been waited for, so the stage memory is free
    if (!PLX_DBG(p.dbg, 1) || tile == blockIdx.x) phase_a3<false>(p, tile, Kslab, s0, tail->mean_s, tid);
    __threadfence();
    asm volatile("fence.proxy.async;\n" ::: "memory");     // generic-proxy stores -> visible to TMA
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_slab);

    double sv[4] = {0.0, 0.0, 0.0, 0.0};
    WarpAcc acc;
    for (int ib = 0; ib < nib; ++ib) {
      const int kend = min((ib + 1) * 128, p.n16);
      acc.zero();
      for (int j0 = 0; j0 < kend; j0 += KCH, ++it) {
        const uint32_t s = it % SWEEP_STAGES, ph = (it / SWEEP_STAGES) & 1;
        mbar_wait(bar0 + 8 * s, ph);
        const int skip = tri_skip(j0, ib, wn);
        if (skip < 8 && !PLX_DBG(p.dbg, 2)) mma_chunk_pf(acc, s0 + s * 2 * TILE_BYTES, s0 + s * 2 * TILE_BYTES + TILE_BYTES, warp, lane, skip);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar0 + 8 * (SWEEP_STAGES + s));
      }
      add_squares(sv, acc);
    }
    fold_rowsums(sv, tail->part, warp, lane);
    named_bar_sync(1, 256);
    if (tid < 128) tile_epilogue(p, tile, tail->part, tail->mean_s, tid, best);
    named_bar_sync(1, 256);                // part / mean_s free for the next tile
  }
  block_best(p, best, tail->sval, tail->sidx, tid);
}

// ------------------------------------------------------------------------------------------
// Variant 0: same algorithm, operands fetched with plain coherent loads (register-prefetched
// double buffer).  256 threads.  Kept as the cross-check for the TMA ring.
// ------------------------------------------------------------------------------------------
constexpr int SWEEP_SIMPLE_SMEM = 4 * TILE_BYTES + 1024 + 8192;

__global__ void __launch_bounds__(256, 1) sweep_simple_kernel(const SweepParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  SweepTail* tail = reinterpret_cast<SweepTail*>(smem + 4 * TILE_BYTES);
  double* mean_s = tail->mean_s;
  double* part = tail->part;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long ntiles = (p.N + 127) / 128;
  const int nib = (p.n + 127) / 128;
  double* Kslab = p.Ks + static_cast<size_t>(blockIdx.x) * 128 * p.npad;
  Best best = {-INFINITY, 0x7fffffffffffffffLL};
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    phase_a(p, tile, Kslab, reinterpret_cast<double*>(smem), mean_s, tid);
    __syncthreads();
    double sv[4] = {0.0, 0.0, 0.0, 0.0};
    WarpAcc acc;
    for (int ib = 0; ib < nib; ++ib) {
      acc.zero();
      mainloop_nt(acc, Kslab, p.npad, p.Z + static_cast<size_t>(ib) * 128 * p.npad, p.npad, 0,
                  min((ib + 1) * 128, p.n16), smem);
      add_squares(sv, acc);
    }
    fold_rowsums(sv, part, warp, lane);
    __syncthreads();
    if (tid < 128) tile_epilogue(p, tile, part, mean_s, tid, best);
    __syncthreads();
  }
  block_best(p, best, tail->sval, tail->sidx, tid);
}

// X / length_scale, transposed to [d][npad] (zero in the padding), for the builders' coalesced reads
__global__ void scale_transpose_x_kernel(const double* __restrict__ X, int n, int d, int npad, double ell,
                                         double* __restrict__ Xt) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, q = blockIdx.y;
  if (j >= npad) return;
  Xt[static_cast<size_t>(q) * npad + j] = (j < n) ? X[static_cast<size_t>(j) * d + q] / ell : 0.0;
}

// Centre of the scaled observations per dimension (one CTA per dimension, a fixed-shape tree sum: deterministic).
__global__ void __launch_bounds__(256) obs_centre_kernel(const double* __restrict__ X, int n, int d, double ell,
                                                         double* __restrict__ xmean) {
  __shared__ double part[256];
  const int q = blockIdx.x, tid = threadIdx.x;
  double s = 0.0;
  if (q < d)
    for (int j = tid; j < n; j += 256) s += X[static_cast<size_t>(j) * d + q] / ell;
  part[tid] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) part[tid] += part[tid + o];
    __syncthreads();
  }
  if (tid == 0) xmean[q] = (q < d) ? part[0] / n : 0.0;
}

// Rows of centred scaled observations for the tensor-path builders: xrow[j][0..31] = X_j / l - centre (zero beyond d
// and beyond n), columns 32..39 zero (the row stride of 40 doubles is what makes the builders' 16-byte fragment loads
// conflict-free); xnc[j] = |row j|^2.
constexpr int XROW_LD = 40;
__global__ void obs_rows_kernel(const double* __restrict__ X, int n, int d, int npad, double ell,
                                const double* __restrict__ xmean, double* __restrict__ xrow, double* __restrict__ xnc) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= npad) return;
  double s = 0.0;
  double* row = xrow + static_cast<size_t>(j) * XROW_LD;
  for (int q = 0; q < XROW_LD; ++q) {
    double u = 0.0;
    if (q < d && j < n) u = X[static_cast<size_t>(j) * d + q] / ell - xmean[q];
    row[q] = u;
    s = fma(u, u, s);
  }
  xnc[j] = s;
}

// Merge the per-CTA bests into the running best of the whole sweep (one thread; grid <= #SMs).
__global__ void sweep_merge_kernel(const double* __restrict__ blk_val, const long long* __restrict__ blk_idx,
                                   int nblk, double* best_val, long long* best_idx) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double v = *best_val; long long i = *best_idx;
  for (int b = 0; b < nblk; ++b)
    if (better(blk_val[b], blk_idx[b], v, i)) { v = blk_val[b]; i = blk_idx[b]; }
  *best_val = v; *best_idx = i;
}

}  // namespace plx
