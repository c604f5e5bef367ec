// Fused form of the int8-sliced sweep (sweep_i8.cuh): one persistent CTA per SM in which the digit planes of the
// NEXT candidate tile are built on the FP64 pipe while the tensor pipe contracts the CURRENT tile.
//
// 512 threads, four warpgroups:
//   WG0, WG1 (warps 0-7)   builders: scaled squared distances for 8 candidates x 4 consecutive observations per
//                          thread from shared tiles, kernel values straight from those registers (no staging of the
//                          distance block), balanced 8-bit digits packed four to a 32-bit store into the CTA's own
//                          plane buffer in global memory (two tiles deep), posterior mean reduced with shuffles;
//   WG2 (warps 8-11)       accumulator drain + epilogue, one thread per candidate (= TMEM lane);
//   WG3 (warp 12, 13)      TMA producer and tensor-instruction issuer, one elected lane each.
// The roles meet only through mbarriers: planes ready (builders -> producer, drain), planes free (drain ->
// builders), stage full / empty, accumulators full / empty.  Arithmetic is that of kstar_slices_kernel +
// sweep_i8_kernel; only the summation order of the posterior mean differs (per-thread partial sums over the
// observations 4 tx .. 4 tx + 3 of every 64-block, then a 16-lane butterfly).
#pragma once
#include "sweep_i8.cuh"

namespace plx {

constexpr int F_CS_LD = 130;                               // Cs[32 dims][130]: scaled candidates of the tile
constexpr int F_XS_LD = 64;                                // Xs[32 dims][64]: scaled observations of a 64-block
constexpr int F_CS = 0, F_XS = 32 * F_CS_LD;
// Tensor-path builders (build_tile_mma): Xr[64 observations][40] (centred scaled rows as obs_rows_kernel lays them out),
// then |row|^2 and alpha of the 64 observations.  The larger of the two layouts sizes the staging area.
constexpr int F_XR_LD = XROW_LD;
// |row|^2 and alpha sit at o + 2 (o >> 4): 16 bytes of padding per 128, so that the four threads of a quad, which
// read eight consecutive values each (64 bytes apart), fall on distinct banks.
constexpr int F_XN = F_XS + 64 * F_XR_LD, F_AL = F_XN + 72;
constexpr int F_STAGING = (F_AL + 72) * 8;                 // 54 912 bytes
static_assert(F_AL + 72 >= F_XS + 32 * F_XS_LD, "staging area holds either layout");
__device__ __forceinline__ int xn_pos(int o) { return o + 2 * (o >> 4); }
constexpr int I8F_SMEM = I8_STAGES * I8_STAGE + F_STAGING + 1024 /*align*/ + 3072 /*tail*/;

struct I8FTail {
  unsigned long long bars[2 * I8_STAGES + 2 + 4];   // full[ST], empty[ST], acc full, acc empty, planes ready[2], planes free[2]
  uint32_t tmem_base;
  double sval[128];
  long long sidx[128];
};
static_assert(sizeof(I8FTail) <= 3072, "tail budget");

// Four kernel values of one candidate -> 7 planes x 4 packed digits (one 32-bit store per plane).
__device__ __forceinline__ void store_digits4(signed char* dst, size_t plane_stride, const double (&k)[4]) {
  long long q[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) q[j] = __double2ll_rn(scale_2p54(k[j]));
#pragma unroll
  for (int s = I8_S - 1; s >= 0; --s) {
    uint32_t word = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int dgt;
      if (s > 0) { dgt = static_cast<int>((q[j] + 128) & 255) - 128; q[j] = (q[j] - dgt) >> 8; }
      else dgt = static_cast<int>(q[j]);                                              // leading digit, 0 .. 65
      word |= static_cast<uint32_t>(dgt & 255) << (8 * j);
    }
    *reinterpret_cast<uint32_t*>(dst + s * plane_stride) = word;
  }
}

// Kernel values, digits and mean contribution of one thread's 8 x 4 block of squared distances.
template <int KIND>
__device__ __forceinline__ void builder_block(const SweepParams& p, double (&r2)[8][4], signed char* rows, int jo,
                                              int ty, double (&macc)[8]) {
  double al[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) al[j] = (jo + j < p.n) ? p.alpha[jo + j] : 0.0;
#pragma unroll
  for (int i = 0; i < 8; i += 2) {
    double v[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) { v[j] = r2[i][j]; v[4 + j] = r2[i + 1][j]; }
    kern_values_any<KIND, 8>(v, p.mg);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      double k[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) k[j] = (jo + j < p.n) ? v[4 * u + j] : 0.0;
      store_digits4(rows + static_cast<size_t>(ty * 8 + i + u) * p.npad + jo, p.plane_stride, k);
      macc[i + u] += (k[0] * al[0] + k[1] * al[1]) + (k[2] * al[2] + k[3] * al[3]);
    }
  }
}

// Eight kernel values of one candidate -> 7 planes x 8 packed digits (one 64-bit store per plane).
__device__ __forceinline__ void store_digits8(signed char* dst, size_t plane_stride, const double (&k)[8]) {
  long long q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) q[j] = __double2ll_rn(scale_2p54(k[j]));
#pragma unroll
  for (int s = I8_S - 1; s >= 0; --s) {
    uint32_t word[2] = {0, 0};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int dgt;
      if (s > 0) { dgt = static_cast<int>((q[j] + 128) & 255) - 128; q[j] = (q[j] - dgt) >> 8; }
      else dgt = static_cast<int>(q[j]);                                              // leading digit, 0 .. 65
      word[j >> 2] |= static_cast<uint32_t>(dgt & 255) << (8 * (j & 3));
    }
    *reinterpret_cast<uint2*>(dst + s * plane_stride) = make_uint2(word[0], word[1]);
  }
}

// Builder role with the squared distances on the FP64 tensor path (d <= 32, p.xrow set).
//
// The distance loop of build_tile is half of the builders' time and the builders are what bounds the kernel: their
// DFMA / DADD and their shared-memory loads are starved while the int8 tensor instructions execute
// (tools/i8_side_load.cu).  Here r^2 = |u|^2 + |v|^2 - 2 u.v with u = c / l - m, v = x / l - m (m = centre of the
// observations, which keeps |u|^2 + |v|^2 near r^2), and u.v comes from mma.sync.m8n8k4 (DMMA): 16x fewer
// instructions and 6x fewer shared-memory loads than the direct differences for the same pairs.
//   warp w       candidates 16 w .. 16 w + 15 (two 8-row M tiles), all 64 observations of a block (eight N tiles);
//   A fragments  the warp's candidates: 16 doubles per thread and observation block, conflict-free 8-byte loads;
//   B fragments  one 16-byte load per N tile and PAIR of K steps: thread t of a quad takes dimensions 8 P + 2 t
//                (K step 2 P) and 8 P + 2 t + 1 (K step 2 P + 1) -- both operands use this assignment, so the product
//                is unchanged;
//   N tile (h, e), column n <-> observation 32 h + 8 (n >> 1) + 2 e + (n & 1), so that thread (g, t) ends up with
//                the eight CONSECUTIVE observations 32 h + 8 t .. + 7 of its candidate (one 64-bit digit store).
// Where the expansion cancels (r^2 < (|u|^2 + |v|^2) / 8: a candidate near an observation; the expansion's error is
// ~32 ulp of |u|^2 + |v|^2, so what it keeps is good to ~3e-14 relative) the pair is recomputed from direct differences
// of the same u and v, so identical points still give exactly zero.  With few dimensions many pairs fall under that
// bound and the recomputation diverges the warp: the default takes this path from 16 dimensions up only.
template <int KIND>
__device__ __forceinline__ void builder_rows_mma(const SweepParams& p, const double (&acc)[2][2][4][2], const double (&cn)[2],
                                                 signed char* rows, int jb, uint32_t sm, int warp, int g, int t,
                                                 double (&macc)[2]) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    double xn[8], al[8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const double2 a = lds128(sm + static_cast<uint32_t>((F_XN + xn_pos(32 * h + 8 * t) + 2 * u) * 8));
      const double2 b = lds128(sm + static_cast<uint32_t>((F_AL + xn_pos(32 * h + 8 * t) + 2 * u) * 8));
      xn[2 * u] = a.x; xn[2 * u + 1] = a.y; al[2 * u] = b.x; al[2 * u + 1] = b.y;
    }
    const int jo = jb * 64 + 32 * h + 8 * t;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      const int c = 16 * warp + 8 * mi + g;
      double v[8];
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        const double s = cn[mi] + xn[o];
        double r = fma(-2.0, acc[mi][h][o >> 1][o & 1], s);
        // r < s / 8, compared on the bit patterns (integer pipe; a negative r has the sign bit set and is "less")
        if (__double_as_longlong(r) < __double_as_longlong(s) - (3LL << 52)) {      // rare: direct differences
          const uint32_t ca = sm + static_cast<uint32_t>((F_CS + c) * 8);
          const uint32_t xa = sm + static_cast<uint32_t>((F_XS + (32 * h + 8 * t + o) * F_XR_LD) * 8);
          r = 0.0;
          for (int q = 0; q < p.d; ++q) { const double df = lds64(ca + q * F_CS_LD * 8) - lds64(xa + q * 8); r = fma(df, df, r); }
        }
        v[o] = r;
      }
      kern_values_any<KIND, 8>(v, p.mg);
      double m = 0.0;
#pragma unroll
      for (int o = 0; o < 8; ++o) { v[o] = (jo + o < p.n) ? v[o] : 0.0; m = fma(v[o], al[o], m); }
      store_digits8(rows + static_cast<size_t>(c) * p.npad + jo, p.plane_stride, v);
      macc[mi] += m;
    }
  }
}

template <bool GEN>
__device__ __forceinline__ void build_tile_mma(const SweepParams& p, long long tile, signed char* rows, double* mean_out,
                                               uint32_t sm, int tid, int jb0, int jbstep) {
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const long long c_base = tile * 128;
  const int nblk = (p.n + 63) / 64;
  const int npair = (p.d + 7) / 8;
  double2 pre[5];
  double prex = 0.0;
  auto fetch = [&](int jb) {
    const double* src = p.xrow + static_cast<size_t>(jb) * 64 * F_XR_LD;
#pragma unroll
    for (int r = 0; r < 5; ++r) pre[r] = ldg128(src + 2 * (tid + 256 * r));
    if (tid < 64) prex = p.xnc[jb * 64 + tid];
    else if (tid < 128) prex = (jb * 64 + tid - 64 < p.n) ? p.alpha[jb * 64 + tid - 64] : 0.0;
  };
  if (jb0 < nblk) fetch(jb0);
  named_bar_sync(1, 256);                                     // the previous tile's readers of Cs are done
  for (int e = tid; e < 128 * 32; e += 256) {
    const int c = e >> 5, q = e & 31;
    double v = 0.0;
    if (q < p.d && c_base + c < p.N) v = p.Xc[(c_base + c) * p.d + q] / p.ell - p.xmean[q];
    sts64(sm + static_cast<uint32_t>((F_CS + q * F_CS_LD + c) * 8), v);
  }
  named_bar_sync(1, 256);
  // A fragments: re-read from shared memory for every observation block (16 conflict-free 8-byte loads) rather than
  // kept across the kernel-value stage, where their 32 registers are what would spill
  const uint32_t afa = sm + static_cast<uint32_t>((F_CS + 2 * t * F_CS_LD + 16 * warp + g) * 8);
  double cn[2], macc[2] = {0.0, 0.0};
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
    double s = 0.0;
#pragma unroll
    for (int P = 0; P < 4; ++P)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const double a = lds64(afa + ((8 * P + e) * F_CS_LD + 8 * mi) * 8);
        s = fma(a, a, s);
      }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    cn[mi] = s;
  }
  for (int jb = jb0; jb < nblk; jb += jbstep) {
    named_bar_sync(1, 256);                                   // readers of the previous observation block are done
#pragma unroll
    for (int r = 0; r < 5; ++r) sts128(sm + static_cast<uint32_t>(F_XS * 8 + (tid + 256 * r) * 16), pre[r]);
    if (tid < 128) sts64(sm + static_cast<uint32_t>(((tid < 64 ? F_XN : F_AL) + xn_pos(tid & 63)) * 8), prex);
    named_bar_sync(1, 256);
    if (jb + jbstep < nblk) fetch(jb + jbstep);
    double acc[2][2][4][2];
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mi][h][e][0] = acc[mi][h][e][1] = 0.0;
#pragma unroll
    for (int P = 0; P < 4; ++P) {
      if (P < npair) {
        double af[2][2];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
          for (int e = 0; e < 2; ++e) af[mi][e] = lds64(afa + ((8 * P + e) * F_CS_LD + 8 * mi) * 8);
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int ob = 32 * h + 8 * (g >> 1) + 2 * e + (g & 1);
            const double2 b = lds128(sm + static_cast<uint32_t>((F_XS + ob * F_XR_LD + 8 * P + 2 * t) * 8));
#pragma unroll
            for (int mi = 0; mi < 2; ++mi) {
              dmma(acc[mi][h][e][0], acc[mi][h][e][1], af[mi][0], b.x);
              dmma(acc[mi][h][e][0], acc[mi][h][e][1], af[mi][1], b.y);
            }
          }
      }
    }
    if constexpr (GEN) {
      builder_rows_mma<K_GEN>(p, acc, cn, rows, jb, sm, warp, g, t, macc);
    } else {
      switch (p.kind) {
        case K_RBF: builder_rows_mma<K_RBF>(p, acc, cn, rows, jb, sm, warp, g, t, macc); break;
        case K_M05: builder_rows_mma<K_M05>(p, acc, cn, rows, jb, sm, warp, g, t, macc); break;
        case K_M15: builder_rows_mma<K_M15>(p, acc, cn, rows, jb, sm, warp, g, t, macc); break;
        default: builder_rows_mma<K_M25>(p, acc, cn, rows, jb, sm, warp, g, t, macc); break;
      }
    }
  }
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
    double m = macc[mi];
    m += __shfl_xor_sync(0xffffffffu, m, 1);
    m += __shfl_xor_sync(0xffffffffu, m, 2);
    if (t == 0) mean_out[16 * warp + 8 * mi + g] = m;
  }
}

// Builder role for one tile: planes into `rows` (plane 0, first candidate row of the tile buffer), means into mean_out.
// Only the 64-observation blocks jb0, jb0 + jbstep, ... are built (1, 0 = all of them; the cooperative kernel deals
// the blocks to the CTAs of a group, and mean_out then receives this CTA's partial means).
// GEN: the instantiation for the general-nu Matern kernel (its own kernels below, so that the closed forms' code is
// what it was without it)
template <bool GEN>
__device__ __forceinline__ void build_tile(const SweepParams& p, long long tile, signed char* rows, double* mean_out,
                                           uint32_t sm, int tid, int jb0 = 0, int jbstep = 1) {
  if (p.xrow != nullptr) { build_tile_mma<GEN>(p, tile, rows, mean_out, sm, tid, jb0, jbstep); return; }
  const int tx = tid & 15, ty = tid >> 4;
  const long long c_base = tile * 128;
  const int nchunk_d = (p.d + 31) / 32;
  const int nblk = (p.n + 63) / 64;
  double macc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) macc[i] = 0.0;
  double pre[8];                      // element e = tid + 256 r of a 32 x 64 block: q = e >> 6, j = e & 63
  auto fetch = [&](int jb, int dc) {
    const int d0 = dc * 32, dn = min(32, p.d - d0);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int e = tid + 256 * r, q = e >> 6, j = e & 63;
      pre[r] = (q < dn) ? p.Xt[static_cast<size_t>(d0 + q) * p.npad + jb * 64 + j] : 0.0;
    }
  };
  if (jb0 < nblk) fetch(jb0, 0);
  for (int jb = jb0; jb < nblk; jb += jbstep) {
    double r2[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r2[i][j] = 0.0;
    for (int dc = 0; dc < nchunk_d; ++dc) {
      const int d0 = dc * 32, dn = min(32, p.d - d0);
      named_bar_sync(1, 256);                                 // readers of the previous shared tiles are done
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int e = tid + 256 * r;
        sts64(sm + static_cast<uint32_t>((F_XS + (e >> 6) * F_XS_LD + (e & 63)) * 8), pre[r]);
      }
      if (jb == jb0 || nchunk_d > 1) {
        for (int e = tid; e < 128 * 32; e += 256) {
          const int c = e >> 5, q = e & 31;
          double v = 0.0;
          if (q < dn && c_base + c < p.N) v = p.Xc[(c_base + c) * p.d + d0 + q] / p.ell;
          sts64(sm + static_cast<uint32_t>((F_CS + q * F_CS_LD + c) * 8), v);
        }
      }
      named_bar_sync(1, 256);
      {                                                       // prefetch the next (jb, dc)
        int njb = jb, ndc = dc + 1;
        if (ndc == nchunk_d) { ndc = 0; njb += jbstep; }
        if (njb < nblk) fetch(njb, ndc);
      }
      const uint32_t cs = sm + static_cast<uint32_t>((F_CS + ty * 8) * 8);
      const uint32_t xs = sm + static_cast<uint32_t>((F_XS + 4 * tx) * 8);
#pragma unroll 2
      for (int q = 0; q < dn; ++q) {
        double cv[8], xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { const double2 t2 = lds128(cs + (q * F_CS_LD + 2 * i) * 8); cv[2 * i] = t2.x; cv[2 * i + 1] = t2.y; }
        { const double2 a = lds128(xs + q * F_XS_LD * 8), b = lds128(xs + q * F_XS_LD * 8 + 16); xv[0] = a.x; xv[1] = a.y; xv[2] = b.x; xv[3] = b.y; }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) { const double df = cv[i] - xv[j]; r2[i][j] += df * df; }
      }
    }
    const int jo = jb * 64 + 4 * tx;
    if constexpr (GEN) {
      builder_block<K_GEN>(p, r2, rows, jo, ty, macc);
    } else {
      switch (p.kind) {
        case K_RBF: builder_block<K_RBF>(p, r2, rows, jo, ty, macc); break;
        case K_M05: builder_block<K_M05>(p, r2, rows, jo, ty, macc); break;
        case K_M15: builder_block<K_M15>(p, r2, rows, jo, ty, macc); break;
        default: builder_block<K_M25>(p, r2, rows, jo, ty, macc); break;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double m = macc[i];
    m += __shfl_xor_sync(0xffffffffu, m, 1);
    m += __shfl_xor_sync(0xffffffffu, m, 2);
    m += __shfl_xor_sync(0xffffffffu, m, 4);
    m += __shfl_xor_sync(0xffffffffu, m, 8);
    if (tx == 0) mean_out[ty * 8 + i] = m;
  }
}

// p: the builders' view (candidates, observations, planes = this launch's plane buffer of gridDim.x * 2 tiles,
// plane_stride); ip: the contraction's view (ip.mean = mean buffer of gridDim.x * 256 doubles).
template <bool GEN>
__device__ __forceinline__ void sweep_i8f_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const SweepParams& p,
                                               const I8Params& ip, double* meanbuf) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  I8FTail* tail = reinterpret_cast<I8FTail*>(smem + I8_STAGES * I8_STAGE + F_STAGING);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = smem_u32(tail->bars);
  const uint32_t bar_accf = bar0 + 8 * (2 * I8_STAGES), bar_acce = bar_accf + 8;
  const uint32_t bar_ready = bar_acce + 8, bar_free = bar_ready + 16;      // two each, one per plane buffer
  if (tid == 0) {
    for (int s = 0; s < I8_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (I8_STAGES + s), 1); }
    mbar_init(bar_accf, 1);
    mbar_init(bar_acce, 128);
    for (int b = 0; b < 2; ++b) { mbar_init(bar_ready + 8 * b, 8); mbar_init(bar_free + 8 * b, 128); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tail->tmem_base)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tail->tmem_base;

  const long long ntiles = (ip.N + 127) / 128;
  const int nh = (ip.n + 63) / 64;
  const int buf_row0 = blockIdx.x * 256;               // this CTA's two tile buffers: rows buf_row0 + 128 b

  if (warp < 8) {
    // ===== builders =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
    const uint32_t sm = s0 + I8_STAGES * I8_STAGE;
    uint32_t t = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
      if (use > 0) mbar_wait(bar_free + 8 * b, (use - 1) & 1);       // the tile that used this buffer is consumed
      signed char* rows = p.planes + static_cast<size_t>(buf_row0 + 128 * b) * p.npad;
      if (!PLX_DBG(p.dbg, 128) || t < 2) build_tile<GEN>(p, tile, rows, meanbuf + buf_row0 + 128 * b, sm, tid);
      __threadfence();
      asm volatile("fence.proxy.async;\n" ::: "memory");            // generic-proxy stores -> visible to TMA
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ready + 8 * b);
    }
  } else if (warp < 12) {
    // ===== accumulator drain + epilogue: thread = candidate row of the tile = TMEM lane =====
    asm volatile("setmaxnreg.dec.sync.aligned.u32 120;\n");
    const int row = tid - 256;
    Best best = {-INFINITY, 0x7fffffffffffffffLL};
    uint32_t acc_round = 0, t = 0;
    const uint32_t lane_base = static_cast<uint32_t>((warp - 8) * 32) << 16;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
      double sumsq = 0.0;
      for (int h = 0; h < nh; ++h, ++acc_round) {
        mbar_wait(bar_accf, acc_round & 1);
        asm volatile("tcgen05.fence::after_thread_sync;\n");
        drain_halfblock(tmem + lane_base, ip.zscale + h * 64, sumsq, bar_acce);
      }
      // all tensor instructions of the tile have completed, hence all loads of its planes: the means were
      // written before the builders' arrival on `ready`, which the producer (and so this point) is ordered after
      mbar_wait(bar_ready + 8 * b, use & 1);
      const double mean = meanbuf[buf_row0 + 128 * b + row];
      mbar_arrive(bar_free + 8 * b);
      const long long ci = tile * 128 + row;
      if (ci < ip.N) {
        double var = 1.0 - sumsq;
        if (var < 0.0) { var = 0.0; if (ip.clamped) atomicAdd(ip.clamped, 1ULL); }   // sklearn clamps negative variances to zero, and says so
        const double sd = sqrt(var);
        const double a = acq_value(ip.acq, mean, sd, ip.y_max, ip.kappa, ip.eps);
        if (ip.out_mean) ip.out_mean[ci] = mean;
        if (ip.out_std) ip.out_std[ci] = sd;
        if (ip.out_acq) ip.out_acq[ci] = a;
        const long long gi = ip.idx_base + ci;
        if (better(a, gi, best.v, best.i)) { best.v = a; best.i = gi; }
      }
    }
    tail->sval[row] = best.v; tail->sidx[row] = best.i;
    named_bar_sync(2, 128);
    if (row == 0) {
      Best bb = {tail->sval[0], tail->sidx[0]};
      for (int q = 1; q < 128; ++q)
        if (better(tail->sval[q], tail->sidx[q], bb.v, bb.i)) { bb.v = tail->sval[q]; bb.i = tail->sidx[q]; }
      ip.blk_val[blockIdx.x] = bb.v;
      ip.blk_idx[blockIdx.x] = bb.i;
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 48;\n");
    if (warp == 12 && lane == 0) {
      // ===== producer =====
      uint32_t it = 0, t = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
        const uint32_t b = t & 1, use = t >> 1;
        mbar_wait(bar_ready + 8 * b, use & 1);              // the tile's planes are complete and visible to TMA
        for (int h = 0; h < nh; ++h)
          for (int kc = 0; kc < (h + 1) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * (I8_STAGES + s), ph ^ 1);
            const uint32_t full = bar0 + 8 * s;
            mbar_expect_tx(full, I8_STAGE);
            tma_load_3d(s0 + s * I8_STAGE, &tmA, kc * I8_KC, buf_row0 + 128 * static_cast<int>(b), 0, full);
            tma_load_3d(s0 + s * I8_STAGE + I8_A_BYTES, &tmB, kc * I8_KC, h * 64, 0, full);
          }
      }
    } else if (warp == 13) {
      // ===== tensor-instruction issuer: the whole warp walks the loops, one elected lane issues =====
      uint32_t it = 0, acc_round = 0, ring = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
        for (int h = 0; h < nh; ++h, ++acc_round) {
          mbar_wait(bar_acce, (acc_round & 1) ^ 1);        // the previous half-block has been drained
          asm volatile("tcgen05.fence::after_thread_sync;\n");
          for (int kc = 0; kc < (h + 1) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;\n");
            const uint32_t a0 = s0 + s * I8_STAGE;
            const uint64_t da0 = umma_desc_k(a0), db0 = umma_desc_k(a0 + I8_A_BYTES);
            const uint32_t fresh = (kc != 0);              // the first chunk of a half-block overwrites the accumulators
            if (elect_one()) {
              if (!PLX_DBG(p.dbg, 64)) issue_stage(tmem, da0, db0, fresh, ring, ip.a_from_smem != 0);
              umma_commit(bar0 + 8 * (I8_STAGES + s));       // stage free once these instructions have read it
            }
            ring += I8_KSTEPS * I8_S;
            __syncwarp();
          }
          if (elect_one()) umma_commit(bar_accf);          // accumulators of this half-block complete
          __syncwarp();
        }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(512));
}

__global__ void __launch_bounds__(512, 1) sweep_i8f_kernel(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const SweepParams p, const I8Params ip, double* meanbuf) {
  sweep_i8f_body<false>(tmA, tmB, p, ip, meanbuf);
}
// the general-nu Matern kernel (kernel values from the table of gp_kernels.cuh)
__global__ void __launch_bounds__(512, 1) sweep_i8f_gen_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const SweepParams p, const I8Params ip, double* meanbuf) {
  sweep_i8f_body<true>(tmA, tmB, p, ip, meanbuf);
}

}  // namespace plx
