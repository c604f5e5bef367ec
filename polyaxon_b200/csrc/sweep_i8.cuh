// The int8-sliced acquisition sweep: the variance contraction V = K* . Z^T on the int8 tensor cores by exact
// slicing, instead of FP64 DMMA.  This header holds the arithmetic shared by its three forms -- digit planes,
// tensor-instruction and descriptor helpers, the accumulator drain -- and the two-kernel form ("sweep_impl" = 2);
// sweep_i8f.cuh fuses the two kernels into one persistent CTA per SM (4), sweep_i8c.cuh lets four CTAs share a
// candidate tile (5; the default from 1536 observations, the fused form from 256).
//
//   K*[c][j] = 4 * sum_s a_s[c][j] 256^-(s+1)                   a_s balanced 8-bit digits (K* in [0, 1], fixed exponent)
//   Z[i][j]  = 2^f_i * sum_t b_t[i][j] 256^-(t+1)               f_i from the largest entry of row i of L^-1
//   V[c][i]  = 2^(f_i+2) * sum_{s+t<S} 256^-(s+t+2) (a_s[c] . b_t[i])
//
// Every dot product a_s . b_t is an exact int32 (tcgen05.mma kind::i8, accumulators in TMEM); the S scale
// classes s + t = 0 .. S-1 are kept in S separate accumulators and recombined in FP64, smallest first.  With
// S = 7 digits of 8 bits (56 bits per operand, 28 digit-plane pairs) the result is as accurate as an FP64 FMA
// chain (tools/ozaki_probe.cu measured 3.5e-16 relative to sum |K* Z| at n = 2048 for 8 x 7 bits, the same
// 56 bits and the same truncation order 2^-47).  Reference arithmetic being reproduced: sklearn _gpr.py:460-481.
//
// Two-kernel form, per chunk of candidates:
//   kstar_slices_kernel  phase A of the DMMA variant (same distances, same kernel values, same posterior mean) writing
//                        the S digit planes of K* (int8) instead of the FP64 tile;
//   sweep_i8_kernel      persistent, 192 threads: warp 4 streams the digit planes of a 128-candidate tile and of a
//                        64-row half-block of Z through a two-stage ring (one 3-D bulk tensor load per operand and
//                        64-observation chunk, 64-byte swizzle), warp 5 issues the 56 tensor instructions per stage
//                        (28 digit-plane pairs x 2 k-steps) into 7 x 64 TMEM columns, warps 0-3 drain the accumulators of a
//                        finished half-block (tcgen05.ld), recombine in FP64, scale, square and sum per candidate, and
//                        finish with variance -> acquisition -> arg-max exactly as the DMMA variant.
#pragma once
#include "sweep.cuh"

namespace plx {

#ifndef PLX_I8_KC
#define PLX_I8_KC 64
#endif
constexpr int I8_KC = PLX_I8_KC;                      // observations (= bytes) per K chunk: 32 or 64
constexpr int I8_KSTEPS = I8_KC / 32;                 // k32 tensor instructions per chunk and digit pair
constexpr int I8_CPH = 64 / I8_KC;                    // chunks per 64 observations
constexpr int I8_A_SLICE = 128 * I8_KC;               // one plane of the candidate tile chunk
constexpr int I8_B_SLICE = 64 * I8_KC;                // one plane of the Z half-block chunk
constexpr int I8_A_BYTES = I8_S * I8_A_SLICE;         // 64 KiB
constexpr int I8_B_BYTES = I8_S * I8_B_SLICE;         // 32 KiB
constexpr int I8_STAGE = I8_A_BYTES + I8_B_BYTES;
constexpr int I8_STAGES = 128 / I8_KC;                // 2 x 84 KB or 4 x 42 KB in flight
constexpr int I8_SMEM = I8_STAGES * I8_STAGE + 1024 /*align*/ + 3072 /*tail*/;

struct I8Params {
  long long N;                 // candidates of this launch (one chunk)
  long long idx_base;          // global index of candidate 0 of this launch
  int n;                       // observations
  const double* zscale;        // [npad]: 2^(f_i + 2), zero for rows >= n
  const double* mean;          // [N] posterior means written by kstar_slices_kernel
  int acq;
  double kappa, eps, y_max;
  double* out_mean; double* out_std; double* out_acq;       // optional dumps, indexed like the candidates of this launch
  double* blk_val; long long* blk_idx;
  unsigned long long* clamped; // nullable: counts negative variances set to zero
  const signed char* zplanes_dbg;   // PLX_PROFILE experiments only
  int a_from_smem;             // 1: both tensor-core operands from shared memory (A/B measurement), 0: A planes via tensor memory
};

// K-major operand rows of I8_KC bytes, swizzle of the same width, 8-row groups 8 * I8_KC bytes apart
// (layout type 4 = 64-byte, 6 = 32-byte swizzle)
__device__ __forceinline__ uint64_t umma_desc_k(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (static_cast<uint64_t>(8 * I8_KC / 16) << 32) |
         (1ull << 46) | (static_cast<uint64_t>(I8_KC == 64 ? 4 : 6) << 61);
}
// kind::i8: D = s32, A = B = signed 8 bit, both K-major, M = 128, N = 64
constexpr uint32_t I8_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      :: "r"(tmem_c), "l"(da), "l"(db), "r"(I8_IDESC), "r"(accumulate), "r"(0u) : "memory");
}
// One lane of a converged warp (elect.sync): tensor instructions take their operands from uniform registers, and
// issuing them from a branch the compiler knows to be single-lane avoids a per-instruction lane loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}\n" : "=r"(pred));
  return pred != 0;
}
// The same instruction with the A operand (128 candidate rows x 32 bytes of one digit plane) in tensor memory:
// lane = candidate row, 8 columns = the 32 bytes of the k step.
__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      :: "r"(tmem_c), "r"(tmem_a), "l"(db), "r"(I8_IDESC), "r"(accumulate), "r"(0u) : "memory");
}
// shared memory -> tensor memory: 128 rows x 256 bits described by a K-major operand descriptor
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t tmem_dst, uint64_t desc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;\n" :: "r"(tmem_dst), "l"(desc) : "memory");
}

// Tensor instructions of one stage (I8_KSTEPS k32 steps of the 28 digit-plane pairs), issued by ONE lane.
//
// from_smem = false (default): every A plane of a k step is copied once into tensor memory (tcgen05.cp) and the
// 7 - s instructions that use plane s read it from there, so the tensor core fetches only the B planes from shared
// memory: 28 KB (copies) + 56 KB (B) per k step instead of 168 KB.  With both operands in shared memory the
// m128 n64 k32 instruction is bound by that fetch (6 KB per 262 k MAC: 2982 TOPS measured against 4074 for n128);
// with A in tensor memory the same stream runs at 4032 TOPS (tools/i8_umma_ts.cu, profiles/r02_i8_umma_ts_probe.log).
// The planes go through a ring of I8_RING slots of 8 columns in the 64 columns the accumulators leave free; the copy
// of plane p + 1 is issued before the instructions of plane p.  Copies and tensor instructions of one thread execute
// in issue order, so a slot is overwritten only after the instructions that read it (no barrier needed).
// `ring` is the running plane count of the caller (advance it by I8_KSTEPS * I8_S per stage on every lane).
// `fresh` = 0 on the first chunk of a half-block: the first instruction into each class overwrites the accumulator.
constexpr int I8_ACC_COLS = I8_S * 64;
constexpr int I8_RING = (512 - I8_ACC_COLS) / 8;
static_assert(I8_RING >= 2, "the A-plane ring needs two slots");
__device__ __forceinline__ void issue_stage(uint32_t tmem, uint64_t da0, uint64_t db0, uint32_t fresh, uint32_t ring,
                                            bool from_smem) {
  if (from_smem) {
#pragma unroll
    for (int t = 0; t < I8_S; ++t) {
#pragma unroll
      for (int sl = 0; sl + t < I8_S; ++sl) {
        // descriptors differ only in the 16-byte-unit address field: + slice * size / 16, + 2 for the second k-step
        const uint64_t da = da0 + static_cast<uint64_t>(sl * (I8_A_SLICE / 16));
        const uint64_t db = db0 + static_cast<uint64_t>(t * (I8_B_SLICE / 16));
        umma_i8(tmem + (sl + t) * 64, da, db, t == 0 ? fresh : 1u);          // class sl + t is first touched by (sl, 0)
        if (I8_KSTEPS == 2) umma_i8(tmem + (sl + t) * 64, da + 2, db + 2, 1u);
      }
    }
    return;
  }
  constexpr int P = I8_KSTEPS * I8_S;
  const uint32_t ring0 = tmem + I8_ACC_COLS;
  tmem_cp_128x256b(ring0 + 8 * (ring % I8_RING), da0);
#pragma unroll
  for (int pl = 0; pl < P; ++pl) {
    const int k = pl / I8_S, sl = pl % I8_S;
    if (pl + 1 < P) {
      const int k1 = (pl + 1) / I8_S, s1 = (pl + 1) % I8_S;
      tmem_cp_128x256b(ring0 + 8 * ((ring + pl + 1) % I8_RING), da0 + static_cast<uint64_t>(s1 * (I8_A_SLICE / 16) + 2 * k1));
    }
    const uint32_t ta = ring0 + 8 * ((ring + pl) % I8_RING);
#pragma unroll
    for (int t = 0; sl + t < I8_S; ++t)                                       // class t is first touched by (0, t) of k step 0
      umma_i8_ts(tmem + (sl + t) * 64, ta, db0 + static_cast<uint64_t>(t * (I8_B_SLICE / 16) + 2 * k),
                 (sl == 0 && k == 0) ? fresh : 1u);
  }
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" :: "r"(bar) : "memory");
}

// tcgen05.ld of 16 consecutive 32-bit columns of this thread's TMEM lane (no wait)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&u)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
        "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
      : "r"(taddr));
}

// Drain of one finished half-block for one candidate (= TMEM lane): the I8_S class accumulators of 64 columns are
// recombined, scaled by zs[.] and their squares added to sumsq.  The recombination is done in INTEGER arithmetic:
// classes 0..3 and classes 4..6 are each folded into one 64-bit integer (|sum of a class| < 2^28, so 2^52 and 2^44:
// exact), and only those two are converted to FP64 and added (one rounding, against seven in an FMA chain over the
// classes).  This matters for speed, not only accuracy: the FP64 pipe is shared with the eight builder warps, and the
// accumulators can only be handed back to the tensor pipe (bar_acce) when the last tcgen05.ld has landed, which the
// FP64 work of the earlier columns used to delay.  The loads of several classes are in flight together (one wait per
// batch: the round trip of tcgen05.ld, not its bandwidth, is what the tensor pipe waits for while TMEM cannot be
// double-buffered).
__device__ __forceinline__ void drain_halfblock(uint32_t taddr, const double* __restrict__ zs, double& sumsq,
                                                uint32_t bar_acce) {
  static_assert(I8_S == 7, "the integer recombination below is written for 7 digit planes");
  const double w_hi = __longlong_as_double(static_cast<long long>(1023 - 40) << 52);   // 256^-(3+2): class 3 is the unit of `hi`
  const double w_lo = __longlong_as_double(static_cast<long long>(1023 - 64) << 52);   // 256^-(6+2): class 6 is the unit of `lo`
#pragma unroll 1
  for (int c0 = 0; c0 < 64; c0 += 16) {
    long long hi[16], lo[16];
    {
      uint32_t u[4][16];
#pragma unroll
      for (int k = 0; k < 4; ++k) tmem_ld16(taddr + k * 64 + c0, u[k]);
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
      for (int i = 0; i < 16; ++i)
        hi[i] = ((static_cast<long long>(static_cast<int>(u[0][i])) * 256 + static_cast<int>(u[1][i])) * 256 +
                 static_cast<int>(u[2][i])) * 256 + static_cast<int>(u[3][i]);
    }
    {
      uint32_t u[3][16];
#pragma unroll
      for (int k = 0; k < 3; ++k) tmem_ld16(taddr + (4 + k) * 64 + c0, u[k]);
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      if (c0 == 48) {                                     // every column of every class has been read
        asm volatile("tcgen05.fence::before_thread_sync;\n");
        mbar_arrive(bar_acce);
      }
#pragma unroll
      for (int i = 0; i < 16; ++i)
        lo[i] = (static_cast<long long>(static_cast<int>(u[0][i])) * 256 + static_cast<int>(u[1][i])) * 256 +
                static_cast<int>(u[2][i]);
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const double v = fma(static_cast<double>(hi[i]), w_hi, static_cast<double>(lo[i]) * w_lo);
      const double x = v * zs[c0 + i];
      sumsq = fma(x, x, sumsq);
    }
  }
}

struct I8Tail {
  unsigned long long bars[2 * I8_STAGES + 2];       // full[ST], empty[ST], accumulators full, accumulators empty
  uint32_t tmem_base;
  double sval[128];
  long long sidx[128];
};
static_assert(sizeof(I8Tail) <= 3072, "tail budget");

__global__ void __launch_bounds__(192, 1) sweep_i8_kernel(const __grid_constant__ CUtensorMap tmA,
                                                          const __grid_constant__ CUtensorMap tmB, const I8Params p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  I8Tail* tail = reinterpret_cast<I8Tail*>(smem + I8_STAGES * I8_STAGE);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = smem_u32(tail->bars);
  const uint32_t bar_accf = bar0 + 8 * (2 * I8_STAGES), bar_acce = bar_accf + 8;
  if (tid == 0) {
    for (int s = 0; s < I8_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (I8_STAGES + s), 1); }
    mbar_init(bar_accf, 1);
    mbar_init(bar_acce, 128);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tail->tmem_base)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tail->tmem_base;

  const long long ntiles = (p.N + 127) / 128;
  const int nh = (p.n + 63) / 64;                    // half-blocks of 64 rows of Z that hold observations

  if (warp == 4) {
    // ===== producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
        for (int h = 0; h < nh; ++h)
          for (int kc = 0; kc < (h + 1) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * (I8_STAGES + s), ph ^ 1);
            const uint32_t full = bar0 + 8 * s;
            mbar_expect_tx(full, I8_STAGE);
            tma_load_3d(s0 + s * I8_STAGE, &tmA, kc * I8_KC, static_cast<int>(tile * 128), 0, full);
            tma_load_3d(s0 + s * I8_STAGE + I8_A_BYTES, &tmB, kc * I8_KC, h * 64, 0, full);
          }
    }
  } else if (warp == 5) {
    // ===== tensor-instruction issuer: the whole warp walks the loops, one elected lane issues =====
    {
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
            const uint32_t fresh = (kc != 0);                // the first chunk of a half-block overwrites the accumulators
            if (elect_one()) {
              issue_stage(tmem, da0, db0, fresh, ring, p.a_from_smem != 0);
              umma_commit(bar0 + 8 * (I8_STAGES + s));       // stage free once these instructions have read it
            }
            ring += I8_KSTEPS * I8_S;
            __syncwarp();
          }
          if (elect_one()) umma_commit(bar_accf);          // accumulators of this half-block complete
          __syncwarp();
        }
    }
  } else if (warp < 4) {
    // ===== accumulator drain + epilogue: thread = candidate row of the tile = TMEM lane =====
    Best best = {-INFINITY, 0x7fffffffffffffffLL};
    uint32_t acc_round = 0;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      double sumsq = 0.0;
      for (int h = 0; h < nh; ++h, ++acc_round) {
        mbar_wait(bar_accf, acc_round & 1);
        asm volatile("tcgen05.fence::after_thread_sync;\n");
        drain_halfblock(tmem + lane_base, p.zscale + h * 64, sumsq, bar_acce);
      }
      const long long ci = tile * 128 + tid;
      if (ci < p.N) {
        double var = 1.0 - sumsq;
        if (var < 0.0) { var = 0.0; if (p.clamped) atomicAdd(p.clamped, 1ULL); }   // sklearn clamps negative variances to zero, and says so
        const double sd = sqrt(var), mean = p.mean[ci];
        const double a = acq_value(p.acq, mean, sd, p.y_max, p.kappa, p.eps);
        if (p.out_mean) p.out_mean[ci] = mean;
        if (p.out_std) p.out_std[ci] = sd;
        if (p.out_acq) p.out_acq[ci] = a;
        const long long gi = p.idx_base + ci;
        if (better(a, gi, best.v, best.i)) { best.v = a; best.i = gi; }
      }
    }
    tail->sval[tid] = best.v; tail->sidx[tid] = best.i;
    named_bar_sync(1, 128);
    if (tid == 0) {
      Best b = {tail->sval[0], tail->sidx[0]};
      for (int t = 1; t < 128; ++t)
        if (better(tail->sval[t], tail->sidx[t], b.v, b.i)) { b.v = tail->sval[t]; b.i = tail->sidx[t]; }
      p.blk_val[blockIdx.x] = b.v;
      p.blk_idx[blockIdx.x] = b.i;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(512));
}

// Phase A for a chunk of candidates: digit planes of K* (rows = candidates of the chunk) and posterior means.
constexpr int KSL_SMEM = PA_DOUBLES * 8 + 1024 + 1024;

__global__ void __launch_bounds__(256, 1) kstar_slices_kernel(const SweepParams p, double* __restrict__ mean_out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  double* mean_s = reinterpret_cast<double*>(smem + PA_DOUBLES * 8);
  const uint32_t s0 = smem_u32(smem);
  const int tid = threadIdx.x;
  const long long ntiles = (p.N + 127) / 128;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    signed char* rows = p.planes + static_cast<size_t>(tile) * 128 * p.npad;
    phase_a3<true>(p, tile, reinterpret_cast<double*>(rows), s0, mean_s, tid);
    if (tid < 128 && tile * 128 + tid < p.N) mean_out[tile * 128 + tid] = mean_s[tid];
    __syncthreads();                      // the staging area is reused by the next tile
  }
}

// Digit planes and row scales of Z = L^-1 (lower triangular, npad x npad).  One CTA per row.
// planes[s][i][j]; zscale[i] = 2^(f_i + 2) for i < n, else 0 (the extra factor 4 is K*'s fixed exponent).
__global__ void __launch_bounds__(128) zslice_kernel(const double* __restrict__ Z, int npad, int n,
                                                     int8_t* __restrict__ planes, double* __restrict__ zscale) {
  __shared__ double red[128];
  const int i = blockIdx.x, tid = threadIdx.x;
  const double* row = Z + static_cast<size_t>(i) * npad;
  double m = 0.0;
  if (i < n)
    for (int j = tid; j <= i; j += 128) m = fmax(m, fabs(row[j]));
  red[tid] = m;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) { if (tid < o) red[tid] = fmax(red[tid], red[tid + o]); __syncthreads(); }
  m = red[0];
  int e = 0;
  if (m > 0.0) { frexp(m, &e); e += 2; }          // m * 2^-e in [1/8, 1/4): the leading digit stays below 65
  if (tid == 0) zscale[i] = (i < n && m > 0.0) ? ldexp(1.0, e + 2) : 0.0;
  const size_t pstride = static_cast<size_t>(npad) * npad;
  for (int j = tid; j < npad; j += 128) {
    const double x = (i < n && j <= i) ? ldexp(row[j], 56 - e) : 0.0;          // |x| < 2^54
    store_digits_q(reinterpret_cast<signed char*>(planes) + static_cast<size_t>(i) * npad + j, pstride, __double2ll_rn(x));
  }
}

// Issue-rate probe for the roofline denominator of the int8 path: every SM repeats the four k32 steps of one
// m128 n128 tile product on operands resident in shared memory (contents irrelevant), accumulating in TMEM.
__global__ void __launch_bounds__(128, 1) i8_peak_kernel(int repeats) {
  __shared__ __align__(1024) uint8_t sA[128 * 128];
  __shared__ __align__(1024) uint8_t sB[128 * 128];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 128 * 128 / 4; i += 128) {
    reinterpret_cast<uint32_t*>(sA)[i] = 0x01010101u * (i & 3);
    reinterpret_cast<uint32_t*>(sB)[i] = 0x01010101u;
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tmem_base_s)), "n"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    // 128-byte rows, 128-byte swizzle, 8-row groups 1024 bytes apart; N = 128
    const uint64_t hi = (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
    const uint64_t da = hi | ((smem_u32(sA) & 0x3FFFF) >> 4), db = hi | ((smem_u32(sB) & 0x3FFFF) >> 4);
    constexpr uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
    for (int rep = 0; rep < repeats; ++rep)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
            :: "r"(tmem), "l"(da + 2 * k), "l"(db + 2 * k), "r"(idesc), "r"(static_cast<uint32_t>((rep | k) != 0)), "r"(0u) : "memory");
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(128));
}

}  // namespace plx
