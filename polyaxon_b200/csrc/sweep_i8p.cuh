// Paired form of the cooperative int8-sliced sweep (sweep_i8c.cuh): the four CTAs of a group are two thread-block
// clusters of two, and the two CTAs of a cluster walk the SAME k chunks in lock step -- CTA 0 of the pair contracts
// half-block 2q of L^-1, CTA 1 half-block 2q + 1 -- so every 56 KB chunk of the candidate tile's digit planes is
// fetched from L2 once per pair instead of once per CTA: each CTA loads half of the planes (even / odd) with a bulk
// tensor load that is multicast to both CTAs of the cluster, and its own 28 KB chunk of Z planes.
//
// Why it was built: with the A planes going to the tensor core through tensor memory the contraction is no longer
// bound by the shared-memory operand reads, and the operand stream (84 KB per 56 tensor instructions and CTA, ~9 TB/s
// from L2 over the whole chip) looked like the next limit; sharing A in a pair cuts what L2 has to read by a third.
// What was measured: the same results at the same speed (8.28 against 8.22 ms per 151 552 candidates; the stream alone
// 3.86 against 3.94 ms).  The kernel is bound by its builders, which the tensor instructions starve of the FP64 pipe
// and the shared-memory port, not by the stream.  Kept as `sweep_impl` 6, not the default.
//
// Lock step needs equal chunk counts: half-block 2q needs chunks 0 .. 2q, half-block 2q + 1 chunks 0 .. 2q + 1.  CTA 0
// also takes chunk 2q + 1; there its rows of Z are above the diagonal, i.e. zero planes (zslice_kernel writes zeros for
// j > i), so the products are exact zeros: 1 extra chunk in 2q + 2 (3 % more tensor work at n = 2048).  Double
// half-blocks q are dealt to the two pairs 0 1 1 0 0 1 1 0 ... so that the triangular work balances.
//
// Stage protocol inside a pair: the `full` barrier of a stage is armed by the CTA's own producer with the bytes of the
// whole stage (its own loads plus the peer's multicast halves land on the same barrier offset in each CTA); the
// `empty` barrier of a stage counts TWO arrivals, the tcgen05.commit of the CTA's own tensor instructions and of the
// peer's (one commit, multicast to both CTAs), because a producer overwrites the stage in both CTAs.  The tensor maps
// cover this GPU's plane buffers only (TMA multicast loads here read only this GPU's memory).
// Everything else -- builders, accumulator drain, the group's hand-over through global counters, arithmetic -- is
// sweep_i8c.cuh's.
#pragma once
#include "sweep_i8c.cuh"

namespace plx {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d_mc(uint32_t dst, const CUtensorMap* map, int x, int y, int z, uint32_t bar,
                                               uint16_t cta_mask) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%2, %3, %4}], [%5], %6;\n"
      :: "r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar), "h"(cta_mask) : "memory");
}
// completion of this thread's earlier tensor instructions -> one arrival on the barrier at this offset in every CTA of the mask
__device__ __forceinline__ void umma_commit_mc(uint32_t bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
               :: "r"(bar), "h"(cta_mask) : "memory");
}
// which pair of the group contracts double half-block q
__device__ __forceinline__ int pair_of(int q) { return (((q & 3) + 1) >> 1) & 1; }

// tmA1: the plane buffer with a box of ONE plane (64 bytes x 128 rows); tmB as in sweep_i8c_kernel.
// Launched with cluster dimension 2; gridDim.x is a multiple of 4.
__global__ void __launch_bounds__(512, 1) sweep_i8p_kernel(const __grid_constant__ CUtensorMap tmA1,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const SweepParams p, const I8Params ip, double* gpart,
                                                           unsigned* gsync) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  I8FTail* tail = reinterpret_cast<I8FTail*>(smem + I8_STAGES * I8_STAGE + F_STAGING);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = smem_u32(tail->bars);
  const uint32_t bar_accf = bar0 + 8 * (2 * I8_STAGES), bar_acce = bar_accf + 8;
  const uint32_t bar_ready = bar_acce + 8;                 // two: this CTA's share of a plane buffer is built
  if (tid == 0) {
    for (int s = 0; s < I8_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (I8_STAGES + s), 2); }
    mbar_init(bar_accf, 1);
    mbar_init(bar_acce, 128);
    for (int b = 0; b < 2; ++b) mbar_init(bar_ready + 8 * b, 8);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tail->tmem_base)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  cluster_sync_all();                                      // the peer's barriers exist before anything is sent to them
  const uint32_t tmem = tail->tmem_base;

  const int groups = gridDim.x / I8C_R, g = blockIdx.x / I8C_R, r = blockIdx.x % I8C_R;
  const int pr = r >> 1, cr = static_cast<int>(cluster_ctarank());           // pair of the group, rank in the pair
  const long long ntiles = (ip.N + 127) / 128;
  const int nh = (ip.n + 63) / 64, nq = (nh + 1) / 2;
  const int buf_row0 = g * 256;
  unsigned* sync_g = gsync + static_cast<size_t>(g) * 8;                     // [b][4]
  double* part_g = gpart + static_cast<size_t>(g) * 2 * I8C_R * 256;         // [b][r][256]

  if (warp < 8) {
    // ===== builders: every I8C_R-th 64-observation block of the tile's planes =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
    const uint32_t sm = s0 + I8_STAGES * I8_STAGE;
    uint32_t t = 0;
    for (long long tile = g; tile < ntiles; tile += groups, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
      if (use > 0) {                                   // every CTA of the group has consumed the tile that used this buffer
        if (tid == 0) spin_until(sync_g + 4 * b + 2, I8C_R * use);
        named_bar_sync(1, 256);
      }
      signed char* rows = p.planes + static_cast<size_t>(buf_row0 + 128 * b) * p.npad;
      if (!PLX_DBG(p.dbg, 128) || t < 2) build_tile<false>(p, tile, rows, part_g + (b * I8C_R + r) * 256 + 128, sm, tid, r, I8C_R);
      __threadfence();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ready + 8 * b);
    }
  } else if (warp < 12) {
    // ===== accumulator drain; CTA 0 of the group also finishes the tile =====
    asm volatile("setmaxnreg.dec.sync.aligned.u32 120;\n");
    const int row = tid - 256;
    Best best = {-INFINITY, 0x7fffffffffffffffLL};
    uint32_t acc_round = 0, t = 0;
    const uint32_t lane_base = static_cast<uint32_t>((warp - 8) * 32) << 16;
    for (long long tile = g; tile < ntiles; tile += groups, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
      mbar_wait(bar_ready + 8 * b, use & 1);           // ties a CTA without half-blocks to the tile (see sweep_i8c.cuh)
      double sumsq = 0.0;
      for (int q = 0; q < nq; ++q) {
        if (pair_of(q) != pr) continue;
        mbar_wait(bar_accf, acc_round & 1);
        ++acc_round;
        asm volatile("tcgen05.fence::after_thread_sync;\n");
        drain_halfblock(tmem + lane_base, ip.zscale + (2 * q + cr) * 64, sumsq, bar_acce);   // zscale is 0 beyond n
      }
      double* mine = part_g + (b * I8C_R + r) * 256;
      mine[row] = sumsq;
      __threadfence();
      named_bar_sync(2, 128);
      if (row == 0) red_release_gpu(sync_g + 4 * b + 1, 1u);
      if (r != 0) {
        if (row == 0) red_release_gpu(sync_g + 4 * b + 2, 1u);        // nothing more to read from this buffer
        continue;
      }
      if (row == 0) spin_until(sync_g + 4 * b + 1, I8C_R * (use + 1));
      named_bar_sync(2, 128);
      double ss = 0.0, mean = 0.0;
#pragma unroll
      for (int k = 0; k < I8C_R; ++k) {
        ss += __ldcg(part_g + (b * I8C_R + k) * 256 + row);
        mean += __ldcg(part_g + (b * I8C_R + k) * 256 + 128 + row);
      }
      named_bar_sync(2, 128);                                          // everyone has read the partials
      if (row == 0) red_release_gpu(sync_g + 4 * b + 2, 1u);
      const long long ci = tile * 128 + row;
      if (ci < ip.N) {
        double var = 1.0 - ss;
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
      for (int k = 1; k < 128; ++k)
        if (better(tail->sval[k], tail->sidx[k], bb.v, bb.i)) { bb.v = tail->sval[k]; bb.i = tail->sidx[k]; }
      ip.blk_val[blockIdx.x] = bb.v;
      ip.blk_idx[blockIdx.x] = bb.i;
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 48;\n");
    if (warp == 12 && lane == 0) {
      // ===== producer: its half of the A planes to both CTAs of the pair, its own Z chunk =====
      uint32_t it = 0, t = 0;
      for (long long tile = g; tile < ntiles; tile += groups, ++t) {
        const uint32_t b = t & 1, use = t >> 1;
        mbar_wait(bar_ready + 8 * b, use & 1);              // this CTA's blocks of the planes are written ...
        red_release_gpu(sync_g + 4 * b, 1u);
        spin_until(sync_g + 4 * b, I8C_R * (use + 1));      // ... and so are the other CTAs'
        asm volatile("fence.proxy.async;\n" ::: "memory");  // bulk tensor loads issued below see them
        for (int q = 0; q < nq; ++q) {
          if (pair_of(q) != pr) continue;
          for (int kc = 0; kc < (2 * q + 2) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * (I8_STAGES + s), ph ^ 1);  // both CTAs of the pair have finished with this stage
            const uint32_t full = bar0 + 8 * s;
            mbar_expect_tx(full, I8_STAGE);
            for (int pl = cr; pl < I8_S; pl += 2)
              tma_load_3d_mc(s0 + s * I8_STAGE + pl * I8_A_SLICE, &tmA1, kc * I8_KC, buf_row0 + 128 * static_cast<int>(b), pl,
                             full, static_cast<uint16_t>(3));
            tma_load_3d(s0 + s * I8_STAGE + I8_A_BYTES, &tmB, kc * I8_KC, (2 * q + cr) * 64, 0, full);
          }
        }
      }
    } else if (warp == 13) {
      // ===== tensor-instruction issuer: the whole warp walks the loops, one elected lane issues =====
      uint32_t it = 0, acc_round = 0, ring = 0;
      for (long long tile = g; tile < ntiles; tile += groups)
        for (int q = 0; q < nq; ++q) {
          if (pair_of(q) != pr) continue;
          mbar_wait(bar_acce, (acc_round & 1) ^ 1);        // the previous half-block has been drained
          ++acc_round;
          asm volatile("tcgen05.fence::after_thread_sync;\n");
          for (int kc = 0; kc < (2 * q + 2) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;\n");
            const uint32_t a0 = s0 + s * I8_STAGE;
            const uint64_t da0 = umma_desc_k(a0), db0 = umma_desc_k(a0 + I8_A_BYTES);
            const uint32_t fresh = (kc != 0);              // the first chunk of a half-block overwrites the accumulators
            if (elect_one()) {
              if (!PLX_DBG(p.dbg, 64)) issue_stage(tmem, da0, db0, fresh, ring, ip.a_from_smem != 0);
              umma_commit_mc(bar0 + 8 * (I8_STAGES + s), static_cast<uint16_t>(3));   // stage free in BOTH CTAs once read here
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
  cluster_sync_all();                                      // the peer may still be sending to this CTA's barriers and stages
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(512));
}

}  // namespace plx
