// Cooperative form of the fused int8-sliced sweep (sweep_i8f.cuh): groups of I8C_R = 4 CTAs share one candidate
// tile.  The fused kernel is bound by its operand stream: every tile's digit planes (14 KB per candidate at
// n = 2048) are re-read once per 64-row half-block of L^-1, and with one tile per SM in flight (148 x 1.8 MB) they
// fall out of the 126 MB L2 -- 34 GB of DRAM reads per 151 552 candidates.  Here CTA r of a group
//   * builds the planes of every fourth 64-observation block of the tile (its builders do a quarter of phase A),
//   * contracts every fourth half-block of L^-1 (dealt 0 1 2 3 3 2 1 0 ... so that the triangular work balances),
//   * hands its partial sum of squares and partial posterior mean to CTA 0 of the group, which adds the four in
//     fixed order and finishes the tile (variance -> acquisition -> arg-max),
// so only 37 tiles are in flight and their planes stay in L2.  The CTAs of a group meet through three counters in
// global memory per plane buffer (planes ready, partials done, buffer free; release / acquire at gpu scope); all
// CTAs of the grid are resident (one per SM, grid <= SM count), so spinning on them cannot deadlock.
// Arithmetic per candidate is that of the fused kernel except for the order in which the posterior mean and the
// sum of squares are added up.
#pragma once
#include "sweep_i8f.cuh"
#ifdef PLX_PROFILE
#include <cstdio>
#endif

namespace plx {

constexpr int I8C_R = 4;

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;\n" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void spin_until(const unsigned* p, unsigned target) {
  while (ld_acquire_gpu(p) < target) __nanosleep(64);
}
// which CTA of the group contracts half-block h
__device__ __forceinline__ int part_of(int h) { return ((h >> 2) & 1) ? 3 - (h & 3) : (h & 3); }

// gsync: [groups][2 buffers][4] counters {ready, done, free, unused}, zero at launch.
// gpart: [groups][2 buffers][I8C_R][256]: 128 partial sums of squares, then 128 partial means.
template <bool GEN>
__device__ __forceinline__ void sweep_i8c_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const SweepParams& p,
                                               const I8Params& ip, double* gpart, unsigned* gsync) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = align1k(smem_raw);
  I8FTail* tail = reinterpret_cast<I8FTail*>(smem + I8_STAGES * I8_STAGE + F_STAGING);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  const uint32_t bar0 = smem_u32(tail->bars);
  const uint32_t bar_accf = bar0 + 8 * (2 * I8_STAGES), bar_acce = bar_accf + 8;
  const uint32_t bar_ready = bar_acce + 8;                 // two: this CTA's share of a plane buffer is built
  if (tid == 0) {
    for (int s = 0; s < I8_STAGES; ++s) { mbar_init(bar0 + 8 * s, 1); mbar_init(bar0 + 8 * (I8_STAGES + s), 1); }
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
  const uint32_t tmem = tail->tmem_base;

  const int groups = gridDim.x / I8C_R, g = blockIdx.x / I8C_R, r = blockIdx.x % I8C_R;
  const long long ntiles = (ip.N + 127) / 128;
  const int nh = (ip.n + 63) / 64;
  const int buf_row0 = g * 256;                        // the group's two tile buffers: rows buf_row0 + 128 b
  unsigned* sync_g = gsync + static_cast<size_t>(g) * 8;                     // [b][4]
  double* part_g = gpart + static_cast<size_t>(g) * 2 * I8C_R * 256;         // [b][r][256]

  if (warp < 8) {
    // ===== builders: every I8C_R-th 64-observation block of the tile's planes =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
    const uint32_t sm = s0 + I8_STAGES * I8_STAGE;
    uint32_t t = 0;
#ifdef PLX_PROFILE
    long long t_wait = 0, t_build = 0;
#endif
    for (long long tile = g; tile < ntiles; tile += groups, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
#ifdef PLX_PROFILE
      const long long c0 = clock64();
#endif
      if (use > 0) {                                   // every CTA of the group has consumed the tile that used this buffer
        if (tid == 0) spin_until(sync_g + 4 * b + 2, I8C_R * use);
        named_bar_sync(1, 256);
      }
#ifdef PLX_PROFILE
      const long long c1 = clock64();
#endif
      signed char* rows = p.planes + static_cast<size_t>(buf_row0 + 128 * b) * p.npad;
      if (!PLX_DBG(p.dbg, 128) || t < 2) build_tile<GEN>(p, tile, rows, part_g + (b * I8C_R + r) * 256 + 128, sm, tid, r, I8C_R);
      __threadfence();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ready + 8 * b);
#ifdef PLX_PROFILE
      t_wait += c1 - c0; t_build += clock64() - c1;
#endif
    }
#ifdef PLX_PROFILE
    if (PLX_DBG(p.dbg, 2048) && tid == 0 && (blockIdx.x < 4 || blockIdx.x == 77))
      printf("cta %d builders: %u tiles, build %lld clk/tile, wait for a free buffer %lld clk/tile\n", blockIdx.x, t, t_build / max(t, 1u), t_wait / max(t, 1u));
#endif
  } else if (warp < 12) {
    // ===== accumulator drain; CTA 0 of the group also finishes the tile =====
    asm volatile("setmaxnreg.dec.sync.aligned.u32 120;\n");
    const int row = tid - 256;
    Best best = {-INFINITY, 0x7fffffffffffffffLL};
    uint32_t acc_round = 0, t = 0;
    const uint32_t lane_base = static_cast<uint32_t>((warp - 8) * 32) << 16;
    for (long long tile = g; tile < ntiles; tile += groups, ++t) {
      const uint32_t b = t & 1, use = t >> 1;
      // a CTA whose share of the half-blocks is empty (fewer than four of them) has no accumulators to wait for:
      // tie it to the tile through its own builders, or it would run ahead and announce buffers free too early
      mbar_wait(bar_ready + 8 * b, use & 1);
      double sumsq = 0.0;
      for (int h = 0; h < nh; ++h) {
        if (part_of(h) != r) continue;
        mbar_wait(bar_accf, acc_round & 1);
        ++acc_round;
        asm volatile("tcgen05.fence::after_thread_sync;\n");
        drain_halfblock(tmem + lane_base, ip.zscale + h * 64, sumsq, bar_acce);
      }
      // this CTA's share of the tile is done: publish the partial sum of squares (the partial means were
      // written by the builders before the planes were announced)
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
      for (int q = 0; q < I8C_R; ++q) {
        ss += __ldcg(part_g + (b * I8C_R + q) * 256 + row);
        mean += __ldcg(part_g + (b * I8C_R + q) * 256 + 128 + row);
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
#ifdef PLX_PROFILE
      long long t_own = 0, t_others = 0, t_stage = 0;
#endif
      for (long long tile = g; tile < ntiles; tile += groups, ++t) {
        const uint32_t b = t & 1, use = t >> 1;
#ifdef PLX_PROFILE
        const long long c0 = clock64();
#endif
        mbar_wait(bar_ready + 8 * b, use & 1);              // this CTA's blocks of the planes are written ...
#ifdef PLX_PROFILE
        const long long c1 = clock64();
#endif
        red_release_gpu(sync_g + 4 * b, 1u);
        spin_until(sync_g + 4 * b, I8C_R * (use + 1));      // ... and so are the other CTAs'
#ifdef PLX_PROFILE
        const long long c2 = clock64(); t_own += c1 - c0; t_others += c2 - c1;
#endif
        asm volatile("fence.proxy.async;\n" ::: "memory");  // bulk tensor loads issued below see them
        for (int h = 0; h < nh; ++h) {
          if (part_of(h) != r) continue;
          for (int kc = 0; kc < (h + 1) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
            mbar_wait(bar0 + 8 * (I8_STAGES + s), ph ^ 1);
            const uint32_t full = bar0 + 8 * s;
#ifdef PLX_PROFILE
            mbar_expect_tx(full, (PLX_DBG(p.dbg, 256) && PLX_DBG(p.dbg, 1024)) ? I8_STAGE / 2 : I8_STAGE);
#else
            mbar_expect_tx(full, I8_STAGE);
#endif
#ifdef PLX_PROFILE
            if (PLX_DBG(p.dbg, 256)) {
              // experiment: the same bytes as two contiguous bulk copies (what a chunk-major plane layout would allow)
              const int nch = p.npad / I8_KC;
              const signed char* srcA = p.planes + (static_cast<size_t>(g * 2 + b) * nch + kc) * I8_A_BYTES;
              const signed char* srcB = ip.zplanes_dbg + (static_cast<size_t>(h) * nch + kc) * I8_B_BYTES;
              // 1024: only half of the bytes are fetched (the rest of the stage is stale): does the incoming stream
              // compete with the tensor core's operand reads for the shared-memory port?
              const uint32_t ab = PLX_DBG(p.dbg, 1024) ? I8_A_BYTES / 2 : I8_A_BYTES, bb = PLX_DBG(p.dbg, 1024) ? I8_B_BYTES / 2 : I8_B_BYTES;
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                           :: "r"(s0 + s * I8_STAGE), "l"(srcA), "r"(ab), "r"(full) : "memory");
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                           :: "r"(s0 + s * I8_STAGE + I8_A_BYTES), "l"(srcB), "r"(bb), "r"(full) : "memory");
              continue;
            }
#endif
            tma_load_3d(s0 + s * I8_STAGE, &tmA, kc * I8_KC, buf_row0 + 128 * static_cast<int>(b), 0, full);
            tma_load_3d(s0 + s * I8_STAGE + I8_A_BYTES, &tmB, kc * I8_KC, h * 64, 0, full);
          }
        }
#ifdef PLX_PROFILE
        t_stage += clock64() - c2;
#endif
      }
#ifdef PLX_PROFILE
      if (PLX_DBG(p.dbg, 2048) && (blockIdx.x < 4 || blockIdx.x == 77))
        printf("cta %d producer: per tile %lld clk waiting for its own builders, %lld for the group's, %lld feeding the ring\n",
               blockIdx.x, t_own / max(t, 1u), t_others / max(t, 1u), t_stage / max(t, 1u));
#endif
    } else if (warp == 13) {
      // ===== tensor-instruction issuer: the whole warp walks the loops, one elected lane issues =====
      uint32_t it = 0, acc_round = 0, ring = 0;
#ifdef PLX_PROFILE
      long long t_acce = 0, t_full = 0, t_all = clock64(), n_tiles = 0;
#endif
      for (long long tile = g; tile < ntiles; tile += groups) {
#ifdef PLX_PROFILE
        ++n_tiles;
#endif
        for (int h = 0; h < nh; ++h) {
          if (part_of(h) != r) continue;
#ifdef PLX_PROFILE
          const long long c0 = clock64();
#endif
          mbar_wait(bar_acce, (acc_round & 1) ^ 1);        // the previous half-block has been drained
#ifdef PLX_PROFILE
          t_acce += clock64() - c0;
#endif
          ++acc_round;
          asm volatile("tcgen05.fence::after_thread_sync;\n");
          for (int kc = 0; kc < (h + 1) * I8_CPH; ++kc, ++it) {
            const uint32_t s = it % I8_STAGES, ph = (it / I8_STAGES) & 1;
#ifdef PLX_PROFILE
            const long long c1 = clock64();
#endif
            mbar_wait(bar0 + 8 * s, ph);
#ifdef PLX_PROFILE
            t_full += clock64() - c1;
#endif
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
#ifdef PLX_PROFILE
      if (PLX_DBG(p.dbg, 2048) && lane == 0 && (blockIdx.x < 4 || blockIdx.x == 77) && n_tiles > 0)
        printf("cta %d issuer: per tile %lld clk in all, %lld waiting for the drain, %lld waiting for a full stage\n",
               blockIdx.x, (clock64() - t_all) / n_tiles, t_acce / n_tiles, t_full / n_tiles);
#endif
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(512));
}

__global__ void __launch_bounds__(512, 1) sweep_i8c_kernel(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const SweepParams p, const I8Params ip, double* gpart,
                                                           unsigned* gsync) {
  sweep_i8c_body<false>(tmA, tmB, p, ip, gpart, gsync);
}
// the general-nu Matern kernel (kernel values from the table of gp_kernels.cuh)
__global__ void __launch_bounds__(512, 1) sweep_i8c_gen_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const SweepParams p, const I8Params ip, double* gpart,
                                                               unsigned* gsync) {
  sweep_i8c_body<true>(tmA, tmB, p, ip, gpart, gsync);
}

}  // namespace plx
