// TMA + mbarrier plumbing shared by the sweep kernel and the fit's tile kernels: barrier helpers, bulk
// tensor loads, and the DMMA consumer of one 128 x 128 x 16 chunk staged with the 128-byte swizzle.
#pragma once
#include "mma_f64.cuh"
#include <cuda.h>

namespace plx {

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
      :: "r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" :: "r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int x, int y, int z, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n"
      :: "r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}

// mma_chunk with the next B fragment in flight while the current one feeds the tensor pipe
__device__ __forceinline__ void mma_chunk_pf(WarpAcc& acc, uint32_t sA, uint32_t sB, int warp, int lane, int nf_begin) {
  const int g = lane >> 2, t = lane & 3;
  const int wm = (warp & 3) * 32, wn = (warp >> 2) * 64;
  const uint32_t rowA = sA + static_cast<uint32_t>((wm + g) * 128);
  const uint32_t rowB = sB + static_cast<uint32_t>((wn + g) * 128);
  const uint32_t c0 = static_cast<uint32_t>(((2 * t) ^ g) << 4), c1 = c0 ^ 16u;
  double2 a0[4], a1[4];
#pragma unroll
  for (int mf = 0; mf < 4; ++mf) { a0[mf] = lds128(rowA + mf * 1024 + c0); a1[mf] = lds128(rowA + mf * 1024 + c1); }
  const int first = nf_begin < 8 ? nf_begin : 7;
  double2 b0 = lds128(rowB + first * 1024 + c0), b1 = lds128(rowB + first * 1024 + c1);
#pragma unroll
  for (int nf = 0; nf < 8; ++nf) {
    if (nf < nf_begin) continue;
    double2 n0 = b0, n1 = b1;
    if (nf + 1 < 8) { n0 = lds128(rowB + (nf + 1) * 1024 + c0); n1 = lds128(rowB + (nf + 1) * 1024 + c1); }
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) dmma(acc.c[mf][nf][0], acc.c[mf][nf][1], a0[mf].x, b0.x);
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) dmma(acc.c[mf][nf][0], acc.c[mf][nf][1], a0[mf].y, b0.y);
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) dmma(acc.c[mf][nf][0], acc.c[mf][nf][1], a1[mf].x, b1.x);
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) dmma(acc.c[mf][nf][0], acc.c[mf][nf][1], a1[mf].y, b1.y);
    b0 = n0; b1 = n1;
  }
}

}  // namespace plx
