// FP64 tensor-core (DMMA) building blocks shared by the sweep and the fit kernels.
//
// Everything on this path is float64 (the reference computes in numpy float64).  tcgen05 has no
// f64 kind, so the FP64 tensor path on sm_100a is mma.sync.m8n8k4 (SASS: DMMA.8x8x4).
//
// Tile convention ("NT" product): C[m][n] += sum_k A[m][k] * B[n][k], both operands K-major in
// global memory.  A CTA tile is 128 x 128, staged in shared memory as K-chunks of 16 doubles:
// one row = 16 doubles = 128 bytes, rows stored densely, and the 16-byte chunk index XOR-ed with
// (row & 7) -- exactly the layout a TMA load with CU_TENSOR_MAP_SWIZZLE_128B produces, so the
// TMA-fed sweep kernel and the plain-load fit kernels share one consumer.
//
// K permutation: within a 16-wide chunk, MMA step s (0..3) assigns k = 4*t + s to fragment slot
// t = lane & 3 (instead of k = 4*s + t).  Both operands use the same permutation, so the product
// is unchanged, and each thread's four A (or B) values for the chunk are 32 contiguous bytes: two
// conflict-free 16-byte shared loads.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace plx {

constexpr int TILE = 128;          // CTA tile edge (rows of A, rows of B)
constexpr int KCH = 16;            // doubles per K-chunk (one 128-byte swizzle row)
constexpr int TILE_BYTES = TILE * KCH * 8;   // 16 KiB per operand per stage

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ double2 lds128(uint32_t addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "r"(addr));
  return v;
}

__device__ __forceinline__ double lds64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts64(uint32_t addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;\n" :: "r"(addr), "d"(v));
}

// coherent 16-byte global load (never the non-coherent path: several kernels re-read scratch that
// the same CTA wrote earlier in the launch)
__device__ __forceinline__ double2 ldg128(const double* p) {
  double2 v;
  asm volatile("ld.global.v2.f64 {%0,%1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

__device__ __forceinline__ void sts128(uint32_t addr, double2 v) {
  asm volatile("st.shared.v2.f64 [%0], {%1,%2};\n" :: "r"(addr), "d"(v.x), "d"(v.y));
}

// byte offset of 16-byte chunk `chunk` (0..7) of row `row` inside a swizzled tile
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
  return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4));
}

// Accumulators of one warp: 32 (M) x 64 (N) as 4 x 8 DMMA tiles.
struct WarpAcc {
  double c[4][8][2];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) { c[i][j][0] = 0.0; c[i][j][1] = 0.0; }
  }
};

// One 16-wide K-chunk of the CTA tile product.  8 warps: warp&3 selects the 32-row M slab,
// warp>>2 the 64-row N slab.  `nf_begin`: number of leading 8-wide N fragments whose B rows are
// structurally zero for this chunk (0 = none); lets triangular callers skip them.
__device__ __forceinline__ void mma_chunk(WarpAcc& acc, uint32_t sA, uint32_t sB, int warp, int lane,
                                          int nf_begin = 0) {
  const int g = lane >> 2, t = lane & 3;
  const int wm = (warp & 3) * 32, wn = (warp >> 2) * 64;
  const uint32_t c0 = static_cast<uint32_t>(((2 * t) ^ g) << 4);
  const uint32_t c1 = c0 ^ 16u;
  double a[4][4];
#pragma unroll
  for (int mf = 0; mf < 4; ++mf) {
    const uint32_t base = sA + static_cast<uint32_t>((wm + mf * 8 + g) * 128);
    const double2 v0 = lds128(base + c0), v1 = lds128(base + c1);
    a[mf][0] = v0.x; a[mf][1] = v0.y; a[mf][2] = v1.x; a[mf][3] = v1.y;
  }
#pragma unroll
  for (int nf = 0; nf < 8; ++nf) {
    if (nf < nf_begin) continue;
    const uint32_t base = sB + static_cast<uint32_t>((wn + nf * 8 + g) * 128);
    const double2 v0 = lds128(base + c0), v1 = lds128(base + c1);
    const double b[4] = {v0.x, v0.y, v1.x, v1.y};
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int mf = 0; mf < 4; ++mf) dmma(acc.c[mf][nf][0], acc.c[mf][nf][1], a[mf][s], b[s]);
  }
}

// Element (m, n) coordinates held by this thread: acc.c[mf][nf][e] is
//   m = (warp&3)*32 + mf*8 + (lane>>2),   n = (warp>>2)*64 + nf*8 + (lane&3)*2 + e.
__device__ __forceinline__ int acc_row(int warp, int lane, int mf) { return (warp & 3) * 32 + mf * 8 + (lane >> 2); }
__device__ __forceinline__ int acc_col(int warp, int lane, int nf, int e) { return (warp >> 2) * 64 + nf * 8 + (lane & 3) * 2 + e; }

// ---------------------------------------------------------------------------------------------
// Plain-load main loop (fit kernels and the non-TMA sweep variant): 256 threads stream K-chunks
// of A[rowA0.., k] and B[rowB0.., k] (row-major, leading dimensions lda/ldb in doubles, both 16-byte
// aligned with even ld) through two swizzled shared stages, prefetching the next chunk into
// registers while the tensor pipe works on the current one.  k0 and k1 are multiples of 16.
// `smem` needs 4 * TILE_BYTES bytes, 1024-byte aligned.
// ---------------------------------------------------------------------------------------------
struct ChunkRegs { double2 a[4], b[4]; };

__device__ __forceinline__ void chunk_fetch(ChunkRegs& r, const double* __restrict__ A, size_t lda,
                                            const double* __restrict__ B, size_t ldb, int k, int tid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = tid + i * 256;          // 1024 16-byte chunks per operand
    const int row = idx >> 3, ch = idx & 7;
    r.a[i] = ldg128(A + static_cast<size_t>(row) * lda + k + ch * 2);
    r.b[i] = ldg128(B + static_cast<size_t>(row) * ldb + k + ch * 2);
  }
}

__device__ __forceinline__ void chunk_store(const ChunkRegs& r, uint32_t sA, uint32_t sB, int tid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = tid + i * 256;
    const int row = idx >> 3, ch = idx & 7;
    sts128(sA + swz(row, ch), r.a[i]);
    sts128(sB + swz(row, ch), r.b[i]);
  }
}

__device__ __forceinline__ void mainloop_nt(WarpAcc& acc, const double* __restrict__ A, size_t lda,
                                            const double* __restrict__ B, size_t ldb, int k0, int k1,
                                            unsigned char* smem) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  if (k0 >= k1) return;
  ChunkRegs r;
  chunk_fetch(r, A, lda, B, ldb, k0, tid);
  int buf = 0;
  for (int k = k0; k < k1; k += KCH) {
    const uint32_t sA = s0 + buf * 2 * TILE_BYTES, sB = sA + TILE_BYTES;
    chunk_store(r, sA, sB, tid);
    __syncthreads();
    if (k + KCH < k1) chunk_fetch(r, A, lda, B, ldb, k + KCH, tid);
    mma_chunk(acc, sA, sB, warp, lane);
    buf ^= 1;
  }
  __syncthreads();
}

}  // namespace plx
