// Stand-alone probe (not part of the library): one 128 x 128 x 128 int8 tile product on the 5th-generation
// tensor cores (tcgen05.mma kind::i8, int32 accumulators in TMEM), checked against the host, and the issue
// rate of the same instruction with operands resident in shared memory.  It exists to size the
// "FP64 by exact int8 slices" route described in DESIGN.md (next round), nothing in the product uses it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/i8_umma_gemm tools/i8_umma_gemm.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// shared-memory matrix descriptor: K-major rows of 128 bytes, 128-byte swizzle, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// instruction descriptor, kind::i8: D = s32, A = B = signed 8 bit, both K-major, N = 128, M = 128
constexpr uint32_t IDESC_I8 = (2u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      :: "r"(tmem_c), "l"(da), "l"(db), "r"(IDESC_I8), "r"(accumulate), "r"(0u) : "memory");
}

__global__ void __launch_bounds__(128, 1) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                int32_t* __restrict__ C, int repeats) {
  __shared__ __align__(1024) uint8_t sA[128 * 128];
  __shared__ __align__(1024) uint8_t sB[128 * 128];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // operands: row r, 16-byte chunk c -> r * 128 + ((c ^ (r & 7)) << 4)   (the layout a SWIZZLE_128B TMA load writes)
  for (int idx = tid; idx < 128 * 8; idx += 128) {
    const int r = idx >> 3, c = idx & 7;
    const int4 va = *reinterpret_cast<const int4*>(A + r * 128 + c * 16);
    const int4 vb = *reinterpret_cast<const int4*>(B + r * 128 + c * 16);
    *reinterpret_cast<int4*>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = va;
    *reinterpret_cast<int4*>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = vb;
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tmem_base_s)), "n"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_base_s;
  if (warp == 0 && lane == 0) {
    const uint64_t da = umma_desc(smem_u32(sA)), db = umma_desc(smem_u32(sB));
    for (int rep = 0; rep < repeats; ++rep)
      for (int k = 0; k < 4; ++k)              // 4 x K=32 bytes inside the 128-byte swizzle row
        umma_i8(tmem, da + 2 * k, db + 2 * k, (rep | k) != 0);
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n" :: "l"(__cvta_generic_to_shared(&bar)) : "memory");
  }
  {
    const uint32_t b = smem_u32(&bar);
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n"
        :: "r"(b), "r"(0u) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  // accumulator row m lives in TMEM lane m; warp w may read lanes 32 w .. 32 w + 31
  if (C != nullptr) {
    int32_t* out = C + static_cast<size_t>(blockIdx.x) * 128 * 128 + static_cast<size_t>(warp * 32 + lane) * 128;
    for (int c0 = 0; c0 < 128; c0 += 16) {
      uint32_t v[16];
      const uint32_t taddr = tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
            "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      for (int i = 0; i < 16; ++i) out[c0 + i] = static_cast<int32_t>(v[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(128));
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main() {
  std::vector<int8_t> A(128 * 128), B(128 * 128);
  srand(7);
  for (auto& v : A) v = static_cast<int8_t>(rand() % 129 - 64);
  for (auto& v : B) v = static_cast<int8_t>(rand() % 129 - 64);
  int8_t *dA, *dB; int32_t* dC;
  CK(cudaMalloc(&dA, A.size())); CK(cudaMalloc(&dB, B.size())); CK(cudaMalloc(&dC, 128 * 128 * 4));
  CK(cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice));
  CK(cudaMemset(dC, 0xff, 128 * 128 * 4));
  probe<<<1, 128>>>(dA, dB, dC, 1);
  CK(cudaDeviceSynchronize());
  std::vector<int32_t> C(128 * 128);
  CK(cudaMemcpy(C.data(), dC, C.size() * 4, cudaMemcpyDeviceToHost));
  long long bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 128; ++n) {
      int32_t s = 0;
      for (int k = 0; k < 128; ++k) s += static_cast<int32_t>(A[m * 128 + k]) * B[n * 128 + k];
      if (s != C[m * 128 + n]) { if (bad < 5) printf("mismatch (%d,%d): host %d device %d\n", m, n, s, C[m * 128 + n]); ++bad; }
    }
  printf("{\"check\": \"int8 128x128x128 tile, C[m][n] = sum_k A[m][k] B[n][k]\", \"mismatches\": %lld}\n", bad);
  // issue rate: every SM repeats the 4-instruction K loop on resident operands
  int dev = 0, sms = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int repeats = 20000;
  probe<<<sms, 128>>>(dA, dB, nullptr, 1000);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  probe<<<sms, 128>>>(dA, dB, nullptr, repeats);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double ops = 2.0 * 128 * 128 * 128 * static_cast<double>(repeats) * sms;
  printf("{\"metric\": \"tcgen05.mma kind::i8 m128n128k32, operands resident in shared memory\", \"sms\": %d, \"ms\": %.3f, \"tops\": %.1f}\n",
         sms, ms, ops / (ms * 1e-3) / 1e12);
  return bad != 0;
}
