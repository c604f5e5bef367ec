// Probe variant: m128 n64, operand rows of 64 bytes in the SWIZZLE_64B layout (two k32 steps per row).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/i8_umma_n64 tools/i8_umma_n64.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
// K-major, 64-byte rows, 64-byte swizzle: 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t umma_desc64(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
constexpr uint32_t IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ void umma_i8(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
               :: "r"(tmem_c), "l"(da), "l"(db), "r"(IDESC), "r"(accumulate), "r"(0u) : "memory");
}
__global__ void __launch_bounds__(128, 1) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                int32_t* __restrict__ C, int repeats) {
  __shared__ __align__(1024) uint8_t sA[128 * 64];
  __shared__ __align__(1024) uint8_t sB[64 * 64];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int idx = tid; idx < 128 * 4; idx += 128) {
    const int r = idx >> 2, c = idx & 3;
    *reinterpret_cast<int4*>(sA + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) = *reinterpret_cast<const int4*>(A + r * 64 + c * 16);
  }
  for (int idx = tid; idx < 64 * 4; idx += 128) {
    const int r = idx >> 2, c = idx & 3;
    *reinterpret_cast<int4*>(sB + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) = *reinterpret_cast<const int4*>(B + r * 64 + c * 16);
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tmem_base_s)), "n"(64));
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
  if (tid == 0) {
    const uint64_t da = umma_desc64(smem_u32(sA)), db = umma_desc64(smem_u32(sB));
    for (int rep = 0; rep < repeats; ++rep)
      for (int k = 0; k < 2; ++k) umma_i8(tmem, da + 2 * k, db + 2 * k, (rep | k) != 0);
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n" :: "l"(__cvta_generic_to_shared(&bar)) : "memory");
  }
  {
    const uint32_t b = smem_u32(&bar);
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n"
                 :: "r"(b), "r"(0u) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  if (C != nullptr) {
    int32_t* out = C + static_cast<size_t>(warp * 32 + lane) * 64;
    for (int c0 = 0; c0 < 64; c0 += 16) {
      uint32_t v[16];
      const uint32_t taddr = tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                     "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                   : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      for (int i = 0; i < 16; ++i) out[c0 + i] = static_cast<int32_t>(v[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(64));
}
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
int main() {
  std::vector<int8_t> A(128 * 64), B(64 * 64);
  srand(7);
  for (auto& v : A) v = static_cast<int8_t>(rand() % 129 - 64);
  for (auto& v : B) v = static_cast<int8_t>(rand() % 129 - 64);
  int8_t *dA, *dB; int32_t* dC;
  CK(cudaMalloc(&dA, A.size())); CK(cudaMalloc(&dB, B.size())); CK(cudaMalloc(&dC, 128 * 64 * 4));
  CK(cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice));
  probe<<<1, 128>>>(dA, dB, dC, 1);
  CK(cudaDeviceSynchronize());
  std::vector<int32_t> C(128 * 64);
  CK(cudaMemcpy(C.data(), dC, C.size() * 4, cudaMemcpyDeviceToHost));
  long long bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 64; ++n) {
      int32_t s = 0;
      for (int k = 0; k < 64; ++k) s += static_cast<int32_t>(A[m * 64 + k]) * B[n * 64 + k];
      if (s != C[m * 64 + n]) { if (bad < 5) printf("mismatch (%d,%d): host %d device %d\n", m, n, s, C[m * 64 + n]); ++bad; }
    }
  printf("{\"check\": \"int8 m128 n64 k64, SWIZZLE_64B rows\", \"mismatches\": %lld}\n", bad);
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int repeats = 40000;
  probe<<<sms, 128>>>(dA, dB, nullptr, 1000); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0)); probe<<<sms, 128>>>(dA, dB, nullptr, repeats); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
  printf("{\"metric\": \"kind::i8 m128n64k32 resident operands\", \"ms\": %.3f, \"tops\": %.1f}\n", ms,
         2.0 * 128 * 64 * 64 * static_cast<double>(repeats) * sms / (ms * 1e-3) / 1e12);
  return bad != 0;
}
