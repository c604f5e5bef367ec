// Probe: A operand of tcgen05.mma kind::i8 from tensor memory.  One stage of the int8-sliced sweep is emulated:
// 7 digit planes of a 128-candidate tile (A, 128 x 64 B rows, SWIZZLE_64B) and 7 planes of a 64-row half-block of
// L^-1 (B, 64 x 64 B rows) resident in shared memory, 28 plane pairs (s + t < 7) x 2 k32 steps into 7 x 64 TMEM
// columns.  Modes:
//   0  both operands from shared memory (the shipped kernel's instruction stream)
//   1  per k32 step: 7 x tcgen05.cp.128x256b of the A planes into TMEM columns 448..503, then 28 MMAs with A in TMEM
//   2  the same with an 8-slot ring of A planes in columns 448..511: the copy of plane s + 1 is issued before the
//      MMAs of plane s, slots rotate across k steps (no waits: copies and MMAs execute in issue order)
//   3  mode 2 with a tcgen05.commit to an mbarrier after every stage (no waits)
//   4  mode 3 with a window: stage i is issued only after the commit of stage i - 2 has arrived (what a two-stage
//      operand ring imposes, with the refill itself free)
//   5  the same with a window of three stages, 6 with four
// Checks every mode against the host, then times `repeats` stages per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/i8_umma_ts tools/i8_umma_ts.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

constexpr int S = 7;
constexpr int A_PLANE = 128 * 64, B_PLANE = 64 * 64;
constexpr int SMEM_BYTES = S * A_PLANE + S * B_PLANE + 1024;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t umma_desc64(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
constexpr uint32_t IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ void umma_ss(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
               :: "r"(tmem_c), "l"(da), "l"(db), "r"(IDESC), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void umma_ts(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
               :: "r"(tmem_c), "r"(tmem_a), "l"(db), "r"(IDESC), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t tmem_dst, uint64_t desc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;\n" :: "r"(tmem_dst), "l"(desc) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
  return pred != 0;
}

__global__ void __launch_bounds__(128, 1) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                int32_t* __restrict__ C, int32_t* __restrict__ Adump, int mode, int repeats) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + S * A_PLANE;
  __shared__ __align__(8) unsigned long long bar;
  __shared__ __align__(8) unsigned long long sbar[4];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int idx = tid; idx < S * 128 * 4; idx += 128) {
    const int pl = idx / 512, r = (idx >> 2) & 127, c = idx & 3;
    *reinterpret_cast<int4*>(sA + pl * A_PLANE + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) =
        *reinterpret_cast<const int4*>(A + pl * A_PLANE + r * 64 + c * 16);
  }
  for (int idx = tid; idx < S * 64 * 4; idx += 128) {
    const int pl = idx / 256, r = (idx >> 2) & 63, c = idx & 3;
    *reinterpret_cast<int4*>(sB + pl * B_PLANE + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) =
        *reinterpret_cast<const int4*>(B + pl * B_PLANE + r * 64 + c * 16);
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tmem_base_s)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&bar)));
    for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&sbar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_base_s;
  if (warp == 0) {
    const uint64_t da0 = umma_desc64(smem_u32(sA)), db0 = umma_desc64(smem_u32(sB));
    uint32_t ring = 0;
    const int window = mode == 4 ? 2 : (mode == 5 ? 3 : (mode == 6 ? 4 : 0));
    for (int rep = 0; rep < repeats; ++rep) {
      if (window > 0 && rep >= window) {
        // the commit of stage rep - window: barrier (rep % window), its (rep / window - 1)-th completion
        const uint32_t b = smem_u32(&sbar[rep % window]), parity = ((rep / window) - 1) & 1;
        asm volatile("{\n.reg .pred p;\nWAITS_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONES_%=;\nbra WAITS_%=;\nDONES_%=:\n}\n"
                     :: "r"(b), "r"(parity) : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;\n");
      }
      if (elect_one()) {
        if (mode == 0) {
#pragma unroll
          for (int t = 0; t < S; ++t)
#pragma unroll
            for (int s = 0; s + t < S; ++s) {
              const uint64_t da = da0 + s * (A_PLANE / 16), db = db0 + t * (B_PLANE / 16);
              umma_ss(tmem + (s + t) * 64, da, db, t == 0 ? (rep != 0) : 1u);
              umma_ss(tmem + (s + t) * 64, da + 2, db + 2, 1u);
            }
        } else if (mode == 1) {
#pragma unroll
          for (int k = 0; k < 2; ++k) {
#pragma unroll
            for (int s = 0; s < S; ++s) tmem_cp_128x256b(tmem + 448 + 8 * s, da0 + s * (A_PLANE / 16) + 2 * k);
#pragma unroll
            for (int s = 0; s < S; ++s)
#pragma unroll
              for (int t = 0; s + t < S; ++t)
                umma_ts(tmem + (s + t) * 64, tmem + 448 + 8 * s, db0 + t * (B_PLANE / 16) + 2 * k,
                        (s == 0 && k == 0) ? (rep != 0) : 1u);
          }
        } else {
          // ring: plane p of the flattened (k, s) sequence lives in slot (ring + p) % 8; the copy runs one plane ahead
          tmem_cp_128x256b(tmem + 448 + 8 * (ring & 7), da0);
#pragma unroll
          for (int p = 0; p < 2 * S; ++p) {
            const int k = p / S, s = p % S;
            if (p + 1 < 2 * S) {
              const int k1 = (p + 1) / S, s1 = (p + 1) % S;
              tmem_cp_128x256b(tmem + 448 + 8 * ((ring + p + 1) & 7), da0 + s1 * (A_PLANE / 16) + 2 * k1);
            }
            const uint32_t ta = tmem + 448 + 8 * ((ring + p) & 7);
#pragma unroll
            for (int t = 0; s + t < S; ++t)
              umma_ts(tmem + (s + t) * 64, ta, db0 + t * (B_PLANE / 16) + 2 * k, (s == 0 && k == 0) ? (rep != 0) : 1u);
          }
          ring += 2 * S;
          if (mode >= 3) {
            const int w = window > 0 ? window : 4;
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n"
                         :: "l"(__cvta_generic_to_shared(&sbar[rep % w])) : "memory");
          }
        }
      }
      __syncwarp();
    }
    if (elect_one())
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n" :: "l"(__cvta_generic_to_shared(&bar)) : "memory");
    __syncwarp();
  }
  {
    const uint32_t b = smem_u32(&bar);
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n"
                 :: "r"(b), "r"(0u) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  if (C != nullptr) {
    for (int c0 = 0; c0 < 512; c0 += 16) {
      uint32_t v[16];
      const uint32_t taddr = tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                     "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                   : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      const size_t row = static_cast<size_t>(warp * 32 + lane);
      for (int i = 0; i < 16; ++i) {
        if (c0 < 448) C[row * 448 + c0 + i] = static_cast<int32_t>(v[i]);
        else Adump[row * 64 + (c0 - 448) + i] = static_cast<int32_t>(v[i]);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(512));
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
int main() {
  std::vector<int8_t> A(S * A_PLANE), B(S * B_PLANE);
  srand(7);
  for (auto& v : A) v = static_cast<int8_t>(rand() % 255 - 127);
  for (auto& v : B) v = static_cast<int8_t>(rand() % 255 - 127);
  int8_t *dA, *dB; int32_t *dC, *dAd;
  CK(cudaMalloc(&dA, A.size())); CK(cudaMalloc(&dB, B.size())); CK(cudaMalloc(&dC, 128 * 448 * 4)); CK(cudaMalloc(&dAd, 128 * 64 * 4));
  CK(cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  // host: class c accumulates every plane pair s + t = c over the 64-byte rows
  std::vector<int32_t> want(128 * 448, 0);
  for (int s = 0; s < S; ++s)
    for (int t = 0; s + t < S; ++t)
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n) {
          int32_t acc = 0;
          for (int k = 0; k < 64; ++k) acc += static_cast<int32_t>(A[s * A_PLANE + m * 64 + k]) * B[t * B_PLANE + n * 64 + k];
          want[m * 448 + (s + t) * 64 + n] += acc;
        }
  int rc = 0;
  for (int mode = 0; mode < 7; ++mode) {
    for (int reps : {1, 3}) {
      CK(cudaMemset(dC, 0xff, 128 * 448 * 4));
      probe<<<1, 128, SMEM_BYTES>>>(dA, dB, dC, dAd, mode, reps);
      CK(cudaDeviceSynchronize());
      std::vector<int32_t> C(128 * 448), Ad(128 * 64);
      CK(cudaMemcpy(C.data(), dC, C.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(Ad.data(), dAd, Ad.size() * 4, cudaMemcpyDeviceToHost));
      long long bad = 0;
      for (size_t i = 0; i < C.size(); ++i)
        if (C[i] != reps * want[i]) { if (bad < 3) printf("mode %d reps %d mismatch at row %zu col %zu: host %d device %d\n", mode, reps, i / 448, i % 448, reps * want[i], C[i]); ++bad; }
      printf("{\"check\": \"mode %d, %d stage(s)\", \"mismatches\": %lld}\n", mode, reps, bad);
      if (mode == 1 && reps == 1) {
        // slot s holds plane s of the second k32 step: lane r, column j = bytes 32 + 4j .. 32 + 4j + 3 of row r
        long long lay_bad = 0;
        for (int s = 0; s < S; ++s)
          for (int r = 0; r < 128; ++r)
            for (int j = 0; j < 8; ++j) {
              int32_t w;
              memcpy(&w, &A[s * A_PLANE + r * 64 + 32 + 4 * j], 4);
              if (Ad[r * 64 + 8 * s + j] != w) ++lay_bad;
            }
        printf("{\"check\": \"tcgen05.cp.128x256b layout: lane = row, column j = bytes 4j..4j+3 of the k32 slice\", \"mismatches\": %lld}\n", lay_bad);
        if (lay_bad) { printf("row0 slot0 words:"); for (int j = 0; j < 8; ++j) printf(" %08x", Ad[j]); printf("\n"); }
      }
      if (bad) rc = 1;
    }
  }
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int repeats = 4000;
  for (int mode = 0; mode < 7; ++mode) {
    probe<<<sms, 128, SMEM_BYTES>>>(dA, dB, nullptr, nullptr, mode, 200); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); probe<<<sms, 128, SMEM_BYTES>>>(dA, dB, nullptr, nullptr, mode, repeats); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("{\"metric\": \"28 plane pairs x 2 k32 steps, m128n64k32, mode %d\", \"ms\": %.3f, \"tops\": %.1f}\n", mode, ms,
           2.0 * 128 * 64 * 32 * 56 * static_cast<double>(repeats) * sms / (ms * 1e-3) / 1e12);
  }
  return rc;
}
