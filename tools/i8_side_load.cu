// Probe: what slows the int8 tensor-instruction stream of the sweep when the digit-plane builders run beside it?
// Warp 0 issues the sweep's stage (tools/i8_umma_ts.cu mode 4: A planes through the tensor-memory ring, commit per stage,
// two-stage window) on resident operands; eight more warps run a side load until the stream has finished:
//   0 nothing   1 FP64 FMA chains in registers   2 128-bit shared-memory loads (conflict-free)   3 both (the builders' mix:
//   6 loads per 64 FP64 operations)   4 integer ALU   5 global (L2) stores of 32-bit words   6 FP64 tensor (mma.sync m8n8k4)
//   7 FP32 FMA chains
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/i8_side_load tools/i8_side_load.cu
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


constexpr int SIDE_BYTES = 49152;
__global__ void __launch_bounds__(384, 1) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B, int side, int repeats,
                                                double* sink, unsigned* gbuf, unsigned long long* side_iters, int nomma) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + S * A_PLANE;
  double* sS = reinterpret_cast<double*>(smem + S * A_PLANE + S * B_PLANE);       // side buffer
  __shared__ __align__(8) unsigned long long bar;
  __shared__ __align__(8) unsigned long long sbar[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int done;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int idx = tid; idx < S * 128 * 4; idx += 384) {
    const int pl = idx / 512, r = (idx >> 2) & 127, c = idx & 3;
    *reinterpret_cast<int4*>(sA + pl * A_PLANE + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) =
        *reinterpret_cast<const int4*>(A + pl * A_PLANE + r * 64 + c * 16);
  }
  for (int idx = tid; idx < S * 64 * 4; idx += 384) {
    const int pl = idx / 256, r = (idx >> 2) & 63, c = idx & 3;
    *reinterpret_cast<int4*>(sB + pl * B_PLANE + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) =
        *reinterpret_cast<const int4*>(B + pl * B_PLANE + r * 64 + c * 16);
  }
  for (int i = tid; i < SIDE_BYTES / 8; i += 384) sS[i] = 1.0 + 1e-9 * i;
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_u32(&tmem_base_s)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    done = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&bar)));
    for (int i = 0; i < 2; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&sbar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_base_s;
  if (warp == 0) {
    const uint64_t da0 = umma_desc64(smem_u32(sA)), db0 = umma_desc64(smem_u32(sB));
    uint32_t ring = 0;
    if (nomma) {                                            // the same wall time without tensor instructions
      const long long t0 = clock64();
      while (clock64() - t0 < 1800LL * repeats) { }
    }
    for (int rep = 0; rep < (nomma ? 0 : repeats); ++rep) {
      if (rep >= 2) {
        const uint32_t b = smem_u32(&sbar[rep & 1]), parity = ((rep >> 1) - 1) & 1;
        asm volatile("{\n.reg .pred p;\nWAITS_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONES_%=;\nbra WAITS_%=;\nDONES_%=:\n}\n"
                     :: "r"(b), "r"(parity) : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;\n");
      }
      if (elect_one()) {
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
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n"
                     :: "l"(__cvta_generic_to_shared(&sbar[rep & 1])) : "memory");
      }
      ring += 2 * S;
      __syncwarp();
    }
    if (elect_one())
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n" :: "l"(__cvta_generic_to_shared(&bar)) : "memory");
    __syncwarp();
    const uint32_t b = smem_u32(&bar);
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n"
                 :: "r"(b), "r"(0u) : "memory");
    if (lane == 0) done = 1;
  } else if (warp >= 4 && side != 0) {
    const int st = tid - 128;                               // 0 .. 255
    double a0 = 1.0, a1 = 1.1, a2 = 1.2, a3 = 1.3, a4 = 0.9, a5 = 0.8, a6 = 0.7, a7 = 0.6;
    unsigned iacc = st;
    float f0 = 1.f, f1 = 1.1f, f2 = 1.2f, f3 = 1.3f, f4 = .9f, f5 = .8f, f6 = .7f, f7 = .6f;
    const double* mine = sS + 2 * (st & 255);
    long long it = 0;
    while (!done) {
      for (int u = 0; u < 64; ++u) {
        if (side == 1 || side == 3) {
          a0 = fma(a0, 1.0000001, 1e-9); a1 = fma(a1, 1.0000001, 1e-9); a2 = fma(a2, 0.9999999, 1e-9); a3 = fma(a3, 0.9999999, 1e-9);
          a4 = fma(a4, 1.0000001, 1e-9); a5 = fma(a5, 1.0000001, 1e-9); a6 = fma(a6, 0.9999999, 1e-9); a7 = fma(a7, 0.9999999, 1e-9);
        }
        if (side == 2 || (side == 3 && (u & 7) < 6 * 8 / 64 + 1)) {
          double vx, vy;
          asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];\n" : "=d"(vx), "=d"(vy)
                       : "r"(smem_u32(mine + ((u * 512) % (SIDE_BYTES / 8 - 512)))));
          a0 += vx * 1e-30; a1 += vy * 1e-30;
        }
        if (side == 6) {
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};\n" : "+d"(a0), "+d"(a1) : "d"(a4), "d"(a5));
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};\n" : "+d"(a2), "+d"(a3) : "d"(a6), "d"(a7));
        }
        if (side == 7) {
          f0 = fmaf(f0, 1.0000001f, 1e-9f); f1 = fmaf(f1, 1.0000001f, 1e-9f); f2 = fmaf(f2, 0.9999999f, 1e-9f); f3 = fmaf(f3, 0.9999999f, 1e-9f);
          f4 = fmaf(f4, 1.0000001f, 1e-9f); f5 = fmaf(f5, 1.0000001f, 1e-9f); f6 = fmaf(f6, 0.9999999f, 1e-9f); f7 = fmaf(f7, 0.9999999f, 1e-9f);
        }
        if (side == 4) { iacc = iacc * 1664525u + 1013904223u; iacc ^= iacc >> 7; iacc += u; iacc = (iacc << 3) | (iacc >> 29); }
        if (side == 5) gbuf[(blockIdx.x * 256 + st) + 65536 * ((it * 64 + u) & 63)] = iacc + u;
      }
      ++it;
    }
    if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + iacc + f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7 == 12345.678) sink[0] = a0;
    if (lane == 0) atomicAdd(side_iters, static_cast<unsigned long long>(it));
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
  int8_t *dA, *dB; double* sink; unsigned* gbuf; unsigned long long* iters;
  CK(cudaMalloc(&dA, A.size())); CK(cudaMalloc(&dB, B.size())); CK(cudaMalloc(&sink, 64)); CK(cudaMalloc(&gbuf, 4 * 65536 * 64)); CK(cudaMalloc(&iters, 8));
  CK(cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice));
  const int smem = S * A_PLANE + S * B_PLANE + SIDE_BYTES + 1024;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int repeats = 4000;
  const char* names[] = {"none", "fp64 fma", "lds.128", "fp64 + lds (builder mix)", "integer alu", "global stores", "fp64 tensor (dmma)", "fp32 fma"};
  for (int nomma = 0; nomma < 2; ++nomma)
    for (int side = 0; side < 8; ++side) {
      probe<<<sms, 384, smem>>>(dA, dB, side, 200, sink, gbuf, iters, nomma); CK(cudaDeviceSynchronize());
      CK(cudaMemset(iters, 0, 8));
      CK(cudaEventRecord(e0)); probe<<<sms, 384, smem>>>(dA, dB, side, repeats, sink, gbuf, iters, nomma); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
      unsigned long long h = 0; CK(cudaMemcpy(&h, iters, 8, cudaMemcpyDeviceToHost));
      printf("{\"tensor stream\": %s, \"side load\": \"%s\", \"ms\": %.3f, \"tops\": %.1f, \"side iterations per warp per ms\": %.0f}\n",
             nomma ? "false" : "true", names[side], ms, nomma ? 0.0 : 2.0 * 128 * 64 * 32 * 56 * static_cast<double>(repeats) * sms / (ms * 1e-3) / 1e12,
             h / (8.0 * sms) / ms);
    }
  return 0;
}
