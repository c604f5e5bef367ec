// Stand-alone numerics probe (not part of the library) for the exact-slicing route of DESIGN.md section 8:
//   V[c][i] = sum_j K[c][j] * Z[i][j]   (128 candidates x 128 rows of L^-1, contraction over n observations)
// computed from signed 7-bit slices of both operands with tcgen05.mma kind::i8 (int32 accumulators in TMEM),
// recombined in FP64, and compared with a long-double host reference and with plain FP64 accumulation.
//   K[c][.] = 2^ek[c] * sum_s a_s[c][.] 2^(-7(s+1)),   Z[i][.] = 2^ez[i] * sum_t b_t[i][.] 2^(-7(t+1)),
//   V[c][i] = 2^(ek[c]+ez[i]) * sum_{s+t<S} 2^(-7(s+t+2)) * (a_s[c] . b_t[i])      every dot product an exact int32.
// One CTA, operands staged with plain loads (this measures accuracy, not speed).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ozaki_probe tools/ozaki_probe.cu
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int SMAX = 10;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
constexpr uint32_t IDESC_I8 = (2u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ void umma_i8(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      :: "r"(tmem_c), "l"(da), "l"(db), "r"(IDESC_I8), "r"(accumulate), "r"(0u) : "memory");
}

// Row-wise slicing: exponent e with max|x| * 2^-e in [1/4, 1/2], then S signed digits of 7 bits.
// planes[s][row][col]; every operation below is exact in FP64.
__global__ void slice_rows(const double* __restrict__ X, int rows, int cols, int S, int8_t* __restrict__ planes,
                           int* __restrict__ expo) {
  const int r = blockIdx.x;
  __shared__ double red[128];
  double m = 0.0;
  for (int j = threadIdx.x; j < cols; j += blockDim.x) m = fmax(m, fabs(X[static_cast<size_t>(r) * cols + j]));
  red[threadIdx.x] = m;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + o]); __syncthreads(); }
  m = red[0];
  int e = 0;
  if (m > 0.0) { frexp(m, &e); e += 1; }          // m = f * 2^(e-1), f in [1/2, 1)  ->  m * 2^-e in [1/4, 1/2)
  if (threadIdx.x == 0) expo[r] = e;
  for (int j = threadIdx.x; j < cols; j += blockDim.x) {
    double x = ldexp(X[static_cast<size_t>(r) * cols + j], -e);
    for (int s = 0; s < S; ++s) {
      x *= 128.0;
      const double a = rint(x);
      planes[(static_cast<size_t>(s) * rows + r) * cols + j] = static_cast<int8_t>(a);
      x -= a;
    }
  }
}

// V tile (128 x 128) from the planes.  128 threads; thread t owns accumulator row t (TMEM lane t).
__global__ void __launch_bounds__(128, 1) sliced_product(const int8_t* __restrict__ Ap, const int* __restrict__ ea,
                                                         const int8_t* __restrict__ Bp, const int* __restrict__ eb,
                                                         int n, int S, double* __restrict__ V) {
  __shared__ __align__(1024) uint8_t sA[128 * 128];
  __shared__ __align__(1024) uint8_t sB[128 * 128];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
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
  double acc[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) acc[i] = 0.0;
  uint32_t phase = 0;
  for (int cls = S - 1; cls >= 0; --cls) {            // smallest contributions first
    bool first = true;
    for (int s = 0; s <= cls; ++s) {
      const int t = cls - s;
      const int8_t* A = Ap + static_cast<size_t>(s) * 128 * n;
      const int8_t* B = Bp + static_cast<size_t>(t) * 128 * n;
      for (int k0 = 0; k0 < n; k0 += 128) {
        __syncthreads();                                // previous chunk's instructions have been waited for below
        for (int idx = tid; idx < 128 * 8; idx += 128) {
          const int r = idx >> 3, c = idx & 7;
          *reinterpret_cast<int4*>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const int4*>(A + static_cast<size_t>(r) * n + k0 + c * 16);
          *reinterpret_cast<int4*>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const int4*>(B + static_cast<size_t>(r) * n + k0 + c * 16);
        }
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        __syncthreads();
        if (tid == 0) {
          asm volatile("tcgen05.fence::after_thread_sync;\n");
          const uint64_t da = umma_desc(smem_u32(sA)), db = umma_desc(smem_u32(sB));
          for (int k = 0; k < 4; ++k) umma_i8(tmem, da + 2 * k, db + 2 * k, (first && k == 0) ? 0u : 1u);
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];\n" :: "l"(__cvta_generic_to_shared(&bar)) : "memory");
        }
        first = false;
        {
          const uint32_t b = smem_u32(&bar);
          asm volatile(
              "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n"
              :: "r"(b), "r"(phase) : "memory");
        }
        phase ^= 1;
      }
    }
    asm volatile("tcgen05.fence::after_thread_sync;\n");
    const double w = ldexp(1.0, -7 * (cls + 2));
    for (int c0 = 0; c0 < 128; c0 += 16) {
      uint32_t v[16];
      const uint32_t taddr = tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
            "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[c0 + i] = fma(static_cast<double>(static_cast<int32_t>(v[i])), w, acc[c0 + i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n");
  }
  const int row = warp * 32 + lane;
  const int er = ea[row];
  for (int i = 0; i < 128; ++i) V[static_cast<size_t>(row) * 128 + i] = ldexp(acc[i], er + eb[i]);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(tmem), "n"(128));
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 2048;          // contraction length (observations), multiple of 128
  srand(11);
  auto uni = [] { return (rand() + 0.5) / (RAND_MAX + 1.0); };
  // K: kernel values in (0, 1] spread over many decades; Z: signed, rows of very different magnitude (1e-2 .. 1e5),
  // entries inside a row spread over ~6 decades -- the shape of L^-1 for an ill-conditioned kernel matrix
  std::vector<double> K(128 * static_cast<size_t>(n)), Z(128 * static_cast<size_t>(n));
  for (int c = 0; c < 128; ++c)
    for (int j = 0; j < n; ++j) K[static_cast<size_t>(c) * n + j] = exp(-30.0 * uni() * uni());
  for (int i = 0; i < 128; ++i) {
    const double row_scale = pow(10.0, -2.0 + 7.0 * i / 127.0);
    for (int j = 0; j < n; ++j) Z[static_cast<size_t>(i) * n + j] = row_scale * (2.0 * uni() - 1.0) * pow(10.0, -6.0 * uni());
  }
  double *dK, *dZ, *dV; int8_t *pK, *pZ; int *eK, *eZ;
  CK(cudaMalloc(&dK, K.size() * 8)); CK(cudaMalloc(&dZ, Z.size() * 8)); CK(cudaMalloc(&dV, 128 * 128 * 8));
  CK(cudaMalloc(&pK, static_cast<size_t>(SMAX) * 128 * n)); CK(cudaMalloc(&pZ, static_cast<size_t>(SMAX) * 128 * n));
  CK(cudaMalloc(&eK, 128 * 4)); CK(cudaMalloc(&eZ, 128 * 4));
  CK(cudaMemcpy(dK, K.data(), K.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dZ, Z.data(), Z.size() * 8, cudaMemcpyHostToDevice));
  // references: long double (taken as exact) and plain FP64 accumulation in index order (what DMMA/DFMA deliver)
  std::vector<long double> ref(128 * 128);
  std::vector<double> f64(128 * 128), absum(128 * 128);
  for (int c = 0; c < 128; ++c)
    for (int i = 0; i < 128; ++i) {
      long double s = 0; double d = 0, a = 0;
      for (int j = 0; j < n; ++j) {
        const double k = K[static_cast<size_t>(c) * n + j], z = Z[static_cast<size_t>(i) * n + j];
        s += static_cast<long double>(k) * z; d = fma(k, z, d); a += fabs(k * z);
      }
      ref[c * 128 + i] = s; f64[c * 128 + i] = d; absum[c * 128 + i] = a;
    }
  std::vector<double> V(128 * 128);
  for (int S = 6; S <= SMAX; ++S) {
    slice_rows<<<128, 128>>>(dK, 128, n, S, pK, eK);
    slice_rows<<<128, 128>>>(dZ, 128, n, S, pZ, eZ);
    sliced_product<<<1, 128>>>(pK, eK, pZ, eZ, n, S, dV);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(V.data(), dV, V.size() * 8, cudaMemcpyDeviceToHost));
    std::vector<int> hz(128), hk(128);
    CK(cudaMemcpy(hz.data(), eZ, 128 * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hk.data(), eK, 128 * 4, cudaMemcpyDeviceToHost));
    // errors relative to sum_j |k z| (the scale FP64 accumulation is accurate to) and to the a-priori bound
    double worst_sliced = 0, worst_f64 = 0, worst_vs_bound = 0;
    for (int c = 0; c < 128; ++c)
      for (int i = 0; i < 128; ++i) {
        const int q = c * 128 + i;
        const double es = fabs(static_cast<double>(V[q] - ref[q])), ef = fabs(static_cast<double>(f64[q] - ref[q]));
        worst_sliced = fmax(worst_sliced, es / absum[q]);
        worst_f64 = fmax(worst_f64, ef / absum[q]);
        const double bound = n * (S + 1.0) * ldexp(1.0, -7 * S - 2) * ldexp(1.0, hk[c] + hz[i]);
        worst_vs_bound = fmax(worst_vs_bound, es / bound);
      }
    printf("{\"n\": %d, \"slices\": %d, \"pairs\": %d, \"max_err_over_sum_abs_sliced\": %.3e, \"max_err_over_sum_abs_fp64\": %.3e, "
           "\"max_err_over_apriori_bound\": %.3f}\n", n, S, S * (S + 1) / 2, worst_sliced, worst_f64, worst_vs_bound);
  }
  return 0;
}
