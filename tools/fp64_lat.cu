// Dependent-issue latencies that bound the serial column steps of the diagonal-block factorisation:
// DFMA chain, DMUL chain, rsqrt.approx.f64, shared-memory round trip, __syncthreads with 8 warps, shuffle of a double.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_lat fp64_lat.cu && ./fp64_lat
#include <cstdio>
#include <cuda_runtime.h>

__global__ void lat_kernel(double* out, long long* clk, double seed) {
  __shared__ double sh[256];
  const int tid = threadIdx.x;
  double a = seed + tid * 1e-9, b = 1.0000001, c = 1e-9;
  long long t0, t1;
  sh[tid] = a;
  __syncthreads();
  // 1. dependent DFMA
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; ++i) a = fma(a, b, c);
  t1 = clock64();
  if (tid == 0) clk[0] = t1 - t0;
  // 2. dependent DMUL
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; ++i) a = a * b;
  t1 = clock64();
  if (tid == 0) clk[1] = t1 - t0;
  // 3. dependent rsqrt.approx
  double y = fabs(a) + 1.0;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; ++i) { asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(y)); }
  t1 = clock64();
  if (tid == 0) clk[2] = t1 - t0;
  // 4. shared-memory round trip (store, load dependent)
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; ++i) {
    volatile double* p = sh;
    p[tid] = y;
    y = p[(tid + 1) & 255] + 1.0;
  }
  t1 = clock64();
  if (tid == 0) clk[3] = t1 - t0;
  // 5. __syncthreads
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; ++i) __syncthreads();
  t1 = clock64();
  if (tid == 0) clk[4] = t1 - t0;
  // 6. shuffle of a double, dependent
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; ++i) y = __shfl_sync(0xffffffffu, y, (tid + 1) & 31) + 1.0;
  t1 = clock64();
  if (tid == 0) clk[5] = t1 - t0;
  out[tid] = a + y;
}

int main() {
  double* out; long long* clk;
  cudaMalloc(&out, 256 * sizeof(double));
  cudaMalloc(&clk, 8 * sizeof(long long));
  for (int threads : {32, 256}) {
    lat_kernel<<<1, threads>>>(out, clk, 1.0);
    lat_kernel<<<1, threads>>>(out, clk, 1.0);
    long long h[8];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%d threads: DFMA %.1f clk, DMUL %.1f, rsqrt.approx.f64 %.1f, smem store+load(+DADD) %.1f, __syncthreads %.1f, shfl double(+DADD) %.1f\n",
           threads, h[0] / 256.0, h[1] / 256.0, h[2] / 64.0, h[3] / 64.0, h[4] / 64.0, h[5] / 64.0);
  }
  return 0;
}
