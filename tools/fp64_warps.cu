// Probe: FP64 FMA throughput of one SM as a function of the number of resident warps (8 independent chains per thread).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_warps tools/fp64_warps.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void fma_chains(double* out, int iters) {
  double a[8];
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + 1e-3 * (threadIdx.x + i);
  const double m = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fma(a[i], m, c);
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += a[i];
  if (s == 12345.678) out[0] = s;
}
int main() {
  double* out; cudaMalloc(&out, 8);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps : {1, 2, 4, 8, 12, 16, 32}) {
    fma_chains<<<sms, 32 * warps>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); fma_chains<<<sms, 32 * warps>>>(out, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * 64 * iters * 32.0 * warps * sms;
    printf("{\"warps per SM\": %d, \"ms\": %.3f, \"tflops\": %.2f, \"fma per clk per SM at 1.965 GHz\": %.1f}\n", warps, ms, flops / (ms * 1e-3) / 1e12,
           flops / 2 / (ms * 1e-3) / sms / 1.965e9);
  }
  return 0;
}
