// Growing the posterior state by one observation at a fixed length scale, in O(n^2).
//
// Reference context: every BO iteration of a group adds the finished experiments to the observation
// set and calls GaussianProcessRegressor.fit on all of them again (optimizer.py:21, sklearn
// _gpr.py:299-368).  With theta held, the factor of the bordered matrix
//     [ K   k ]        [ L    0      ]
//     [ k^T c ]   is   [ l^T  lambda ]     l = L^-1 k,  lambda^2 = c - l^T l,  c = 1 + jitter,
// and the inverse factor Z = L^-1 gains the row [ -(l^T Z) / lambda , 1 / lambda ].
// alpha = Z^T (Z y) is rebuilt from the extended Z (two triangular mat-vecs).
// The kernel column uses the same arithmetic as kmat_kernel (operands divided by the length scale,
// squared differences summed in dimension order, library sqrt / exp).
#pragma once
#include "gp_kernels.cuh"

namespace plx {

// kv[j] = k(x_j, x_new), j < n; x_new is row n of X
__global__ void append_kvec_kernel(const double* __restrict__ X, int n, int d, double ell, int kind, MaternGen mg,
                                   double* __restrict__ kv) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const double* xj = X + static_cast<size_t>(j) * d;
  const double* xn = X + static_cast<size_t>(n) * d;
  double r2 = 0.0;
  for (int q = 0; q < d; ++q) { const double df = xj[q] / ell - xn[q] / ell; r2 += df * df; }
  kv[j] = kern_value(kind, r2, mg);
}

// y[i] = sum_{j <= i} Z[i][j] x[j], i < n.  One warp per row, lanes stride the columns, fixed-order
// butterfly: the result does not depend on the launch shape.
__global__ void __launch_bounds__(256) trmv_lower_kernel(const double* __restrict__ Z, int ld, int n,
                                                         const double* __restrict__ x, double* __restrict__ y) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= n) return;
  const double* row = Z + static_cast<size_t>(i) * ld;
  double s = 0.0;
  for (int j = lane; j <= i; j += 32) s += row[j] * x[j];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) y[i] = s;
}

// y[j] = scale * sum_{j <= i < n} Z[i][j] x[i], j < n.  A block owns 32 columns; its 8 warps deal the
// rows among themselves and the 8 partial sums are added in warp order.
__global__ void __launch_bounds__(256) trmv_lower_t_kernel(const double* __restrict__ Z, int ld, int n,
                                                           const double* __restrict__ x, double scale,
                                                           double* __restrict__ y) {
  __shared__ double part[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j0 = blockIdx.x * 32, j = j0 + tx;
  double s = 0.0;
  if (j < n)
    for (int i = j0 + ty; i < n; i += 8)
      if (i >= j) s += Z[static_cast<size_t>(i) * ld + j] * x[i];
  part[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && j < n) {
    double t = part[0][tx];
#pragma unroll
    for (int w = 1; w < 8; ++w) t += part[w][tx];
    y[j] = scale * t;
  }
}

// out[0] = c - sum_j l[j]^2 (one block; strided partial sums, then a fixed-order tree)
__global__ void __launch_bounds__(256) append_pivot_kernel(const double* __restrict__ l, int n, double c,
                                                           double* __restrict__ out) {
  __shared__ double red[256];
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += 256) s += l[j] * l[j];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = c - red[0];
}

// row n of a lower-triangular factor: dst[n][0..n) = src (left alone when src is null), dst[n][n] = diag
__global__ void append_row_kernel(double* __restrict__ dst, int ld, int n, const double* __restrict__ src,
                                  double diag) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double* row = dst + static_cast<size_t>(n) * ld;
  if (j < n) { if (src) row[j] = src[j]; }
  else if (j == n) row[j] = diag;
}

}  // namespace plx
