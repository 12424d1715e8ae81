// K4 -- MatMul for map graphs: out[M, N] = op(a)[M, K] * op(b)[K, N] (+ bias[N]), float32 or float64.
//
// This file holds the portable SIMT kernel: FMA accumulation in k order at the operand precision, any shape,
// transposes handled by index arithmetic.  It serves odd shapes and is the correctness anchor for
// the tensor-core path (kernels_matmul_tc.cu), which takes over for the DataFrame-shaped case
// [rows, K] x [K, N] with a constant right-hand side.
// Reference: the TF MatMul kernel reached from Session.run (impl/DebugRowOps.scala:794); the
// reference's own use is the k-means demo (src/main/python/tensorframes_snippets/kmeans_demo.py:33).
#include "kernels.h"

namespace tfs {
namespace {

constexpr int kTile = 32;

// element (r, c) of op(x) where x is stored row-major [rows, cols] (before the optional transpose)
template <class T>
__device__ __forceinline__ T at(const T* __restrict__ x, int64_t ld, bool trans, int64_t r, int64_t c) {
  return trans ? x[c * ld + r] : x[r * ld + c];
}

// rows on grid.x (up to 2^31 - 1 tiles), columns on grid.y
template <class T>
__global__ void __launch_bounds__(kTile * kTile)
matmul_simt_kernel(const T* __restrict__ a, const T* __restrict__ b, const T* __restrict__ bias, T* __restrict__ out,
                   const int64_t M, const int64_t N, const int64_t K, const int64_t lda, const int64_t ldb, const bool ta,
                   const bool tb) {
  __shared__ T sa[kTile][kTile + 1];
  __shared__ T sb[kTile][kTile + 1];
  const int tx = threadIdx.x % kTile, ty = threadIdx.x / kTile;
  const int64_t row = int64_t(blockIdx.x) * kTile + ty;
  const int64_t col = int64_t(blockIdx.y) * kTile + tx;
  T acc = T(0);
  for (int64_t k0 = 0; k0 < K; k0 += kTile) {
    sa[ty][tx] = (row < M && k0 + tx < K) ? at(a, lda, ta, row, k0 + tx) : T(0);
    sb[ty][tx] = (k0 + ty < K && col < N) ? at(b, ldb, tb, k0 + ty, col) : T(0);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kTile; ++k) acc = fma(sa[ty][k], sb[k][tx], acc);
    __syncthreads();
  }
  if (row < M && col < N) out[row * N + col] = bias ? acc + bias[col] : acc;
}

template <class T>
int launch_t(const void* a, const void* b, const void* bias, void* out, int64_t M, int64_t N, int64_t K, bool ta, bool tb,
             cudaStream_t stream) {
  const int64_t lda = ta ? M : K;   // a is stored [K, M] when transposed
  const int64_t ldb = tb ? K : N;   // b is stored [N, K] when transposed
  dim3 grid((unsigned)((M + kTile - 1) / kTile), (unsigned)((N + kTile - 1) / kTile));
  TFS_CHECK(grid.y <= 65535u, TFS_ERR_UNSUPPORTED, "MatMul: more than 2M output columns are not supported");
  matmul_simt_kernel<T><<<grid, kTile * kTile, 0, stream>>>(static_cast<const T*>(a), static_cast<const T*>(b),
                                                           static_cast<const T*>(bias), static_cast<T*>(out), M, N, K, lda, ldb, ta, tb);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace

int launch_matmul_simt(int dtype, const void* a, const void* b, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                       bool ta, bool tb, const DeviceInfo& dev, cudaStream_t stream) {
  (void)dev;
  if (M <= 0 || N <= 0) return 0;
  if (dtype == TFS_DOUBLE) return launch_t<double>(a, b, bias, out, M, N, K, ta, tb, stream);
  return launch_t<float>(a, b, bias, out, M, N, K, ta, tb, stream);
}

}  // namespace tfs
