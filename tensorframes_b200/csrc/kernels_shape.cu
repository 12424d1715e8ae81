// Shape-changing steps of map graphs (sm_100a): Tile, ArgMin / ArgMax, and the scatter that finishes
// UnsortedSegmentSum.  They serve the reference's k-means snippet
// (src/main/python/tensorframes_snippets/kmeans_demo.py:26-43, 76-79, 136-144), whose per-block graph is
// distances = tile(|c|^2) + tile(|x|^2) - 2 x c^T, argmin / min over the centroid axis, and a segment sum of the
// points by nearest centroid.  All are HBM-bound copies or row scans: coalesced along the fastest output
// dimension, persistent grids, no shared memory needed.
#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;

struct TileDims {
  int rank;
  int64_t out_d[TFS_MAX_RANK];
  int64_t in_d[TFS_MAX_RANK];
  int64_t in_stride[TFS_MAX_RANK];
};

// out[coords] = in[coords mod in_dims]
template <class T>
__global__ void __launch_bounds__(kThreads)
tile_kernel(const T* __restrict__ in, T* __restrict__ out, const int64_t n, const TileDims d) {
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    int64_t rem = i, src = 0;
#pragma unroll 1
    for (int k = d.rank - 1; k >= 0; --k) {
      const int64_t c = rem % d.out_d[k];
      rem /= d.out_d[k];
      src += (c % d.in_d[k]) * d.in_stride[k];
    }
    out[i] = in[src];
  }
}

// out[o, i] = index of the extreme of in[o, :, i]; ties go to the smallest index (TF / Eigen argmin order).
template <class T, class I, bool IS_MIN>
__global__ void __launch_bounds__(kThreads)
argminmax_kernel(const T* __restrict__ in, I* __restrict__ out, const int64_t outer, const int64_t len, const int64_t inner) {
  const int64_t n = outer * inner;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += stride) {
    const int64_t o = t / inner, i = t % inner;
    const T* p = in + o * len * inner + i;
    T best = p[0];
    int64_t arg = 0;
    for (int64_t a = 1; a < len; ++a) {
      const T v = p[a * inner];
      if (IS_MIN ? (v < best) : (best < v)) { best = v; arg = a; }
    }
    out[t] = I(arg);
  }
}

// inner == 1 and a short axis: one warp per 32 rows would read strided; instead each warp takes rows one at a time
// with its lanes across the axis (coalesced), then an arg-reduction by shuffles.
template <class T, class I, bool IS_MIN>
__global__ void __launch_bounds__(kThreads)
argminmax_rows_kernel(const T* __restrict__ in, I* __restrict__ out, const int64_t rows, const int64_t len) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += warps) {
    const T* p = in + r * len;
    bool have = false;
    T best = T(0);
    int64_t arg = 0;
    for (int64_t a = lane; a < len; a += 32) {
      const T v = p[a];
      if (!have || (IS_MIN ? (v < best) : (best < v))) { best = v; arg = a; have = true; }
    }
    for (int m = 16; m > 0; m >>= 1) {
      const T ob = __shfl_xor_sync(0xffffffffu, best, m);
      const int64_t oa = __shfl_xor_sync(0xffffffffu, arg, m);
      const int oh = __shfl_xor_sync(0xffffffffu, int(have), m);
      if (oh) {
        const bool better = IS_MIN ? (ob < best) : (best < ob);
        const bool tie = !(ob < best) && !(best < ob);
        if (!have || better || (tie && oa < arg)) { best = ob; arg = oa; have = true; }
      }
    }
    if (lane == 0) out[r] = I(arg);
  }
}

template <class K>
__global__ void __launch_bounds__(kThreads)
widen_ids_kernel(const K* __restrict__ in, long long* __restrict__ out, const int64_t n) {
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (long long)in[i];
}

// out[key[g], :] = sums[g, :]; a key outside [0, segments) raises the error flag (code 2)
template <class T>
__global__ void __launch_bounds__(kThreads)
segment_scatter_kernel(const T* __restrict__ sums, const long long* __restrict__ keys, const int64_t groups, const int64_t cell,
                       const int64_t segments, T* __restrict__ out, int* __restrict__ err) {
  const int64_t n = groups * cell;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t g = i / cell, c = i % cell;
    const long long k = keys[g];
    if (k < 0 || k >= segments) { if (c == 0) atomicMax(err, 2); continue; }
    out[k * cell + c] = sums[i];
  }
}

int64_t grid_for(int64_t n, const DeviceInfo& dev) {
  int64_t g = (n + kThreads - 1) / kThreads;
  g = std::min<int64_t>(g, int64_t(dev.sm_count) * 8);
  return g < 1 ? 1 : g;
}

template <class T, class I>
int launch_arg_t(bool is_min, const T* in, int64_t outer, int64_t len, int64_t inner, I* out, const DeviceInfo& dev, cudaStream_t stream) {
  const int64_t n = outer * inner;
  if (n <= 0) return 0;
  if (inner == 1 && len >= 8) {
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((outer * 32 + kThreads - 1) / kThreads, int64_t(dev.sm_count) * 8));
    if (is_min) argminmax_rows_kernel<T, I, true><<<grid, kThreads, 0, stream>>>(in, out, outer, len);
    else argminmax_rows_kernel<T, I, false><<<grid, kThreads, 0, stream>>>(in, out, outer, len);
  } else {
    const unsigned grid = (unsigned)grid_for(n, dev);
    if (is_min) argminmax_kernel<T, I, true><<<grid, kThreads, 0, stream>>>(in, out, outer, len, inner);
    else argminmax_kernel<T, I, false><<<grid, kThreads, 0, stream>>>(in, out, outer, len, inner);
  }
  TFS_CUDA(cudaGetLastError());
  return 1;
}

template <class T>
int launch_arg_i(bool is_min, int out_dtype, const void* in, int64_t outer, int64_t len, int64_t inner, void* out,
                 const DeviceInfo& dev, cudaStream_t stream) {
  if (out_dtype == TFS_INT32)
    return launch_arg_t<T, int>(is_min, static_cast<const T*>(in), outer, len, inner, static_cast<int*>(out), dev, stream);
  return launch_arg_t<T, long long>(is_min, static_cast<const T*>(in), outer, len, inner, static_cast<long long*>(out), dev, stream);
}

}  // namespace

int launch_tile(int elem_size, const void* in, const Shape& in_shape, void* out, const Shape& out_shape, const DeviceInfo& dev,
                cudaStream_t stream) {
  const int64_t n = out_shape.numel();
  if (n <= 0) return 0;
  TileDims d{};
  d.rank = out_shape.rank;
  int64_t st = 1;
  for (int k = in_shape.rank - 1; k >= 0; --k) {
    d.out_d[k] = out_shape.d[k];
    d.in_d[k] = in_shape.d[k];
    d.in_stride[k] = st;
    st *= in_shape.d[k];
  }
  const unsigned grid = (unsigned)grid_for(n, dev);
  if (elem_size == 8)
    tile_kernel<unsigned long long><<<grid, kThreads, 0, stream>>>(static_cast<const unsigned long long*>(in),
                                                                   static_cast<unsigned long long*>(out), n, d);
  else
    tile_kernel<unsigned int><<<grid, kThreads, 0, stream>>>(static_cast<const unsigned int*>(in), static_cast<unsigned int*>(out), n, d);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

int launch_argminmax(int dtype, int out_dtype, bool is_min, const void* in, int64_t outer, int64_t len, int64_t inner, void* out,
                     const DeviceInfo& dev, cudaStream_t stream) {
  TFS_CHECK(len >= 1, TFS_ERR_INVALID, "ArgMin / ArgMax over an empty axis");
  switch (dtype) {
    case TFS_FLOAT: return launch_arg_i<float>(is_min, out_dtype, in, outer, len, inner, out, dev, stream);
    case TFS_DOUBLE: return launch_arg_i<double>(is_min, out_dtype, in, outer, len, inner, out, dev, stream);
    case TFS_INT32: return launch_arg_i<int>(is_min, out_dtype, in, outer, len, inner, out, dev, stream);
    case TFS_INT64: return launch_arg_i<long long>(is_min, out_dtype, in, outer, len, inner, out, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "ArgMin / ArgMax: unsupported dtype");
}

int launch_widen_ids(const int* in, int64_t n, int64_t* out, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  widen_ids_kernel<int><<<(unsigned)grid_for(n, dev), kThreads, 0, stream>>>(in, reinterpret_cast<long long*>(out), n);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

int launch_segment_scatter(int dtype, const void* sums, const int64_t* keys, int64_t groups, int64_t cell, int64_t segments,
                           void* out, int32_t* err, const DeviceInfo& dev, cudaStream_t stream) {
  const int64_t n = groups * cell;
  if (n <= 0) return 0;
  const unsigned grid = (unsigned)grid_for(n, dev);
  const long long* k = reinterpret_cast<const long long*>(keys);
  switch (dtype) {
    case TFS_FLOAT: segment_scatter_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(sums), k, groups, cell, segments, static_cast<float*>(out), err); break;
    case TFS_DOUBLE: segment_scatter_kernel<double><<<grid, kThreads, 0, stream>>>(static_cast<const double*>(sums), k, groups, cell, segments, static_cast<double*>(out), err); break;
    case TFS_INT32: segment_scatter_kernel<int><<<grid, kThreads, 0, stream>>>(static_cast<const int*>(sums), k, groups, cell, segments, static_cast<int*>(out), err); break;
    case TFS_INT64: segment_scatter_kernel<long long><<<grid, kThreads, 0, stream>>>(static_cast<const long long*>(sums), k, groups, cell, segments, static_cast<long long*>(out), err); break;
    default: fail(TFS_ERR_UNSUPPORTED, "UnsortedSegmentSum: unsupported dtype");
  }
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace tfs
