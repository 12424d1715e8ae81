// tfs_reduce_axis0_noalloc -- a block reduction that allocates nothing.
//
// Who needs it.  Some hosts of this library are not allowed to allocate device memory at the
// moment they call it: a Spark executor whose GPU memory has been handed, whole, to a pool
// allocator (any cudaMalloc issued behind the pool's back then fails or fragments the device);
// a task thread that must not trigger the device-wide synchronisation cudaMalloc/cudaFree imply
// while other tasks' kernels are in flight; a caller running with a fixed, audited memory
// budget.  The regular entry points (tfs_session_run) grow a workspace with cudaMalloc on first
// use.  This one does not: the caller passes the input block and the output cell as device
// pointers it already owns, and the only scratch the kernel needs -- per-CTA partial results
// and an arrival counter -- lives in a fixed-size array that is part of the library image.
//
// What that implies.
//   * The array (kScratchBytes, a little over half a MiB) is resident on every GPU that loads
//     the library, from load until the process exits, whether or not this entry point is used.
//   * Its size bounds the problem: cells wider than kMaxCell elements are refused with
//     TFS_ERR_UNSUPPORTED (use tfs_session_run for those).  The row count is not limited.
//   * All calls on one device share the array, and it keeps whatever the last call left in
//     it.  The arrival counter is therefore cleared on the stream before every launch; the
//     partials need no clearing because each launch overwrites the slots it reads.  Two calls
//     whose kernels overlap on the same GPU (different streams, or different threads) would
//     use the same slots at once and corrupt each other: serialise them.
//   * The array's device address is looked up on the first call made for a device (one
//     cudaGetSymbolAddress with that device current; in host code the array's name is only a
//     handle, not an address a kernel could use).  Make that first call during start-up,
//     ahead of any stream capture.
//   * No cudaMalloc, cudaFree, cudaMallocAsync or allocating library routine is called.
//     The work is enqueued on the given stream and the call returns without synchronising.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "common.h"
#include "graph.h"
#include "kernels.h"

namespace tfs {
void set_last_error(const std::string& msg);   // capi.cpp
}

namespace {

constexpr int kThreads = 256;
constexpr int kJ = 4;
constexpr int kMaxCell = 256;                       // widest cell accepted, in elements
constexpr size_t kHeaderBytes = 256;                // arrival counter, padded
constexpr size_t kPartialBytes = 512 * 1024;        // per-CTA partial results
constexpr size_t kScratchBytes = kHeaderBytes + kPartialBytes;
constexpr int kMaxDevices = 64;

// The scratch itself: one copy in the global memory of every device that loads this library.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

template <class T> struct Lim;
template <> struct Lim<double> { static __device__ __forceinline__ double hi() { return __longlong_as_double(0x7ff0000000000000LL); } static __device__ __forceinline__ double lo() { return -hi(); } };
template <> struct Lim<float> { static __device__ __forceinline__ float hi() { return __int_as_float(0x7f800000); } static __device__ __forceinline__ float lo() { return -hi(); } };
template <> struct Lim<int> { static __device__ __forceinline__ int hi() { return 0x7fffffff; } static __device__ __forceinline__ int lo() { return int(0x80000000u); } };
template <> struct Lim<long long> { static __device__ __forceinline__ long long hi() { return 0x7fffffffffffffffLL; } static __device__ __forceinline__ long long lo() { return (long long)0x8000000000000000ULL; } };

template <class T> __device__ __forceinline__ T add_rn(T a, T b);
template <> __device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
template <> __device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ int add_rn(int a, int b) { return int((unsigned)a + (unsigned)b); }
template <> __device__ __forceinline__ long long add_rn(long long a, long long b) { return (long long)((unsigned long long)a + (unsigned long long)b); }

template <class T, int OP> struct R {
  static __device__ __forceinline__ T identity() { return OP == tfs::R_SUM ? T(0) : (OP == tfs::R_MIN ? Lim<T>::hi() : Lim<T>::lo()); }
  static __device__ __forceinline__ T op(T a, T b) {
    if (OP == tfs::R_SUM) return add_rn<T>(a, b);
    if (OP == tfs::R_MIN) return (b < a) ? b : a;
    return (a < b) ? b : a;
  }
};

template <class T, int P> struct VecOf;
template <> struct VecOf<double, 2> { using type = double2; };
template <> struct VecOf<long long, 2> { using type = longlong2; };
template <> struct VecOf<float, 4> { using type = float4; };
template <> struct VecOf<int, 4> { using type = int4; };

template <class T, int P>
__device__ __forceinline__ void load_p(const T* p, T (&v)[P]) {
  if constexpr (P == 1) {
    v[0] = __ldcs(p);
  } else {
    using V = typename VecOf<T, P>::type;
    const V t = __ldcs(reinterpret_cast<const V*>(p));
    const T* tv = reinterpret_cast<const T*>(&t);
#pragma unroll
    for (int e = 0; e < P; ++e) v[e] = tv[e];
  }
}

// in[n] = a row-major [outer, inner] block.  L = lcm(inner, P), q = L / P, blockDim.x % q == 0, so a
// thread's lane e always holds column ((threadIdx.x % q) * P + e) % inner.  `scratch` is the device
// address of g_static_scratch: [0, 256) arrival counter (zero on entry), then gridDim.x * L partials.
template <class T, int OP, int P>
__global__ void __launch_bounds__(kThreads)
reduce_axis0_static_scratch_kernel(const T* __restrict__ in, const int64_t n, const int inner, const int L, const int q,
                                   unsigned char* __restrict__ scratch, T* __restrict__ out) {
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
  T* partials = reinterpret_cast<T*>(scratch + kHeaderBytes);
  __shared__ T sm[kThreads * P];
  __shared__ bool is_last;
  const int b = blockDim.x;
  const int64_t n_vec = n / P;
  const int64_t stride = int64_t(gridDim.x) * b;
  T acc[kJ][P];
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) acc[j][e] = R<T, OP>::identity();
  int64_t v = int64_t(blockIdx.x) * b + threadIdx.x;
  for (; v + (kJ - 1) * stride < n_vec; v += kJ * stride) {
    T x[kJ][P];
#pragma unroll
    for (int j = 0; j < kJ; ++j) load_p<T, P>(in + (v + j * stride) * P, x[j]);
#pragma unroll
    for (int j = 0; j < kJ; ++j)
#pragma unroll
      for (int e = 0; e < P; ++e) acc[j][e] = R<T, OP>::op(acc[j][e], x[j][e]);
  }
  for (; v < n_vec; v += stride) {
    T x[P];
    load_p<T, P>(in + v * P, x);
#pragma unroll
    for (int e = 0; e < P; ++e) acc[0][e] = R<T, OP>::op(acc[0][e], x[e]);
  }
#pragma unroll
  for (int e = 0; e < P; ++e)
    sm[threadIdx.x * P + e] = R<T, OP>::op(R<T, OP>::op(acc[0][e], acc[1][e]), R<T, OP>::op(acc[2][e], acc[3][e]));
  __syncthreads();
  if (threadIdx.x < L) {   // fold the threads of each column class, fixed order
    const int tc = threadIdx.x / P, e = threadIdx.x % P;
    T r = R<T, OP>::identity();
    for (int t = tc; t < b; t += q) r = R<T, OP>::op(r, sm[t * P + e]);
    partials[int64_t(blockIdx.x) * L + threadIdx.x] = r;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // the last CTA to arrive folds every CTA's partials; slot f belongs to column f % inner
  const int bl = (b / L) * L;
  T r = R<T, OP>::identity();
  if (threadIdx.x < bl) {
    const int64_t total = int64_t(gridDim.x) * L;
    for (int64_t i = threadIdx.x; i < total; i += bl) r = R<T, OP>::op(r, __ldcg(partials + i));
  }
  __syncthreads();
  sm[threadIdx.x] = r;
  __syncthreads();
  if (threadIdx.x < inner) {
    T s = R<T, OP>::identity();
    for (int t = threadIdx.x; t < bl; t += inner) s = R<T, OP>::op(s, sm[t]);   // bl % inner == 0
    for (int64_t i = n_vec * P; i < n; ++i)                                       // elements past the last full vector
      if (int(i % inner) == int(threadIdx.x)) s = R<T, OP>::op(s, in[i]);
    out[threadIdx.x] = s;
  }
}

int64_t gcd64(int64_t a, int64_t b) { while (b) { int64_t t = a % b; a = b; b = t; } return a; }

struct DeviceSlot { bool ready = false; unsigned char* scratch = nullptr; int sm_count = 0; };
DeviceSlot g_slots[kMaxDevices];
std::mutex g_slots_mutex;

// First use on a device: resolve the device address of the static array (the symbol's name is a
// host-side handle; only cudaGetSymbolAddress, called with that device current, yields a pointer a
// kernel can dereference).
const DeviceSlot& device_slot(int device) {
  std::lock_guard<std::mutex> lock(g_slots_mutex);
  DeviceSlot& slot = g_slots[device];
  if (!slot.ready) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      tfs::fail(TFS_ERR_CUDA, std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e));
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || sms <= 0)
      tfs::fail(TFS_ERR_CUDA, std::string("cudaDeviceGetAttribute(SM count): ") + cudaGetErrorString(e));
    slot.scratch = static_cast<unsigned char*>(ptr);
    slot.sm_count = sms;
    slot.ready = true;
  }
  return slot;
}

template <class T, int OP>
void launch_t(const T* in, int64_t outer, int64_t inner, T* out, const DeviceSlot& slot, cudaStream_t stream) {
  const int64_t n = outer * inner;
  constexpr int PV = 16 / int(sizeof(T));
  int P = (reinterpret_cast<uintptr_t>(in) & 15) ? 1 : PV;
  int64_t L = inner / gcd64(inner, P) * P;
  if (L > kThreads) { P = 1; L = inner; }          // keeps L <= 256 for every cell up to kMaxCell
  const int q = int(L / P);
  const int threads = (kThreads / q) * q;
  const int64_t n_vec = n / P;
  int64_t grid = (n_vec + int64_t(threads) * kJ - 1) / (int64_t(threads) * kJ);
  grid = std::min<int64_t>(grid, int64_t(slot.sm_count) * 4);
  grid = std::min<int64_t>(grid, int64_t(kPartialBytes / (size_t(L) * sizeof(T))));   // partials must fit the array
  if (grid < 1) grid = 1;
  // the counter must start at zero whatever an earlier (possibly failed) call left behind
  cudaError_t e = cudaMemsetAsync(slot.scratch, 0, kHeaderBytes, stream);
  if (e != cudaSuccess) tfs::fail(TFS_ERR_CUDA, std::string("cudaMemsetAsync: ") + cudaGetErrorString(e));
  if (P == 1)
    reduce_axis0_static_scratch_kernel<T, OP, 1><<<(unsigned)grid, threads, 0, stream>>>(in, n, int(inner), int(L), q, slot.scratch, out);
  else
    reduce_axis0_static_scratch_kernel<T, OP, PV><<<(unsigned)grid, threads, 0, stream>>>(in, n, int(inner), int(L), q, slot.scratch, out);
  e = cudaGetLastError();
  if (e != cudaSuccess) tfs::fail(TFS_ERR_CUDA, std::string("kernel launch: ") + cudaGetErrorString(e));
}

template <class T>
void launch_op(int red_op, const void* in, int64_t outer, int64_t inner, void* out, const DeviceSlot& slot, cudaStream_t stream) {
  const T* i = static_cast<const T*>(in);
  T* o = static_cast<T*>(out);
  switch (red_op) {
    case tfs::R_SUM: launch_t<T, tfs::R_SUM>(i, outer, inner, o, slot, stream); break;
    case tfs::R_MIN: launch_t<T, tfs::R_MIN>(i, outer, inner, o, slot, stream); break;
    default: launch_t<T, tfs::R_MAX>(i, outer, inner, o, slot, stream); break;
  }
}

}  // namespace

extern "C" int tfs_reduce_axis0_noalloc(int device, void* cuda_stream, int32_t dtype, int32_t red_op, const void* in,
                                        int64_t outer, int64_t inner, void* out) {
  try {
    TFS_CHECK(device >= 0 && device < kMaxDevices, TFS_ERR_INVALID, "device ordinal out of range");
    TFS_CHECK(tfs::dtype_supported(dtype), TFS_ERR_UNSUPPORTED, "dtype must be float, double, int32 or int64");
    TFS_CHECK(red_op == TFS_REDUCE_SUM || red_op == TFS_REDUCE_MIN || red_op == TFS_REDUCE_MAX, TFS_ERR_INVALID,
              "red_op must be TFS_REDUCE_SUM, TFS_REDUCE_MIN or TFS_REDUCE_MAX");
    TFS_CHECK(outer >= 0 && inner >= 1, TFS_ERR_INVALID, "the block must be [outer >= 0, inner >= 1]");
    TFS_CHECK(inner <= kMaxCell, TFS_ERR_UNSUPPORTED,
              "cell of " + std::to_string(inner) + " elements exceeds the " + std::to_string(kMaxCell) +
                  "-element limit of the allocation-free entry point (its scratch is a fixed " +
                  std::to_string(kScratchBytes) + "-byte array); use tfs_session_run for wider cells");
    TFS_CHECK(out != nullptr && (in != nullptr || outer == 0), TFS_ERR_INVALID, "null device pointer");
    int previous = -1;
    cudaError_t e = cudaGetDevice(&previous);
    if (e != cudaSuccess) tfs::fail(TFS_ERR_CUDA, std::string("no CUDA device is available: ") + cudaGetErrorString(e));
    if (previous != device) {
      e = cudaSetDevice(device);
      if (e != cudaSuccess) tfs::fail(TFS_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    }
    struct Restore { int prev, cur; ~Restore() { if (prev != cur && prev >= 0) cudaSetDevice(prev); } } restore{previous, device};
    // input and output must already be memory of this device (a lookup, not an allocation)
    if (outer > 0)
      TFS_CHECK(tfs::on_device_or_refuse(in, device, "`in`"), TFS_ERR_INVALID, "`in` must be a device pointer");
    TFS_CHECK(tfs::on_device_or_refuse(out, device, "`out`"), TFS_ERR_INVALID, "`out` must be a device pointer");
    const DeviceSlot& slot = device_slot(device);
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const int op = red_op == TFS_REDUCE_SUM ? tfs::R_SUM : (red_op == TFS_REDUCE_MIN ? tfs::R_MIN : tfs::R_MAX);
    switch (dtype) {
      case TFS_DOUBLE: launch_op<double>(op, in, outer, inner, out, slot, stream); break;
      case TFS_FLOAT: launch_op<float>(op, in, outer, inner, out, slot, stream); break;
      case TFS_INT32: launch_op<int>(op, in, outer, inner, out, slot, stream); break;
      default: launch_op<long long>(op, in, outer, inner, out, slot, stream); break;
    }
    tfs::set_last_error("");
    return TFS_OK;
  } catch (const tfs::Error& err) {
    tfs::set_last_error(err.what());
    return err.code;
  } catch (const std::exception& err) {
    tfs::set_last_error(std::string("unexpected: ") + err.what());
    return TFS_ERR_RUNTIME;
  }
}
