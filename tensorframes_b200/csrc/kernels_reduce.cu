// K2 -- block reductions Sum / Min / Max (sm_100a).
//
// reduce_axis0: in[outer, inner] -> out[inner], the shape of
// `reduce_sum(x_input, [0])` over a DataFrame block (reference: the TF Sum/Min
// kernels reached from performReduceBlock, impl/DebugRowOps.scala:872-895).
//
// The block is streamed as one flat array with 128-bit coalesced loads.  The
// thread count is chosen so that the stride between a thread's successive
// vectors is a multiple of lcm(inner, P): every register lane then always sees
// the same column, so the hot loop is "load, accumulate" with no index math.
// Combination order is fixed (per-thread unrolled accumulators -> warp shuffle
// -> shared memory -> per-CTA partial in global scratch -> last CTA folds the
// partials), so results are run-to-run deterministic; floating-point sums differ
// from TF-CPU only by summation order.  HBM-bound: bytes read / HBM bandwidth.
#include <cstdlib>
#include <string>

#include "kernels.h"

namespace tfs {
namespace {

constexpr int kMaxThreads = 256;
constexpr int kJ = 4;            // independent vector loads in flight per thread
constexpr int kMaxQ = 64;        // widest thread-class count handled by the flat kernel
constexpr size_t kCounterBytes = 256;

template <class T> struct Limits;
template <> struct Limits<double> {
  static __device__ __forceinline__ double lowest() { return -__longlong_as_double(0x7ff0000000000000LL); }
  static __device__ __forceinline__ double highest() { return __longlong_as_double(0x7ff0000000000000LL); }
};
template <> struct Limits<float> {
  static __device__ __forceinline__ float lowest() { return -__int_as_float(0x7f800000); }
  static __device__ __forceinline__ float highest() { return __int_as_float(0x7f800000); }
};
template <> struct Limits<int> {
  static __device__ __forceinline__ int lowest() { return int(0x80000000u); }
  static __device__ __forceinline__ int highest() { return 0x7fffffff; }
};
template <> struct Limits<long long> {
  static __device__ __forceinline__ long long lowest() { return (long long)(0x8000000000000000ULL); }
  static __device__ __forceinline__ long long highest() { return 0x7fffffffffffffffLL; }
};

template <class T> __device__ __forceinline__ T wrap_add(T a, T b);
template <> __device__ __forceinline__ double wrap_add(double a, double b) { return __dadd_rn(a, b); }
template <> __device__ __forceinline__ float wrap_add(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ int wrap_add(int a, int b) { return int((unsigned)a + (unsigned)b); }
template <> __device__ __forceinline__ long long wrap_add(long long a, long long b) {
  return (long long)((unsigned long long)a + (unsigned long long)b);
}

template <class T, int OP> struct Red {
  static __device__ __forceinline__ T identity() {
    if (OP == R_SUM) return T(0);
    if (OP == R_MIN) return Limits<T>::highest();
    return Limits<T>::lowest();
  }
  static __device__ __forceinline__ T op(T a, T b) {
    if (OP == R_SUM) return wrap_add<T>(a, b);
    if (OP == R_MIN) return (b < a) ? b : a;
    return (a < b) ? b : a;
  }
};

template <class T, int P> struct Vec;
template <> struct Vec<double, 2> { using type = double2; };
template <> struct Vec<long long, 2> { using type = longlong2; };
template <> struct Vec<float, 4> { using type = float4; };
template <> struct Vec<int, 4> { using type = int4; };

template <class T, int P>
__device__ __forceinline__ void load_p(const T* p, T (&v)[P]) {
  if constexpr (P == 1) {
    v[0] = __ldcs(p);
  } else {
    using V = typename Vec<T, P>::type;
    V t = __ldcs(reinterpret_cast<const V*>(p));
    if constexpr (P == 2) { v[0] = t.x; v[1] = t.y; }
    else { v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
  }
}

template <class T>
__device__ __forceinline__ T shfl_xor(T v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

// CTA combine + grid combine shared by the LDG and the TMA variants.  `a` holds the thread's P
// lane accumulators; thread t of the CTA owns flat slots (t % q) * P + e of the L-periodic pattern.
template <class T, int OP, int P>
__device__ __forceinline__ void combine_and_finish(T (&a)[P], const T* __restrict__ in, const int64_t n,
                                                   const int64_t n_vec, const int inner, const int L, const int q,
                                                   T* __restrict__ partials, unsigned int* __restrict__ counter,
                                                   T* __restrict__ out, T* sm, T* sm2, bool* is_last) {
  using R = Red<T, OP>;
  const int b = blockDim.x;
  // lanes whose index differs by a multiple of q hold the same columns
  const bool shuffle_ok = (q <= 32) && ((q & (q - 1)) == 0);
  int contributors;  // threads per CTA whose `a` still matters, laid out class-major below
  if (shuffle_ok) {
    for (int m = 16; m >= q; m >>= 1) {
#pragma unroll
      for (int e = 0; e < P; ++e) a[e] = R::op(a[e], shfl_xor(a[e], m));
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < q) {
#pragma unroll
      for (int e = 0; e < P; ++e) sm[(warp * q + lane) * P + e] = a[e];
    }
    contributors = (b >> 5) * q;   // b is a multiple of 32 here (q | 32 and b = (256/q)*q)
  } else {
#pragma unroll
    for (int e = 0; e < P; ++e) sm[threadIdx.x * P + e] = a[e];
    contributors = b;
  }
  __syncthreads();
  if (threadIdx.x < L) {
    const int tc = threadIdx.x / P, e = threadIdx.x % P;
    T r = R::identity();
    for (int t = tc; t < contributors; t += q) r = R::op(r, sm[t * P + e]);
    partials[int64_t(blockIdx.x) * L + threadIdx.x] = r;
  }
  // ---- grid combine by the last CTA to finish
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int ticket = atomicAdd(counter, 1u);
    *is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!*is_last) return;
  __threadfence();
  const int bl = (b / L) * L;            // threads used, a multiple of L (L <= b by construction)
  T r = R::identity();
  if (threadIdx.x < bl) {
    const int64_t total = int64_t(gridDim.x) * L;
    for (int64_t i = threadIdx.x; i < total; i += bl) r = R::op(r, __ldcg(partials + i));
  }
  sm2[threadIdx.x] = r;
  __syncthreads();
  if (threadIdx.x < L) {
    T s = R::identity();
    for (int t = threadIdx.x; t < bl; t += L) s = R::op(s, sm2[t]);
    sm[threadIdx.x] = s;   // per flat-slot f totals
  }
  __syncthreads();
  if (threadIdx.x < inner) {
    T s = R::identity();
    for (int f = threadIdx.x; f < L; f += inner) s = R::op(s, sm[f]);
    // elements past the last full vector (n % P of them)
    for (int64_t i = n_vec * P; i < n; ++i)
      if (int(i % inner) == int(threadIdx.x)) s = R::op(s, in[i]);
    out[threadIdx.x] = s;
  }
  if (threadIdx.x == 0) *counter = 0;
}

// in: n = outer*inner elements. L = lcm(inner, P), q = L / P, blockDim.x % q == 0.
// partials: [gridDim.x][L]. counter: zero on entry, zero again on exit.
template <class T, int OP, int P>
__global__ void __launch_bounds__(kMaxThreads, 5)
reduce_axis0_flat_kernel(const T* __restrict__ in, const int64_t n, const int inner, const int L, const int q,
                         T* __restrict__ partials, unsigned int* __restrict__ counter, T* __restrict__ out) {
  using R = Red<T, OP>;
  __shared__ T sm[kMaxThreads * P];
  __shared__ T sm2[kMaxThreads];
  __shared__ bool is_last;

  const int b = blockDim.x;
  const int64_t n_vec = n / P;
  const int64_t stride = int64_t(gridDim.x) * b;   // in vectors; stride * P % L == 0
  int64_t v = int64_t(blockIdx.x) * b + threadIdx.x;

  T acc[kJ][P];
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) acc[j][e] = R::identity();

  for (; v + (kJ - 1) * stride < n_vec; v += kJ * stride) {
    T x[kJ][P];
#pragma unroll
    for (int j = 0; j < kJ; ++j) load_p<T, P>(in + (v + j * stride) * P, x[j]);
#pragma unroll
    for (int j = 0; j < kJ; ++j)
#pragma unroll
      for (int e = 0; e < P; ++e) acc[j][e] = R::op(acc[j][e], x[j][e]);
  }
  for (; v < n_vec; v += stride) {
    T x[P];
    load_p<T, P>(in + v * P, x);
#pragma unroll
    for (int e = 0; e < P; ++e) acc[0][e] = R::op(acc[0][e], x[e]);
  }
  T a[P];
#pragma unroll
  for (int e = 0; e < P; ++e) a[e] = R::op(R::op(acc[0][e], acc[1][e]), R::op(acc[2][e], acc[3][e]));
  combine_and_finish<T, OP, P>(a, in, n, n_vec, inner, L, q, partials, counter, out, sm, sm2, &is_last);
}

// ---- TMA-staged variant -------------------------------------------------------------------------
// One persistent CTA per SM.  An elected thread streams kStageBytes chunks of the block into a
// ring of shared-memory stages with cp.async.bulk (the TMA unit, no registers or LSU slots held
// while the bytes are in flight: kStages * kStageBytes per SM), completion is signalled on an
// mbarrier per stage, and all threads fold the landed stage with conflict-free 128-bit shared
// loads.  The stage size is a multiple of blockDim.x * P elements, so a thread's lanes keep the
// same columns in every stage, exactly like the LDG variant.
constexpr int kStages = 6;
constexpr int kVecPerThread = 8;   // 16-byte vectors per thread per stage: 256 * 8 * 16 B = 32 KB stages

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t"
        "}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
               "r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// Work is handed out in UNITS of `unit_chunks` consecutive stages, by ticket (one atomicAdd per unit): an SM that
// starts late or is shared with another kernel's CTAs -- the exchange of the previous block runs on a second stream
// while this kernel streams the next block -- simply takes fewer units, instead of holding the whole launch back by
// its fixed share.  Which CTA folds a unit changes from run to run; what is folded, and in which order, does not:
// every unit is reduced to its own partial (fixed order inside the CTA), partials are indexed by unit, and the last
// CTA folds them in index order.  So the result bits do not depend on the hand-out.
struct StageMeta { long long unit; int first_in_unit, last_in_unit; long long chunk; };

template <class T, int OP, int P>
__device__ __forceinline__ void cta_fold_to(T (&a)[P], const int L, const int q, T* __restrict__ dst, T* sm) {
  using R = Red<T, OP>;
  const int b = blockDim.x;
  const bool shuffle_ok = (q <= 32) && ((q & (q - 1)) == 0);
  int contributors;
  if (shuffle_ok) {
    for (int m = 16; m >= q; m >>= 1) {
#pragma unroll
      for (int e = 0; e < P; ++e) a[e] = R::op(a[e], shfl_xor(a[e], m));
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < q) {
#pragma unroll
      for (int e = 0; e < P; ++e) sm[(warp * q + lane) * P + e] = a[e];
    }
    contributors = (b >> 5) * q;
  } else {
#pragma unroll
    for (int e = 0; e < P; ++e) sm[threadIdx.x * P + e] = a[e];
    contributors = b;
  }
  __syncthreads();
  if (threadIdx.x < L) {
    const int tc = threadIdx.x / P, e = threadIdx.x % P;
    T r = R::identity();
    for (int t = tc; t < contributors; t += q) r = R::op(r, sm[t * P + e]);
    dst[threadIdx.x] = r;
  }
}

// header: header[0] = CTAs finished, header[1] = next unit ticket; both zero on entry and zero again on exit.
template <class T, int OP, int P>
__global__ void __launch_bounds__(kMaxThreads, 1)
reduce_axis0_tma_kernel(const T* __restrict__ in, const int64_t n, const int inner, const int L, const int q,
                        const int64_t n_units, const int unit_chunks, T* __restrict__ partials,
                        unsigned int* __restrict__ header, T* __restrict__ out) {
  using R = Red<T, OP>;
  extern __shared__ __align__(128) unsigned char dyn_smem[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ StageMeta meta[kStages];
  __shared__ T sm[kMaxThreads * P];
  __shared__ T sm2[kMaxThreads];
  __shared__ bool is_last;

  const int b = blockDim.x;
  const int64_t n_vec = n / P;
  const int64_t chunk_vecs = int64_t(b) * kVecPerThread;          // vectors per stage
  const uint32_t stage_bytes = uint32_t(chunk_vecs) * 16u;
  const int64_t n_chunks = (n_vec + chunk_vecs - 1) / chunk_vecs;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full_bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // producer state (thread 0 only): the unit being issued and the position inside it
  long long p_unit = -1;
  int p_pos = 0, p_len = 0;
  auto issue = [&](int64_t k) {   // thread 0 only: the k-th stage of this CTA's own sequence
    const int st = int(k % kStages);
    if (p_pos == p_len) {         // next unit, by ticket
      const long long t = (long long)atomicAdd(header + 1, 1u);
      if (t >= n_units) {
        meta[st].unit = -1;
        p_unit = -1;
        p_pos = p_len = 0;
        return;
      }
      p_unit = t;
      p_pos = 0;
      const int64_t c0 = t * unit_chunks;
      p_len = int((n_chunks - c0) < unit_chunks ? (n_chunks - c0) : unit_chunks);
    }
    const int64_t chunk = p_unit * unit_chunks + p_pos;
    const int64_t v0 = chunk * chunk_vecs;
    const int64_t vecs = (n_vec - v0) < chunk_vecs ? (n_vec - v0) : chunk_vecs;
    const uint32_t bytes = uint32_t(vecs) * 16u;
    meta[st].unit = p_unit;
    meta[st].chunk = chunk;
    meta[st].first_in_unit = p_pos == 0;
    meta[st].last_in_unit = p_pos == p_len - 1;
    ++p_pos;
    mbar_expect_tx(&full_bar[st], bytes);
    bulk_load(dyn_smem + size_t(st) * stage_bytes, in + v0 * P, bytes, &full_bar[st]);
  };
  if (threadIdx.x == 0) {
    bool open = true;
    for (int64_t k = 0; k < kStages - 1; ++k) {
      if (open) { issue(k); open = meta[k % kStages].unit >= 0; }
      else meta[k % kStages].unit = -1;
    }
  }

  T acc[2][P];
#pragma unroll
  for (int e = 0; e < P; ++e) acc[0][e] = acc[1][e] = R::identity();

  using V = typename Vec<T, P>::type;
  bool producer_open = true;     // thread 0: tickets may be left
  for (int64_t k = 0;; ++k) {
    __syncthreads();   // every thread is done with the stage the next load overwrites; thread 0's meta writes are visible
    if (threadIdx.x == 0) {
      const int st_new = int((k + kStages - 1) % kStages);
      if (k == 0) producer_open = meta[(kStages - 2) % kStages].unit >= 0;
      if (producer_open) { issue(k + kStages - 1); producer_open = meta[st_new].unit >= 0; }
      else meta[st_new].unit = -1;
    }
    const int st = int(k % kStages);
    const StageMeta m = meta[st];
    if (m.unit < 0) break;       // the same for every thread: the hand-out is over for this CTA
    mbar_wait(&full_bar[st], uint32_t((k / kStages) & 1));
    const int64_t left = n_vec - m.chunk * chunk_vecs;   // valid vectors in this stage
    const V* stage = reinterpret_cast<const V*>(dyn_smem + size_t(st) * stage_bytes);
    if (left >= chunk_vecs) {
#pragma unroll
      for (int j = 0; j < kVecPerThread; ++j) {
        const V t = stage[j * b + threadIdx.x];
        const T* tv = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int e = 0; e < P; ++e) acc[j & 1][e] = R::op(acc[j & 1][e], tv[e]);
      }
    } else {
      for (int j = 0; j < kVecPerThread; ++j) {
        if (int64_t(j) * b + threadIdx.x < left) {
          const V t = stage[j * b + threadIdx.x];
          const T* tv = reinterpret_cast<const T*>(&t);
#pragma unroll
          for (int e = 0; e < P; ++e) acc[j & 1][e] = R::op(acc[j & 1][e], tv[e]);
        }
      }
    }
    if (m.last_in_unit) {        // this unit's partial, then start afresh
      T a[P];
#pragma unroll
      for (int e = 0; e < P; ++e) a[e] = R::op(acc[0][e], acc[1][e]);
      cta_fold_to<T, OP, P>(a, L, q, partials + m.unit * L, sm);
#pragma unroll
      for (int e = 0; e < P; ++e) acc[0][e] = acc[1][e] = R::identity();
    }
  }
  // ---- grid combine by the last CTA to finish: partials in unit order
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int done = atomicAdd(header, 1u);
    is_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int bl = (b / L) * L;            // threads used, a multiple of L (L <= b by construction)
  T r = R::identity();
  if (threadIdx.x < bl) {
    const int64_t total = n_units * L;
    for (int64_t i = threadIdx.x; i < total; i += bl) r = R::op(r, __ldcg(partials + i));
  }
  sm2[threadIdx.x] = r;
  __syncthreads();
  if (threadIdx.x < L) {
    T s = R::identity();
    for (int t = threadIdx.x; t < bl; t += L) s = R::op(s, sm2[t]);
    sm[threadIdx.x] = s;   // per flat-slot totals
  }
  __syncthreads();
  if (threadIdx.x < inner) {
    T s = R::identity();
    for (int f = threadIdx.x; f < L; f += inner) s = R::op(s, sm[f]);
    for (int64_t i = n_vec * P; i < n; ++i)   // elements past the last full vector (n % P of them)
      if (int(i % inner) == int(threadIdx.x)) s = R::op(s, in[i]);
    out[threadIdx.x] = s;
  }
  if (threadIdx.x == 0) { header[0] = 0; header[1] = 0; }
}

// Wide cells: one thread per column, a slice of rows per blockIdx.y; partials[split][inner].
template <class T, int OP>
__global__ void __launch_bounds__(kMaxThreads)
reduce_axis0_cols_kernel(const T* __restrict__ in, const int64_t outer, const int64_t inner,
                         T* __restrict__ partials) {
  using R = Red<T, OP>;
  const int64_t c = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= inner) return;
  T acc[kJ];
#pragma unroll
  for (int j = 0; j < kJ; ++j) acc[j] = R::identity();
  int64_t r = blockIdx.y;
  const int64_t step = gridDim.y;
  for (; r + (kJ - 1) * step < outer; r += kJ * step) {
    T x[kJ];
#pragma unroll
    for (int j = 0; j < kJ; ++j) x[j] = __ldcs(in + (r + j * step) * inner + c);
#pragma unroll
    for (int j = 0; j < kJ; ++j) acc[j] = R::op(acc[j], x[j]);
  }
  for (; r < outer; r += step) acc[0] = R::op(acc[0], __ldcs(in + r * inner + c));
  partials[int64_t(blockIdx.y) * inner + c] = R::op(R::op(acc[0], acc[1]), R::op(acc[2], acc[3]));
}

template <class T, int OP>
__global__ void __launch_bounds__(kMaxThreads)
fold_partials_kernel(const T* __restrict__ partials, const int64_t splits, const int64_t inner, T* __restrict__ out) {
  using R = Red<T, OP>;
  const int64_t c = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (c >= inner) return;
  T s = R::identity();
  for (int64_t k = 0; k < splits; ++k) s = R::op(s, partials[k * inner + c]);
  out[c] = s;
}

// Row reduction: one warp per row, lanes stride over the row, fixed shuffle tree.
template <class T, int OP>
__global__ void __launch_bounds__(kMaxThreads)
reduce_rows_kernel(const T* __restrict__ in, const int64_t outer, const int64_t inner, T* __restrict__ out) {
  using R = Red<T, OP>;
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t row = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; row < outer; row += warps) {
    T s = R::identity();
    for (int64_t c = lane; c < inner; c += 32) s = R::op(s, __ldcs(in + row * inner + c));
    for (int m = 16; m >= 1; m >>= 1) s = R::op(s, shfl_xor(s, m));
    if (lane == 0) out[row] = s;
  }
}

// Middle-axis reduction: in[outer, len, inner] -> out[outer, inner] (mapRows over matrix cells: axis 0 of the cell is
// the middle axis of the block).  One thread per output element, neighbouring threads on neighbouring `inner`
// positions so every step of the fold is a coalesced read; the fold order is the axis order.
template <class T, int OP>
__global__ void __launch_bounds__(kMaxThreads)
reduce_mid_kernel(const T* __restrict__ in, const int64_t outer, const int64_t len, const int64_t inner, T* __restrict__ out) {
  using R = Red<T, OP>;
  const int64_t n = outer * inner;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += stride) {
    const int64_t o = t / inner, i = t % inner;
    const T* p = in + o * len * inner + i;
    T s = R::identity();
    for (int64_t a = 0; a < len; ++a) s = R::op(s, __ldcs(p + a * inner));
    out[t] = s;
  }
}

int64_t gcd64(int64_t a, int64_t b) { while (b) { int64_t t = a % b; a = b; b = t; } return a; }

struct FlatGeom { bool ok; int P, L, q, threads; int grid; };

FlatGeom flat_geometry(int es, const void* in, int64_t inner, const DeviceInfo& dev) {
  FlatGeom g{};
  g.P = (reinterpret_cast<uintptr_t>(in) & 15) ? 1 : 16 / es;
  int64_t L = inner / gcd64(inner, g.P) * g.P;
  int64_t q = L / g.P;
  g.ok = q <= kMaxQ && L <= kMaxThreads;
  if (!g.ok) return g;
  g.L = int(L);
  g.q = int(q);
  g.threads = (kMaxThreads / g.q) * g.q;
  g.grid = dev.sm_count * (2048 / kMaxThreads);
  return g;
}

// TFS_REDUCE_IMPL=ldg selects the register-staged kernel for A/B measurements; default is TMA.
int reduce_impl_from_env() {
  const char* e = std::getenv("TFS_REDUCE_IMPL");
  if (e && std::string(e) == "ldg") return 0;
  return 1;
}

// resident CTAs per SM of the LDG kernel (persistent single wave)
template <class T, int OP, int PV>
int blocks_per_sm(int P) {
  static int cached[2] = {0, 0};
  int& c = cached[P == 1 ? 0 : 1];
  if (c == 0) {
    int b = 0;
    if (P == 1) TFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, reduce_axis0_flat_kernel<T, OP, 1>, kMaxThreads, 0));
    else TFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, reduce_axis0_flat_kernel<T, OP, PV>, kMaxThreads, 0));
    c = b > 0 ? b : 1;
  }
  return c;
}

template <class T, int OP>
int launch_axis0_t(const T* in, int64_t outer, int64_t inner, T* out, void* scratch, const DeviceInfo& dev,
                   cudaStream_t stream) {
  const int64_t n = outer * inner;
  FlatGeom g = flat_geometry(int(sizeof(T)), in, inner, dev);
  unsigned int* counter = static_cast<unsigned int*>(scratch);
  T* partials = reinterpret_cast<T*>(static_cast<char*>(scratch) + kCounterBytes);
  if (g.ok) {
    int64_t n_vec = n / g.P;
    constexpr int PV = 16 / int(sizeof(T));
    static const int impl = reduce_impl_from_env();
    const int64_t chunk_vecs = int64_t(g.threads) * kVecPerThread;
    if (impl == 1 && g.P == PV && n_vec >= 4 * chunk_vecs) {
      // TMA-staged: one persistent CTA per SM
      const size_t dyn = size_t(kStages) * size_t(chunk_vecs) * 16;
      static PerDeviceOnce attr;
      attr.run(dev.device, [] {
        TFS_CUDA(cudaFuncSetAttribute(reduce_axis0_tma_kernel<T, OP, PV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      int(size_t(kStages) * kMaxThreads * kVecPerThread * 16)));
      });
      const int64_t n_chunks = (n_vec + chunk_vecs - 1) / chunk_vecs;
      int grid = int(std::max<int64_t>(1, std::min<int64_t>(dev.sm_count, n_chunks)));
      // units per launch: about 32 per CTA so that an SM that falls behind loses ~3 % of its share at most, fewer for
      // wide cells (every unit costs one L-element partial that the last CTA has to fold)
      int64_t want_units = std::min<int64_t>(int64_t(grid) * 32, std::max<int64_t>(grid, 65536 / g.L));
      int64_t unit_chunks = std::max<int64_t>(1, (n_chunks + want_units - 1) / want_units);
      unit_chunks = std::min<int64_t>(unit_chunks, int64_t(1) << 20);
      const int64_t n_units = (n_chunks + unit_chunks - 1) / unit_chunks;
      reduce_axis0_tma_kernel<T, OP, PV><<<grid, g.threads, dyn, stream>>>(in, n, int(inner), g.L, g.q, n_units, int(unit_chunks),
                                                                            partials, counter, out);
      TFS_CUDA(cudaGetLastError());
      return 1;
    }
    int64_t want = (n_vec + int64_t(g.threads) * kJ - 1) / (int64_t(g.threads) * kJ);
    int grid = int(std::max<int64_t>(1, std::min<int64_t>(int64_t(dev.sm_count) * blocks_per_sm<T, OP, PV>(g.P), want)));
    if (g.P == 1)
      reduce_axis0_flat_kernel<T, OP, 1><<<grid, g.threads, 0, stream>>>(in, n, int(inner), g.L, g.q, partials, counter, out);
    else
      reduce_axis0_flat_kernel<T, OP, PV><<<grid, g.threads, 0, stream>>>(in, n, int(inner), g.L, g.q, partials, counter, out);
    TFS_CUDA(cudaGetLastError());
    return 1;
  }
  // wide cells
  const int threads = kMaxThreads;
  const int64_t col_blocks = (inner + threads - 1) / threads;
  int64_t splits = std::max<int64_t>(1, (int64_t(dev.sm_count) * 8 + col_blocks - 1) / col_blocks);
  splits = std::min<int64_t>(splits, std::max<int64_t>(1, outer / 4));
  splits = std::min<int64_t>(splits, 65535);
  reduce_axis0_cols_kernel<T, OP><<<dim3((unsigned)col_blocks, (unsigned)splits), threads, 0, stream>>>(in, outer, inner, partials);
  TFS_CUDA(cudaGetLastError());
  fold_partials_kernel<T, OP><<<(unsigned)col_blocks, threads, 0, stream>>>(partials, splits, inner, out);
  TFS_CUDA(cudaGetLastError());
  return 2;
}

template <class T>
int launch_axis0_op(int red_op, const void* in, int64_t outer, int64_t inner, void* out, void* scratch,
                    const DeviceInfo& dev, cudaStream_t stream) {
  const T* i = static_cast<const T*>(in);
  T* o = static_cast<T*>(out);
  switch (red_op) {
    case R_SUM: return launch_axis0_t<T, R_SUM>(i, outer, inner, o, scratch, dev, stream);
    case R_MIN: return launch_axis0_t<T, R_MIN>(i, outer, inner, o, scratch, dev, stream);
    default: return launch_axis0_t<T, R_MAX>(i, outer, inner, o, scratch, dev, stream);
  }
}

template <class T>
int launch_rows_op(int red_op, const void* in, int64_t outer, int64_t inner, void* out, const DeviceInfo& dev,
                   cudaStream_t stream) {
  const T* i = static_cast<const T*>(in);
  T* o = static_cast<T*>(out);
  const int threads = kMaxThreads;
  int64_t grid = std::min<int64_t>((outer * 32 + threads - 1) / threads, int64_t(dev.sm_count) * 8);
  if (grid < 1) grid = 1;
  switch (red_op) {
    case R_SUM: reduce_rows_kernel<T, R_SUM><<<(unsigned)grid, threads, 0, stream>>>(i, outer, inner, o); break;
    case R_MIN: reduce_rows_kernel<T, R_MIN><<<(unsigned)grid, threads, 0, stream>>>(i, outer, inner, o); break;
    default: reduce_rows_kernel<T, R_MAX><<<(unsigned)grid, threads, 0, stream>>>(i, outer, inner, o); break;
  }
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace

size_t reduce_scratch_bytes(int dtype, int64_t outer, int64_t inner, const DeviceInfo& dev) {
  (void)outer;
  const size_t es = size_t(dtype_size(dtype));
  // flat kernel: grid * L partials with L <= 256; column kernel: splits * inner with
  // splits <= sm_count * 8 + 1.  Take the larger bound of the two layouts.
  size_t flat = size_t(dev.sm_count) * 8 * kMaxThreads * es;
  size_t cols = (size_t(dev.sm_count) * 8 + 1) * size_t(inner) * es;
  return kCounterBytes + (flat > cols ? flat : cols);
}

int launch_reduce_axis0(int dtype, int red_op, const void* in, int64_t outer, int64_t inner, void* out,
                        void* scratch, const DeviceInfo& dev, cudaStream_t stream) {
  switch (dtype) {
    case TFS_DOUBLE: return launch_axis0_op<double>(red_op, in, outer, inner, out, scratch, dev, stream);
    case TFS_FLOAT: return launch_axis0_op<float>(red_op, in, outer, inner, out, scratch, dev, stream);
    case TFS_INT32: return launch_axis0_op<int>(red_op, in, outer, inner, out, scratch, dev, stream);
    case TFS_INT64: return launch_axis0_op<long long>(red_op, in, outer, inner, out, scratch, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "reduce: unsupported dtype");
}

namespace {
template <class T>
int launch_mid_op(int red_op, const void* in, int64_t outer, int64_t len, int64_t inner, void* out, const DeviceInfo& dev,
                  cudaStream_t stream) {
  const T* i = static_cast<const T*>(in);
  T* o = static_cast<T*>(out);
  const int threads = kMaxThreads;
  int64_t grid = std::min<int64_t>((outer * inner + threads - 1) / threads, int64_t(dev.sm_count) * 8);
  if (grid < 1) grid = 1;
  switch (red_op) {
    case R_SUM: reduce_mid_kernel<T, R_SUM><<<(unsigned)grid, threads, 0, stream>>>(i, outer, len, inner, o); break;
    case R_MIN: reduce_mid_kernel<T, R_MIN><<<(unsigned)grid, threads, 0, stream>>>(i, outer, len, inner, o); break;
    default: reduce_mid_kernel<T, R_MAX><<<(unsigned)grid, threads, 0, stream>>>(i, outer, len, inner, o); break;
  }
  TFS_CUDA(cudaGetLastError());
  return 1;
}
}  // namespace

int launch_reduce_mid(int dtype, int red_op, const void* in, int64_t outer, int64_t len, int64_t inner, void* out,
                      const DeviceInfo& dev, cudaStream_t stream) {
  if (outer * inner <= 0) return 0;
  switch (dtype) {
    case TFS_DOUBLE: return launch_mid_op<double>(red_op, in, outer, len, inner, out, dev, stream);
    case TFS_FLOAT: return launch_mid_op<float>(red_op, in, outer, len, inner, out, dev, stream);
    case TFS_INT32: return launch_mid_op<int>(red_op, in, outer, len, inner, out, dev, stream);
    case TFS_INT64: return launch_mid_op<long long>(red_op, in, outer, len, inner, out, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "reduce: unsupported dtype");
}

int launch_reduce_rows(int dtype, int red_op, const void* in, int64_t outer, int64_t inner, void* out,
                       const DeviceInfo& dev, cudaStream_t stream) {
  if (outer <= 0) return 0;
  switch (dtype) {
    case TFS_DOUBLE: return launch_rows_op<double>(red_op, in, outer, inner, out, dev, stream);
    case TFS_FLOAT: return launch_rows_op<float>(red_op, in, outer, inner, out, dev, stream);
    case TFS_INT32: return launch_rows_op<int>(red_op, in, outer, inner, out, dev, stream);
    case TFS_INT64: return launch_rows_op<long long>(red_op, in, outer, inner, out, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "reduce: unsupported dtype");
}

}  // namespace tfs
