// K3 -- keyed block reduction for `groupBy(key).aggregate(reduce_*(x_input, [0]))` (sm_100a).
//
// Stands in for the whole TensorFlowUDAF machinery of the reference
// (impl/DebugRowOps.scala:601-695): there, Spark hashes rows to per-key buffers of <= 11 rows
// and block-reduces each buffer through a fresh TF session.  Here one block is reduced in
// four passes, all HBM-bound:
//
//   1. insert   every key into an open-addressing table (64-bit CAS on the key slot); the row
//               remembers its slot (u32)                                    8 B read + 4 B written / row
//               (small tables, and blocks with a carried scalar fetch, also count the rows of every slot here)
//   2. scan     one pass over the table: dense group id of every occupied slot, out_keys (and the
//               segment offset of every slot where rows were counted)      (table only)
//   3. sort     rows by slot: a stable radix partition of the (slot, row) pairs on the slot bits, three 8-bit passes
//               for tables of up to 2^24 slots (kernels_radix.cu); the first pass takes the row index from the
//               element's position, the last one writes row indices only and records where every slot's
//               segment begins and ends
//   4. gather-reduce  one warp per group walks its segment of the sorted row list and gathers
//               the payload rows -- each a contiguous cell, read with coalesced vector loads --
//               accumulating in registers; one coalesced store per group.      cell bytes / row
//
// The payload (the dominant traffic: 256 B/row at BASELINE config 4) is read exactly once and
// no accumulator table is read-modify-written, unlike a scatter-add formulation.  Keys are
// compared bit-exactly.  Which slot a key lands in depends on who wins a contended probe, so the
// table order of the groups is not reproducible; the (few) groups are therefore sorted by key before
// the gather, and -- the row sort being stable -- the rows of a group are folded in block order:
// output order and values are run-to-run deterministic, floats included.
#include <cstdlib>

#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;
constexpr unsigned long long kEmpty = 0x8000000000000000ULL;

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
  x ^= x >> 27; x *= 0x94d049bb133111ebULL;
  x ^= x >> 31;
  return x;
}

// header[0] = occupied slots so far, header[1] = overflow flag, header[2] = number of groups (after the scan)
// Every row records its slot; its own index needs no column: the first partition pass reads it off the position.
// Slot `slots` (one past the table) belongs to the one key value that equals the empty marker.
//
// Each thread takes kInsU keys of a CTA tile: the first probes of all of them are in flight together, the
// (rare) collisions are walked afterwards, and only when the whole warp is done -- __syncwarp -- are the
// row slots stored.  Without that reconvergence every lane stores as it leaves its own probe loop, the
// warp's 256 B store turns into 32 partial-sector writes, and L2 has to read every sector back from DRAM
// to merge them: measured 85 B/row of DRAM reads instead of 8, 9.2 ms instead of 3.0 ms per 2.5e8 rows.
//
// SMEM_COUNTS: tables of at most kSmemSlots slots count in shared memory and flush once per CTA; with few
// distinct keys the global counters are a handful of addresses that every row would hit (13 ms per
// 2.5e8 rows at 1000 keys, serialised in the L2 atomic units).
constexpr int kInsU = 2;
constexpr int kSmemSlots = 8192;

template <bool SMEM_COUNTS>
__global__ void __launch_bounds__(kThreads)
agg_insert_kernel(const long long* __restrict__ keys, const int64_t n, unsigned long long* __restrict__ slot_key,
                  unsigned int* __restrict__ counts, const unsigned long long slots, unsigned int* __restrict__ header,
                  const unsigned int max_groups, unsigned int* __restrict__ row_slot, const bool count_rows) {
  __shared__ unsigned int sm_counts[SMEM_COUNTS ? kSmemSlots + 1 : 1];
  if (SMEM_COUNTS) {
    for (unsigned int j = threadIdx.x; j <= (unsigned int)slots; j += kThreads) sm_counts[j] = 0u;
    __syncthreads();
  }
  const unsigned long long mask = slots - 1;
  const unsigned long long* ukeys = reinterpret_cast<const unsigned long long*>(keys);
  const int64_t tile = int64_t(kThreads) * kInsU;
  const int64_t tiles = (n + tile - 1) / tile;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {   // CTA-uniform trip count
    const int64_t base = t * tile + threadIdx.x;
    unsigned long long k[kInsU], h[kInsU], cur[kInsU];
#pragma unroll
    for (int u = 0; u < kInsU; ++u) {
      const int64_t i = base + int64_t(u) * kThreads;
      k[u] = i < n ? __ldcs(ukeys + i) : 0ULL;
    }
#pragma unroll
    for (int u = 0; u < kInsU; ++u) {
      h[u] = k[u] == kEmpty ? slots : (mix64(k[u]) & mask);
      cur[u] = *reinterpret_cast<volatile unsigned long long*>(slot_key + h[u]);
    }
    unsigned int claimed = 0;
#pragma unroll
    for (int u = 0; u < kInsU; ++u) {
      if (base + int64_t(u) * kThreads >= n) continue;
      if (k[u] == kEmpty) {   // the marker-key slot starts as 0, not as the marker
        if (cur[u] == 0ULL && atomicCAS(slot_key + slots, 0ULL, kEmpty) == 0ULL) ++claimed;
        continue;
      }
      unsigned long long c = cur[u];
      unsigned int probes = 0;
      while (true) {
        if (c == k[u]) break;
        if (c == kEmpty) {
          const unsigned long long prev = atomicCAS(slot_key + h[u], kEmpty, k[u]);   // rare: once per distinct key
          if (prev == kEmpty) { ++claimed; break; }
          if (prev == k[u]) break;
        }
        h[u] = (h[u] + 1) & mask;
        // The table has at least 2 * max_groups slots, so a run of more than max_groups occupied slots holding other
        // keys proves the block has more distinct keys than max_groups: raise the overflow flag from this lane and
        // stop.  (Waiting for another warp to publish its claims is not enough: with the table full every warp has a
        // lane stuck in this loop, so nobody would ever get to publish.)  Other lanes look at the flag now and then --
        // not on every probe: the whole grid would hammer one address -- and give up when it is set.
        if (++probes > max_groups) { *reinterpret_cast<volatile unsigned int*>(header + 1) = 1u; break; }
        if ((probes & 255u) == 0u && *reinterpret_cast<volatile unsigned int*>(header + 1)) break;
        c = *reinterpret_cast<volatile unsigned long long*>(slot_key + h[u]);
      }
    }
    __syncwarp();   // see above: the stores below must leave as whole-warp, whole-sector writes
    // one counter update per warp: a million first occurrences would otherwise queue on this one address
    const unsigned int warp_claimed = __reduce_add_sync(0xffffffffu, claimed);
    if ((threadIdx.x & 31) == 0 && warp_claimed != 0u && atomicAdd(header, warp_claimed) + warp_claimed > max_groups)
      header[1] = 1u;
#pragma unroll
    for (int u = 0; u < kInsU; ++u) {
      const int64_t i = base + int64_t(u) * kThreads;
      if (i < n) {
        __stcs(row_slot + i, (unsigned int)h[u]);
        if (SMEM_COUNTS) atomicAdd(sm_counts + h[u], 1u);
        else if (count_rows) atomicAdd(counts + h[u], 1u);
      }
    }
  }
  if (SMEM_COUNTS) {
    __syncthreads();
    for (unsigned int j = threadIdx.x; j <= (unsigned int)slots; j += kThreads)
      if (sm_counts[j] != 0u) atomicAdd(counts + j, sm_counts[j]);
  }
}

// counts[i] = 1 for a taken slot, 0 for a free one (the marker-key slot, index `slots`, starts as 0 and holds the marker
// once that key has been seen)
__global__ void __launch_bounds__(kThreads)
agg_mark_occupied_kernel(const unsigned long long* __restrict__ slot_key, const unsigned long long slots, unsigned int* __restrict__ counts) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i <= slots; i += stride)
    counts[i] = i < slots ? (slot_key[i] != kEmpty ? 1u : 0u) : (slot_key[i] == kEmpty ? 1u : 0u);
}

// counts[i] = rows of slot i, from the segment bounds the last partition pass recorded (lo = first position, hi = one past
// the last; untouched slots have lo = 0xffffffff, hi = 0)
__global__ void __launch_bounds__(kThreads)
agg_counts_from_bounds_kernel(const unsigned int* __restrict__ lo, const unsigned int* __restrict__ hi, const uint64_t m,
                              unsigned int* __restrict__ counts) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < m; i += stride) counts[i] = hi[i] != 0u ? hi[i] - lo[i] : 0u;
}

__global__ void __launch_bounds__(kThreads)
fill_u64_kernel(unsigned long long* __restrict__ p, const uint64_t n, const unsigned long long v) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

// ---- one scan over the table: (rows, occupied) prefix sums packed in 64 bits -------------------------
constexpr int kScanTile = 2048;  // slots per CTA

__device__ __forceinline__ unsigned long long pack_count(unsigned int c) {
  return (unsigned long long)c | ((unsigned long long)(c != 0u) << 32);
}

__global__ void __launch_bounds__(kThreads)
scan_tile_sums_kernel(const unsigned int* __restrict__ counts, const int64_t m, unsigned long long* __restrict__ tile_sums) {
  __shared__ unsigned long long sm[kThreads];
  const int64_t base = int64_t(blockIdx.x) * kScanTile;
  unsigned long long s = 0;
  for (int i = threadIdx.x; i < kScanTile; i += kThreads)
    if (base + i < m) s += pack_count(counts[base + i]);
  sm[threadIdx.x] = s;
  __syncthreads();
  for (int w = kThreads / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) sm[threadIdx.x] += sm[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = sm[0];
}

__global__ void __launch_bounds__(1024)
scan_tile_offsets_kernel(unsigned long long* __restrict__ tile_sums, const int64_t tiles, unsigned int* __restrict__ header) {
  // single CTA: in-place exclusive scan, 1024 tiles per round with a running carry
  __shared__ unsigned long long sm[1024];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < tiles; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const unsigned long long v = i < tiles ? tile_sums[i] : 0ULL;
    sm[threadIdx.x] = v;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
      unsigned long long t = threadIdx.x >= d ? sm[threadIdx.x - d] : 0ULL;
      __syncthreads();
      sm[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < tiles) tile_sums[i] = carry + sm[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += sm[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) header[2] = (unsigned int)(carry >> 32);   // number of groups
}

// offsets[slot] = rows before this slot; occupied slots get the next dense group id and publish their key.
__global__ void __launch_bounds__(kThreads)
scan_apply_kernel(const unsigned int* __restrict__ counts, const int64_t m, const unsigned long long* __restrict__ tile_off,
                  const unsigned long long* __restrict__ slot_key, unsigned int* __restrict__ offsets,
                  unsigned int* __restrict__ group_slot, long long* __restrict__ out_keys, const unsigned int max_groups) {
  constexpr int PER = kScanTile / kThreads;
  __shared__ unsigned long long sm[kThreads];
  const int64_t base = int64_t(blockIdx.x) * kScanTile + int64_t(threadIdx.x) * PER;
  unsigned int c[PER];
  unsigned long long s = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    c[j] = base + j < m ? counts[base + j] : 0u;
    s += pack_count(c[j]);
  }
  sm[threadIdx.x] = s;
  __syncthreads();
  for (int d = 1; d < kThreads; d <<= 1) {
    unsigned long long t = threadIdx.x >= d ? sm[threadIdx.x - d] : 0ULL;
    __syncthreads();
    sm[threadIdx.x] += t;
    __syncthreads();
  }
  unsigned long long run = tile_off[blockIdx.x] + sm[threadIdx.x] - s;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    if (base + j < m) {
      offsets[base + j] = (unsigned int)run;
      if (c[j] != 0u) {
        const unsigned int g = (unsigned int)(run >> 32);
        if (g < max_groups) {
          group_slot[g] = (unsigned int)(base + j);
          out_keys[g] = (long long)slot_key[base + j];
        }
      }
    }
    run += pack_count(c[j]);
  }
}

// ---- gather-reduce ---------------------------------------------------------------------------
template <class T> struct Lim;
template <> struct Lim<double> { static __device__ double hi() { return __longlong_as_double(0x7ff0000000000000LL); } };
template <> struct Lim<float> { static __device__ float hi() { return __int_as_float(0x7f800000); } };
template <> struct Lim<int> { static __device__ int hi() { return 0x7fffffff; } };
template <> struct Lim<long long> { static __device__ long long hi() { return 0x7fffffffffffffffLL; } };
template <class T> __device__ __forceinline__ T lowest_of();
template <> __device__ __forceinline__ double lowest_of<double>() { return -Lim<double>::hi(); }
template <> __device__ __forceinline__ float lowest_of<float>() { return -Lim<float>::hi(); }
template <> __device__ __forceinline__ int lowest_of<int>() { return int(0x80000000u); }
template <> __device__ __forceinline__ long long lowest_of<long long>() { return (long long)0x8000000000000000ULL; }

template <class T, int OP> __device__ __forceinline__ T red_identity() {
  if (OP == R_SUM) return T(0);
  if (OP == R_MIN) return Lim<T>::hi();
  return lowest_of<T>();
}
template <class T> __device__ __forceinline__ T add_wrap(T a, T b) { return a + b; }
template <> __device__ __forceinline__ int add_wrap(int a, int b) { return int((unsigned)a + (unsigned)b); }
template <> __device__ __forceinline__ long long add_wrap(long long a, long long b) {
  return (long long)((unsigned long long)a + (unsigned long long)b);
}
template <> __device__ __forceinline__ float add_wrap(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ double add_wrap(double a, double b) { return __dadd_rn(a, b); }
template <class T, int OP> __device__ __forceinline__ T red_op(T a, T b) {
  if (OP == R_SUM) return add_wrap<T>(a, b);
  if (OP == R_MIN) return (b < a) ? b : a;
  return (a < b) ? b : a;
}

template <class T, int V> struct VecT;
template <> struct VecT<float, 4> { using type = float4; };
template <> struct VecT<float, 2> { using type = float2; };
template <> struct VecT<int, 4> { using type = int4; };
template <> struct VecT<int, 2> { using type = int2; };
template <> struct VecT<double, 2> { using type = double2; };
template <> struct VecT<long long, 2> { using type = longlong2; };

template <class T, int V>
__device__ __forceinline__ void load_v(const T* p, T (&x)[V]) {
  if constexpr (V == 1) {
    x[0] = __ldcs(p);
  } else {
    using VT = typename VecT<T, V>::type;
    VT t = __ldcs(reinterpret_cast<const VT*>(p));
    if constexpr (V == 2) { x[0] = t.x; x[1] = t.y; }
    else { x[0] = t.x; x[1] = t.y; x[2] = t.z; x[3] = t.w; }
  }
}

// One warp per work item = (group, part): normally one part per group; when there are few, large groups
// (k-means: 10 segments, a low-cardinality groupBy) every group's segment is cut into `parts` equal pieces so
// that the whole chip has work, the pieces go to a partial buffer and fold_parts_kernel combines them in order.
// The warp is split into R row-lanes x C col-lanes (R * C = 32, C a power of two >= ceil(cell / V) when the
// cell fits, else C = 32 with a loop over column chunks).  kU rows are in flight per row-lane.
template <class T, int OP, int V, bool PERM>
__global__ void __launch_bounds__(kThreads)
agg_gather_reduce_kernel(const T* __restrict__ payload, const int cell, const unsigned int* __restrict__ perm,
                         const unsigned int* __restrict__ offsets, const unsigned int* __restrict__ counts,
                         const unsigned int* __restrict__ group_slot, const int64_t n_groups, const int col_lanes,
                         const int parts, T* __restrict__ out) {
  constexpr int kU = 8;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const int C = col_lanes, R = 32 / C;
  const int cl = lane % C, rl = lane / C;
  const int64_t n_items = n_groups * parts;
  for (int64_t g = warp; g < n_items; g += warps) {          // g indexes the output row: the group, or (group, part)
    const unsigned int slot = group_slot[g / parts];
    const unsigned long long part = (unsigned long long)(g % parts), cnt = counts[slot];
    const unsigned long long lo = offsets[slot] + cnt * part / (unsigned long long)parts;
    const unsigned long long hi = offsets[slot] + cnt * (part + 1) / (unsigned long long)parts;
    for (int cbase = 0; cbase < cell; cbase += C * V) {   // warp-uniform trip count
      const int c0 = cbase + cl * V;
      const bool active = c0 < cell;
      T acc[V];
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] = red_identity<T, OP>();
      if (active) {
        unsigned long long r = lo + rl;
        for (; r + (unsigned long long)(kU - 1) * R < hi; r += (unsigned long long)kU * R) {
          unsigned int idx[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u)
            idx[u] = PERM ? __ldg(perm + r + (unsigned long long)u * R) : (unsigned int)(r + (unsigned long long)u * R);
          T x[kU][V];
#pragma unroll
          for (int u = 0; u < kU; ++u) load_v<T, V>(payload + (int64_t)idx[u] * cell + c0, x[u]);
#pragma unroll
          for (int u = 0; u < kU; ++u)
#pragma unroll
            for (int e = 0; e < V; ++e) acc[e] = red_op<T, OP>(acc[e], x[u][e]);
        }
        for (; r < hi; r += R) {
          T x[V];
          load_v<T, V>(payload + (int64_t)(PERM ? __ldg(perm + r) : (unsigned int)r) * cell + c0, x);
#pragma unroll
          for (int e = 0; e < V; ++e) acc[e] = red_op<T, OP>(acc[e], x[e]);
        }
      }
      // fold the R row-lanes that share this column slice (every lane takes part)
      for (int m = 16; m >= C; m >>= 1) {
#pragma unroll
        for (int e = 0; e < V; ++e) acc[e] = red_op<T, OP>(acc[e], __shfl_xor_sync(0xffffffffu, acc[e], m));
      }
      if (active && rl == 0) {
#pragma unroll
        for (int e = 0; e < V; ++e) out[g * cell + c0 + e] = acc[e];
      }
    }
  }
}

int64_t grid_for(int64_t n, const DeviceInfo& dev);

// out[g, c] = fold over part of partial[(g * parts + part), c], in part order
template <class T, int OP>
__global__ void __launch_bounds__(kThreads)
fold_parts_kernel(const T* __restrict__ partial, const int64_t n_groups, const int64_t cell, const int parts, T* __restrict__ out) {
  const int64_t n = n_groups * cell;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t g = i / cell, c = i % cell;
    T acc = partial[(g * parts) * cell + c];
    for (int p = 1; p < parts; ++p) acc = red_op<T, OP>(acc, partial[(g * parts + p) * cell + c]);
    out[i] = acc;
  }
}

template <class T, int OP>
int launch_gather_t(const T* payload, int64_t cell, const unsigned int* perm, const unsigned int* offsets,
                    const unsigned int* counts, const unsigned int* group_slot, int64_t n_groups, int64_t n_rows, T* out,
                    void* split_scratch, size_t split_bytes, const DeviceInfo& dev, cudaStream_t stream) {
  if (n_groups <= 0 || cell <= 0) return 0;
  constexpr int VMAX = 16 / int(sizeof(T));
  int V = 1;
  const bool aligned16 = (reinterpret_cast<uintptr_t>(payload) & 15) == 0;
  if (aligned16 && cell % VMAX == 0 && cell >= 8 * VMAX) V = VMAX;
  else if (aligned16 && VMAX == 4 && cell % 2 == 0 && cell >= 64) V = 2;
  if (perm == nullptr) V = 1;          // carried values: the scalar-load instantiation serves the (tiny) cell
  int need = int((cell + V - 1) / V);
  int C = 1;
  while (C < need && C < 32) C <<= 1;
  // few large groups: cut every segment into `parts` pieces so that about 32 warps per SM have work
  const int64_t target = int64_t(dev.sm_count) * 32;
  int parts = 1;
  if (split_scratch != nullptr && n_groups < target && n_rows / n_groups >= 2048) {
    int64_t p = std::min<int64_t>((target + n_groups - 1) / n_groups, 1024);
    p = std::min<int64_t>(p, n_rows / n_groups / 256);                       // at least 256 rows per piece
    p = std::min<int64_t>(p, int64_t(split_bytes / (size_t(n_groups) * size_t(cell) * sizeof(T))));
    if (p > 1) parts = int(p);
  }
  T* dst = parts > 1 ? static_cast<T*>(split_scratch) : out;
  const int64_t warps_per_cta = kThreads / 32;
  const int64_t items = n_groups * parts;
  int64_t grid = std::min<int64_t>((items + warps_per_cta - 1) / warps_per_cta, int64_t(dev.sm_count) * 16);
  if (grid < 1) grid = 1;
  const int icell = int(cell);
  if (perm == nullptr) {       // values already in segment order (carried through the sort): scalar cells only
    agg_gather_reduce_kernel<T, OP, 1, false><<<(unsigned)grid, kThreads, 0, stream>>>(payload, icell, perm, offsets, counts, group_slot, n_groups, C, parts, dst);
  } else if (V == 4) {
    if constexpr (VMAX == 4)
      agg_gather_reduce_kernel<T, OP, 4, true><<<(unsigned)grid, kThreads, 0, stream>>>(payload, icell, perm, offsets, counts, group_slot, n_groups, C, parts, dst);
  } else if (V == 2) {
    agg_gather_reduce_kernel<T, OP, 2, true><<<(unsigned)grid, kThreads, 0, stream>>>(payload, icell, perm, offsets, counts, group_slot, n_groups, C, parts, dst);
  } else {
    agg_gather_reduce_kernel<T, OP, 1, true><<<(unsigned)grid, kThreads, 0, stream>>>(payload, icell, perm, offsets, counts, group_slot, n_groups, C, parts, dst);
  }
  TFS_CUDA(cudaGetLastError());
  if (parts == 1) return 1;
  fold_parts_kernel<T, OP><<<(unsigned)grid_for(n_groups * cell, dev), kThreads, 0, stream>>>(dst, n_groups, cell, parts, out);
  TFS_CUDA(cudaGetLastError());
  return 2;
}

template <class T>
int launch_gather_op(int red_op, const void* payload, int64_t cell, const unsigned int* perm,
                     const unsigned int* offsets, const unsigned int* counts, const unsigned int* group_slot,
                     int64_t n_groups, int64_t n_rows, void* out, void* split_scratch, size_t split_bytes, const DeviceInfo& dev,
                     cudaStream_t stream) {
  const T* p = static_cast<const T*>(payload);
  T* o = static_cast<T*>(out);
  switch (red_op) {
    case R_SUM: return launch_gather_t<T, R_SUM>(p, cell, perm, offsets, counts, group_slot, n_groups, n_rows, o, split_scratch, split_bytes, dev, stream);
    case R_MIN: return launch_gather_t<T, R_MIN>(p, cell, perm, offsets, counts, group_slot, n_groups, n_rows, o, split_scratch, split_bytes, dev, stream);
    default: return launch_gather_t<T, R_MAX>(p, cell, perm, offsets, counts, group_slot, n_groups, n_rows, o, split_scratch, split_bytes, dev, stream);
  }
}

int64_t grid_for(int64_t n, const DeviceInfo& dev) {
  int64_t g = (n + kThreads - 1) / kThreads;
  g = std::min<int64_t>(g, int64_t(dev.sm_count) * 8);
  return g < 1 ? 1 : g;
}

int slot_bits(uint64_t slots) {   // bits needed for slot indices 0..slots (inclusive: the marker-key slot)
  int b = 1;
  while ((uint64_t(1) << b) <= slots) ++b;
  return b;
}

}  // namespace

AggLayout agg_layout(int64_t n_rows, int64_t max_groups, const DeviceInfo& dev) {
  AggLayout L{};
  uint64_t cap = 1024;
  // at least 2 slots per group; 4 while the table (key, count and offset per slot) stays well inside L2: fewer and
  // shorter probe walks (insert 3.0 -> 2.65 ms per 2.5e8 rows at 1e6 groups; 8 buys nothing more)
  uint64_t factor = uint64_t(std::max<int64_t>(max_groups, 1)) * 4 * 16 <= (uint64_t(96) << 20) ? 4 : 2;
  if (const char* e = std::getenv("TFS_AGG_TABLE_FACTOR")) {     // (A/B measurements of the table's load factor)
    const long v = std::atol(e);
    if (v >= 2 && v <= 16) factor = uint64_t(v);
  }
  while (cap < factor * uint64_t(std::max<int64_t>(max_groups, 1))) cap <<= 1;
  L.table_slots = cap;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  size_t off = 0;
  L.off_header = off; off += 256;
  L.off_slot_key = off; off += up((cap + 1) * 8);
  L.off_counts = off; off += up((cap + 1) * 4);
  L.off_offsets = off; off += up((cap + 2) * 4);
  L.off_ends = off; off += up((cap + 2) * 4);
  L.rows_counted = true;
  L.off_group_slot = off; off += up(size_t(max_groups + 1) * 4);
  L.off_group_slot_tab = off; off += up(size_t(max_groups + 1) * 4);     // table order, before the sort by key
  L.off_keys_tab = off; off += up(size_t(max_groups + 1) * 8);
  L.off_tile_sums = off; off += up(size_t((cap + 1 + kScanTile) / kScanTile + 1) * 8);
  // row partition: the slot column the insert pass wrote (kept: every carried fetch sorts from it again), the slot
  // column between passes (a second one only when the table needs a third pass) and two row-index columns
  const int passes = radix_passes(n_rows, slot_bits(cap));
  L.bounds_eligible = passes >= 2 && cap > uint64_t(kSmemSlots);      // (small tables count in shared memory: cheap)
  L.off_row_slot[0] = off; off += up(size_t(n_rows) * 4);
  // between passes: two packed (slot, row) columns of 8 bytes per row (the second only when there are three passes or
  // more); the carried path of scalar payloads uses the same memory as 4-byte slot and value columns
  L.off_pairs[0] = off; off += up(size_t(n_rows) * 8);
  L.off_pairs[1] = off; off += passes >= 3 ? up(size_t(n_rows) * 8) : 0;
  L.off_row_slot[1] = L.off_pairs[0];
  L.off_row_slot[2] = passes >= 3 ? L.off_pairs[1] : L.off_pairs[0] + up(size_t(n_rows) * 4);
  L.off_row_id[0] = off; off += up(size_t(n_rows) * 4);
  L.off_row_id[1] = L.off_row_id[0];
  L.off_perm = L.off_row_id[0];   // set by launch_agg_segments to the buffer the partition finished in
  // group sort by key: one more (key, slot) column pair; then the histogram table of a pass
  L.off_group_tmp_keys = off; off += up(size_t(max_groups + 1) * 8);
  L.off_group_tmp_slot = off; off += up(size_t(max_groups + 1) * 4);
  L.sort_temp_bytes = radix_scratch_bytes(dev);
  L.off_sort_temp = off; off += up(L.sort_temp_bytes);
  L.bytes = off;
  return L;
}

int launch_agg_build(const int64_t* keys, int64_t n_rows, int64_t max_groups, char* scratch, const AggLayout& L,
                     const DeviceInfo& dev, cudaStream_t stream) {
  unsigned int* header = reinterpret_cast<unsigned int*>(scratch + L.off_header);
  unsigned long long* slot_key = reinterpret_cast<unsigned long long*>(scratch + L.off_slot_key);
  unsigned int* counts = reinterpret_cast<unsigned int*>(scratch + L.off_counts);
  unsigned int* offsets = reinterpret_cast<unsigned int*>(scratch + L.off_offsets);
  unsigned int* group_slot = reinterpret_cast<unsigned int*>(scratch + L.off_group_slot_tab);
  long long* keys_tab = reinterpret_cast<long long*>(scratch + L.off_keys_tab);
  unsigned long long* tile_sums = reinterpret_cast<unsigned long long*>(scratch + L.off_tile_sums);
  unsigned int* row_slot = reinterpret_cast<unsigned int*>(scratch + L.off_row_slot[0]);
  const int64_t m = int64_t(L.table_slots) + 1;   // table + the marker-key slot
  const bool count_rows = L.rows_counted;
  TFS_CUDA(cudaMemsetAsync(header, 0, 256, stream));
  TFS_CUDA(cudaMemsetAsync(counts, 0, size_t(m) * 4, stream));
  fill_u64_kernel<<<(unsigned)grid_for(int64_t(L.table_slots), dev), kThreads, 0, stream>>>(slot_key, L.table_slots, kEmpty);
  TFS_CUDA(cudaGetLastError());
  TFS_CUDA(cudaMemsetAsync(slot_key + L.table_slots, 0, 8, stream));
  int launches = 1;
  if (n_rows > 0) {
    const int64_t tiles = (n_rows + int64_t(kThreads) * kInsU - 1) / (int64_t(kThreads) * kInsU);
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(tiles, int64_t(dev.sm_count) * 16));
    if (L.table_slots <= uint64_t(kSmemSlots))
      agg_insert_kernel<true><<<grid, kThreads, 0, stream>>>(reinterpret_cast<const long long*>(keys), n_rows, slot_key, counts,
                                                             L.table_slots, header, (unsigned int)max_groups, row_slot, count_rows);
    else
      agg_insert_kernel<false><<<grid, kThreads, 0, stream>>>(reinterpret_cast<const long long*>(keys), n_rows, slot_key, counts,
                                                              L.table_slots, header, (unsigned int)max_groups, row_slot, count_rows);
    TFS_CUDA(cudaGetLastError());
    ++launches;
  }
  if (!count_rows) {
    // the rows are not counted while they are inserted (250 M scattered atomics): the table scan only needs to know
    // which slots are taken, and the segment of every slot comes out of the last partition pass
    agg_mark_occupied_kernel<<<(unsigned)grid_for(m, dev), kThreads, 0, stream>>>(slot_key, L.table_slots, counts);
    TFS_CUDA(cudaGetLastError());
    ++launches;
  }
  const int64_t tiles = (m + kScanTile - 1) / kScanTile;
  scan_tile_sums_kernel<<<(unsigned)tiles, kThreads, 0, stream>>>(counts, m, tile_sums);
  TFS_CUDA(cudaGetLastError());
  scan_tile_offsets_kernel<<<1, 1024, 0, stream>>>(tile_sums, tiles, header);
  TFS_CUDA(cudaGetLastError());
  scan_apply_kernel<<<(unsigned)tiles, kThreads, 0, stream>>>(counts, m, tile_sums, slot_key, offsets, group_slot, keys_tab,
                                                            (unsigned int)max_groups);
  TFS_CUDA(cudaGetLastError());
  return launches + 3;
}

int launch_agg_segments(const int64_t* keys, int64_t n_rows, int64_t n_groups, char* scratch, AggLayout& L,
                        int64_t* out_keys_dev, int group_order, bool sort_rows, const DeviceInfo& dev, cudaStream_t stream) {
  (void)keys;
  if (n_rows <= 0) return 0;
  int launches = 0;
  void* temp = scratch + L.off_sort_temp;
  // groups by key (signed order: the sign bit is flipped for the digit extraction): out_keys ascending, group_slot
  // permuted alongside.  64 key bits are six passes over a column that is small next to the rows.
  if (group_order == kGroupsTableOrder && n_groups > 0) {
    // groups stay in table order (not reproducible from run to run): for partial results that a merge will reorder
    TFS_CUDA(cudaMemcpyAsync(out_keys_dev, scratch + L.off_keys_tab, size_t(n_groups) * 8, cudaMemcpyDeviceToDevice, stream));
    TFS_CUDA(cudaMemcpyAsync(scratch + L.off_group_slot, scratch + L.off_group_slot_tab, size_t(n_groups) * 4, cudaMemcpyDeviceToDevice, stream));
  }
  if (group_order == kGroupsByKey && n_groups > 0) {
    // the passes alternate between two buffer pairs: order them so that the last pass lands in the output
    const bool even = radix_passes(n_groups, 64) % 2 == 0;
    void* tmp_k = scratch + L.off_group_tmp_keys;
    void* tmp_v = scratch + L.off_group_tmp_slot;
    void* out_v = scratch + L.off_group_slot;
    void* kbuf[2] = {even ? tmp_k : static_cast<void*>(out_keys_dev), even ? static_cast<void*>(out_keys_dev) : tmp_k};
    void* vbuf[2] = {even ? tmp_v : out_v, even ? out_v : tmp_v};
    const int at = launch_radix_sort(8, 4, scratch + L.off_keys_tab, scratch + L.off_group_slot_tab, kbuf, vbuf, n_groups,
                                     0x8000000000000000ULL, 64, true, temp, dev, stream, &launches);
    TFS_CHECK(kbuf[at] == static_cast<void*>(out_keys_dev), TFS_ERR_INVALID, "internal: group sort ended in the wrong buffer");
  }
  if (!sort_rows) return launches;       // every fetch carries its values through its own sort (launch_agg_reduce_carried)
  // rows by slot, stable: values = row indices, read off the position in the first pass
  unsigned int* seg_lo = nullptr;
  unsigned int* seg_hi = nullptr;
  const uint64_t m = L.table_slots + 1;
  if (!L.rows_counted) {
    seg_lo = reinterpret_cast<unsigned int*>(scratch + L.off_offsets);
    seg_hi = reinterpret_cast<unsigned int*>(scratch + L.off_ends);
    TFS_CUDA(cudaMemsetAsync(seg_lo, 0xff, size_t(m) * 4, stream));
    TFS_CUDA(cudaMemsetAsync(seg_hi, 0, size_t(m) * 4, stream));
  }
  launches += launch_radix_rows_by_slot(reinterpret_cast<const unsigned int*>(scratch + L.off_row_slot[0]), n_rows, slot_bits(L.table_slots),
                                        reinterpret_cast<unsigned long long*>(scratch + L.off_pairs[0]),
                                        reinterpret_cast<unsigned long long*>(scratch + L.off_pairs[1]),
                                        reinterpret_cast<unsigned int*>(scratch + L.off_row_id[0]), seg_lo, seg_hi, temp, dev, stream);
  if (!L.rows_counted) {
    agg_counts_from_bounds_kernel<<<(unsigned)grid_for(int64_t(m), dev), kThreads, 0, stream>>>(
        seg_lo, seg_hi, m, reinterpret_cast<unsigned int*>(scratch + L.off_counts));
    TFS_CUDA(cudaGetLastError());
    ++launches;
  }
  L.off_perm = L.off_row_id[0];
  return launches;
}

// Scalar payloads (one 4- or 8-byte value per row: groupBy(key).sum(x)): gathering them through the row permutation
// costs a 128-byte DRAM fill per value (measured 32.9 GB of reads for 2 GB of payload).  Instead the values ride
// through the stable slot sort themselves and the per-group reduction reads them sequentially.  Same fold order as the
// gather path, so the same bits.  Must run before the row sort (which reuses the unsorted slot buffer).
size_t agg_carried_bytes(int64_t n_rows, int elem_size, const AggLayout& L) {
  (void)L;
  if (n_rows <= 0) return 0;
  return 2 * ((size_t(n_rows) * size_t(elem_size) + 255) / 256 * 256);      // two value columns, alternating between passes
}

int launch_agg_reduce_carried(int dtype, int red_op, const void* payload, int64_t cell, int64_t n_groups, int64_t n_rows, char* scratch,
                              const AggLayout& L, void* out, char* carried_buf, size_t carried_bytes, void* split_scratch,
                              size_t split_bytes, const DeviceInfo& dev, cudaStream_t stream) {
  const int es = dtype_size(dtype) * int(cell);      // bytes per row: 4 or 8
  TFS_CHECK(es == 4 || es == 8, TFS_ERR_INVALID, "internal: carried rows must be 4 or 8 bytes");
  const size_t values_bytes = (size_t(n_rows) * size_t(es) + 255) / 256 * 256;
  TFS_CHECK(carried_bytes >= 2 * values_bytes, TFS_ERR_INVALID, "internal: carried-sort scratch too small");
  void* kbuf[2] = {scratch + L.off_row_slot[1], scratch + L.off_row_slot[2]};
  void* vbuf[2] = {carried_buf, carried_buf + values_bytes};
  int sort_launches = 0;
  const int at = launch_radix_sort(4, es, scratch + L.off_row_slot[0], payload, kbuf, vbuf, n_rows, 0, slot_bits(L.table_slots), false,
                                   scratch + L.off_sort_temp, dev, stream, &sort_launches);
  const char* carried = static_cast<const char*>(vbuf[at]);
  const unsigned int* offsets = reinterpret_cast<const unsigned int*>(scratch + L.off_offsets);
  const unsigned int* counts = reinterpret_cast<const unsigned int*>(scratch + L.off_counts);
  const unsigned int* group_slot = reinterpret_cast<const unsigned int*>(scratch + L.off_group_slot);
  if (split_bytes == 0) split_scratch = nullptr;
  const unsigned int* no_perm = nullptr;      // the values are already in segment order
  switch (dtype) {
    case TFS_FLOAT: return sort_launches + launch_gather_op<float>(red_op, carried, cell, no_perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_DOUBLE: return sort_launches + launch_gather_op<double>(red_op, carried, cell, no_perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_INT32: return sort_launches + launch_gather_op<int>(red_op, carried, cell, no_perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_INT64: return sort_launches + launch_gather_op<long long>(red_op, carried, cell, no_perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "aggregate: unsupported dtype");
}

size_t agg_split_scratch_bytes(int64_t cell, const DeviceInfo& dev) {
  // room for the partial rows of split groups: at most 2 x (32 warps per SM) work items of one cell each;
  // very wide cells go without (one warp per group then)
  const size_t bytes = size_t(2) * size_t(dev.sm_count) * 32 * size_t(std::max<int64_t>(cell, 1)) * 8;
  return bytes <= (size_t(256) << 20) ? bytes : 0;
}

int launch_agg_reduce(int dtype, int red_op, const void* payload, int64_t cell, int64_t n_groups, int64_t n_rows, char* scratch,
                      const AggLayout& L, void* out, void* split_scratch, size_t split_bytes, const DeviceInfo& dev,
                      cudaStream_t stream) {
  const unsigned int* perm = reinterpret_cast<const unsigned int*>(scratch + L.off_perm);
  const unsigned int* offsets = reinterpret_cast<const unsigned int*>(scratch + L.off_offsets);
  const unsigned int* counts = reinterpret_cast<const unsigned int*>(scratch + L.off_counts);
  const unsigned int* group_slot = reinterpret_cast<const unsigned int*>(scratch + L.off_group_slot);
  if (split_bytes == 0) split_scratch = nullptr;
  switch (dtype) {
    case TFS_FLOAT: return launch_gather_op<float>(red_op, payload, cell, perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_DOUBLE: return launch_gather_op<double>(red_op, payload, cell, perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_INT32: return launch_gather_op<int>(red_op, payload, cell, perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
    case TFS_INT64: return launch_gather_op<long long>(red_op, payload, cell, perm, offsets, counts, group_slot, n_groups, n_rows, out, split_scratch, split_bytes, dev, stream);
  }
  fail(TFS_ERR_UNSUPPORTED, "aggregate: unsupported dtype");
}

// ---- the exchange step of aggregate over several GPUs ---------------------------------------------------------
// owner[i] = (mix64(key[i]) >> 32) mod world -- the rank that merges group i -- and how many groups go to each rank.
// (The high half: the low bits of the same mix pick the table slot of the merge on the owner.)
namespace {
__global__ void __launch_bounds__(kThreads)
agg_owner_kernel(const long long* __restrict__ keys, const int64_t n, const unsigned int world, unsigned int* __restrict__ owner,
                 unsigned long long* __restrict__ counts) {
  __shared__ unsigned int sm[64];
  if (threadIdx.x < 64) sm[threadIdx.x] = 0u;
  __syncthreads();
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned int o = (unsigned int)((mix64((unsigned long long)keys[i]) >> 32) % world);
    owner[i] = o;
    atomicAdd(&sm[o], 1u);
  }
  __syncthreads();
  if (threadIdx.x < world && sm[threadIdx.x] != 0u) atomicAdd(counts + threadIdx.x, (unsigned long long)sm[threadIdx.x]);
}

// dst[i, :] = src[perm[i], :] for rows of `words` 4-byte words
__global__ void __launch_bounds__(kThreads)
agg_gather_rows_kernel(const unsigned int* __restrict__ src, const unsigned int* __restrict__ perm, const int64_t n, const int64_t words,
                       unsigned int* __restrict__ dst) {
  const int64_t total = n * words;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int64_t row = e / words, w = e - row * words;
    dst[e] = src[int64_t(perm[row]) * words + w];
  }
}
}  // namespace

int launch_agg_owner(const int64_t* keys, int64_t n, int world, unsigned int* owner, unsigned long long* counts, const DeviceInfo& dev,
                     cudaStream_t stream) {
  TFS_CHECK(world >= 1 && world <= 64, TFS_ERR_UNSUPPORTED, "aggregate exchange: at most 64 ranks");
  TFS_CUDA(cudaMemsetAsync(counts, 0, size_t(world) * 8, stream));
  if (n <= 0) return 0;
  agg_owner_kernel<<<(unsigned)grid_for(n, dev), kThreads, 0, stream>>>(reinterpret_cast<const long long*>(keys), n, (unsigned int)world,
                                                                        owner, counts);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

int launch_gather_rows(const void* src, const unsigned int* perm, int64_t n, int64_t row_bytes, void* dst, const DeviceInfo& dev,
                       cudaStream_t stream) {
  if (n <= 0 || row_bytes <= 0) return 0;
  TFS_CHECK(row_bytes % 4 == 0, TFS_ERR_INVALID, "internal: rows are whole 4-byte words");
  agg_gather_rows_kernel<<<(unsigned)grid_for(n * (row_bytes / 4), dev), kThreads, 0, stream>>>(
      static_cast<const unsigned int*>(src), perm, n, row_bytes / 4, static_cast<unsigned int*>(dst));
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace tfs
