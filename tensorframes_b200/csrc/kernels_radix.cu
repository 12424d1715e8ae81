// Stable radix partition passes (sm_100a) -- the row-by-slot and group-by-key orderings of the keyed reduction (K3)
// and the by-owner partition of its multi-GPU exchange.
//
// One pass orders n (key, value) pairs by up to 11 bits of the key, stably; a sort is a sequence of passes from the
// low bits up (LSD).  What the callers need is narrow -- 32-bit slot numbers carrying a row index or a 4/8-byte value,
// 64-bit group keys carrying a slot number -- so the pass is written for exactly that:
//
//   * every CTA owns one contiguous range of the input (grid = 2 CTAs per SM).  A histogram kernel counts the digits of
//     each range, two small scan kernels turn the [bin][cta] table into global write positions, and the scatter kernel
//     walks its range tile by tile with its bin cursors in shared memory.  No atomics decide an output position, so the
//     pass is stable and its output is a pure function of its input: equal keys keep their input order, run after run.
//   * 2048 bins instead of a library sort's 256: 22 slot bits are two passes instead of three (9 GB of traffic per
//     2.5e8 rows where three 8-bit passes move 13 GB), the first pass takes its values from the element index (no
//     row-id column is read), the last pass writes no keys.
//   * inside a tile (4096 pairs) ranks come from __match_any_sync over the digit -- lanes with equal digits rank
//     themselves by lane order -- on per-warp counters, then one pass over the 8 x 2048 counters gives every warp's
//     offset inside every bin.  Pairs are staged in shared memory in output order, so the global writes of a tile are
//     runs of consecutive addresses.
#include <cstdlib>

#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;                       // pairs per thread per tile
constexpr int kTile = kThreads * kItems;         // 4096
constexpr int kMaxBits = 11;
constexpr int kMaxBins = 1 << kMaxBits;
constexpr int kCtasPerSm = 2;

template <class K>
__device__ __forceinline__ unsigned int digit_of(K key, K xor_mask, int shift, unsigned int mask) {
  return (unsigned int)((key ^ xor_mask) >> shift) & mask;
}

// hist[bin * grid + cta] = number of keys of CTA `cta`'s range whose digit is `bin`
template <class K>
__global__ void __launch_bounds__(kThreads)
radix_hist_kernel(const K* __restrict__ keys, const int64_t n, const int64_t per_cta, const K xor_mask, const int shift,
                  const int bits, unsigned int* __restrict__ hist) {
  __shared__ unsigned int sm[kMaxBins];
  const int bins = 1 << bits;
  const unsigned int mask = (unsigned int)bins - 1u;
  for (int b = threadIdx.x; b < bins; b += kThreads) sm[b] = 0u;
  __syncthreads();
  const int64_t lo = int64_t(blockIdx.x) * per_cta;
  const int64_t hi = lo + per_cta < n ? lo + per_cta : n;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) atomicAdd(&sm[digit_of<K>(__ldg(keys + i), xor_mask, shift, mask)], 1u);
  __syncthreads();
  for (int b = threadIdx.x; b < bins; b += kThreads) hist[int64_t(b) * gridDim.x + blockIdx.x] = sm[b];
}

// one warp per bin: exclusive prefix over the CTAs (in place) and the bin total
__global__ void __launch_bounds__(kThreads)
radix_scan_bins_kernel(unsigned int* __restrict__ hist, const int grid, const int bins, unsigned int* __restrict__ bin_total) {
  const int warp = (blockIdx.x * kThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= bins) return;
  unsigned int* row = hist + int64_t(warp) * grid;
  unsigned int carry = 0;
  for (int base = 0; base < grid; base += 32) {
    const int i = base + lane;
    const unsigned int v = i < grid ? row[i] : 0u;
    unsigned int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (i < grid) row[i] = carry + incl - v;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) bin_total[warp] = carry;
}

// single CTA: exclusive scan of the bin totals (bins <= 2048) in place
__global__ void __launch_bounds__(1024)
radix_scan_totals_kernel(unsigned int* __restrict__ bin_total, const int bins) {
  __shared__ unsigned int warp_sums[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i0 = threadIdx.x * 2;
  const unsigned int a = i0 < bins ? bin_total[i0] : 0u, b = i0 + 1 < bins ? bin_total[i0 + 1] : 0u;
  const unsigned int v = a + b;
  unsigned int incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const unsigned int w = warp_sums[lane];
    unsigned int wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += t;
    }
    warp_sums[lane] = wi - w;
  }
  __syncthreads();
  const unsigned int excl = warp_sums[warp] + incl - v;
  if (i0 < bins) bin_total[i0] = excl;
  if (i0 + 1 < bins) bin_total[i0 + 1] = excl + a;
}

// Dynamic shared memory of the scatter kernel:
//   cursor[bins] u32 | bin_start[bins + 1] u32 | cnt[kWarps][bins] u16 | staged keys [kTile] | staged values [kTile]
template <class K, class V>
constexpr size_t scatter_smem_bytes() {
  return size_t(kMaxBins) * 4 + size_t(kMaxBins + 8) * 4 + size_t(kWarps) * kMaxBins * 2 + size_t(kTile) * (sizeof(K) + sizeof(V));
}

template <class K, class V, bool IOTA>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
radix_scatter_kernel(const K* __restrict__ keys_in, const V* __restrict__ vals_in, K* __restrict__ keys_out, V* __restrict__ vals_out,
                     const int64_t n, const int64_t per_cta, const K xor_mask, const int shift, const int bits,
                     const unsigned int* __restrict__ hist, const unsigned int* __restrict__ bin_base) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int bins = 1 << bits;
  const unsigned int mask = (unsigned int)bins - 1u;
  unsigned int* cursor = reinterpret_cast<unsigned int*>(smem_raw);
  unsigned int* bin_start = cursor + kMaxBins;
  unsigned short* cnt = reinterpret_cast<unsigned short*>(bin_start + kMaxBins + 8);
  K* st_keys = reinterpret_cast<K*>(cnt + kWarps * kMaxBins);
  V* st_vals = reinterpret_cast<V*>(st_keys + kTile);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int lt_mask = (1u << lane) - 1u;
  for (int b = threadIdx.x; b < bins; b += kThreads)
    cursor[b] = bin_base[b] + hist[int64_t(b) * gridDim.x + blockIdx.x];
  const int64_t lo = int64_t(blockIdx.x) * per_cta;
  const int64_t hi = lo + per_cta < n ? lo + per_cta : n;
  unsigned short* my_cnt = cnt + warp * bins;

  for (int64_t tile = lo; tile < hi; tile += kTile) {
    // (a) counters to zero
    {
      unsigned int* z = reinterpret_cast<unsigned int*>(cnt);
      for (int i = threadIdx.x; i < kWarps * bins / 2; i += kThreads) z[i] = 0u;
    }
    __syncthreads();       // also: the previous tile's write-out and cursor update are done
    // (b) load, rank inside the warp.  Warp w owns items [w * 512, w * 512 + 512) of the tile, round r = 32 of them.
    K key[kItems];
    V val[kItems];
    unsigned short lrank[kItems];
    const int64_t wbase = tile + int64_t(warp) * (32 * kItems) + lane;
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const int64_t i = wbase + r * 32;
      key[r] = i < hi ? __ldcs(keys_in + i) : K(0);
      if constexpr (IOTA) val[r] = V(i);
      else val[r] = i < hi ? __ldcs(vals_in + i) : V(0);
    }
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const bool valid = wbase + r * 32 < hi;
      const unsigned int d = valid ? digit_of<K>(key[r], xor_mask, shift, mask) : 0xffffffffu;
      const unsigned int peers = __match_any_sync(0xffffffffu, d);
      const unsigned int before = __popc(peers & lt_mask);
      unsigned int base = 0;
      if (valid) base = my_cnt[d];
      __syncwarp();
      if (valid && before == 0u) my_cnt[d] = (unsigned short)(base + __popc(peers));
      __syncwarp();
      lrank[r] = (unsigned short)(base + before);
    }
    __syncthreads();
    // (c) per bin: each warp's offset inside the bin (in place) and the bin's start inside the tile
    {
      const int per_thread = (bins + kThreads - 1) / kThreads;       // 8 at 2048 bins
      const int b0 = threadIdx.x * per_thread;
      unsigned int tot[kMaxBins / kThreads];
      unsigned int mine = 0;
#pragma unroll
      for (int j = 0; j < kMaxBins / kThreads; ++j) {
        tot[j] = 0;
        const int b = b0 + j;
        if (j < per_thread && b < bins) {
          unsigned int run = 0;
#pragma unroll
          for (int w = 0; w < kWarps; ++w) {
            const unsigned int c = cnt[w * bins + b];
            cnt[w * bins + b] = (unsigned short)run;
            run += c;
          }
          tot[j] = run;
          mine += run;
        }
      }
      // exclusive scan of `mine` over the CTA
      unsigned int incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      __shared__ unsigned int warp_tot[kWarps];
      if (lane == 31) warp_tot[warp] = incl;
      __syncthreads();
      unsigned int wofs = 0;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) wofs += w < warp ? warp_tot[w] : 0u;
      unsigned int run = wofs + incl - mine;
#pragma unroll
      for (int j = 0; j < kMaxBins / kThreads; ++j) {
        const int b = b0 + j;
        if (j < per_thread && b < bins) {
          bin_start[b] = run;
          run += tot[j];
        }
      }
      if (threadIdx.x == kThreads - 1) bin_start[bins] = run;      // = pairs in this tile
    }
    __syncthreads();
    // (d) stage the pairs in output order
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      if (wbase + r * 32 < hi) {
        const unsigned int d = digit_of<K>(key[r], xor_mask, shift, mask);
        const unsigned int pos = bin_start[d] + my_cnt[d] + lrank[r];
        st_keys[pos] = key[r];
        st_vals[pos] = val[r];
      }
    }
    __syncthreads();
    // (e) write out: position p of the tile goes to cursor[digit] + (p - bin_start[digit])
    const unsigned int in_tile = bin_start[bins];
#pragma unroll 4
    for (int j = 0; j < kItems; ++j) {
      const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
      if (p < in_tile) {
        const K k = st_keys[p];
        const unsigned int d = digit_of<K>(k, xor_mask, shift, mask);
        const unsigned int g = cursor[d] + (p - bin_start[d]);
        if (keys_out != nullptr) keys_out[g] = k;
        vals_out[g] = st_vals[p];
      }
    }
    __syncthreads();
    // (f) advance the cursors
    for (int b = threadIdx.x; b < bins; b += kThreads) cursor[b] += bin_start[b + 1] - bin_start[b];
    // (the next tile's first __syncthreads orders this against its use)
  }
}

struct PassGeom { int grid; int64_t per_cta; };
PassGeom pass_geometry(int64_t n, const DeviceInfo& dev) {
  PassGeom g;
  g.grid = dev.sm_count * kCtasPerSm;
  int64_t per = (n + g.grid - 1) / g.grid;
  per = (per + kTile - 1) / kTile * kTile;
  g.per_cta = per < kTile ? kTile : per;
  return g;
}

template <class K, class V, bool IOTA>
int launch_pass_t(const K* keys_in, const V* vals_in, K* keys_out, V* vals_out, int64_t n, K xor_mask, int shift, int bits,
                  void* scratch, const DeviceInfo& dev, cudaStream_t stream) {
  const PassGeom g = pass_geometry(n, dev);
  const int bins = 1 << bits;
  unsigned int* hist = static_cast<unsigned int*>(scratch);
  unsigned int* bin_total = hist + size_t(kMaxBins) * size_t(g.grid);
  radix_hist_kernel<K><<<g.grid, kThreads, 0, stream>>>(keys_in, n, g.per_cta, xor_mask, shift, bits, hist);
  TFS_CUDA(cudaGetLastError());
  radix_scan_bins_kernel<<<(bins * 32 + kThreads - 1) / kThreads, kThreads, 0, stream>>>(hist, g.grid, bins, bin_total);
  TFS_CUDA(cudaGetLastError());
  radix_scan_totals_kernel<<<1, 1024, 0, stream>>>(bin_total, bins);
  TFS_CUDA(cudaGetLastError());
  constexpr size_t smem = scatter_smem_bytes<K, V>();
  static PerDeviceOnce attr;
  attr.run(dev.device, [] {
    TFS_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<K, V, IOTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  });
  radix_scatter_kernel<K, V, IOTA><<<g.grid, kThreads, smem, stream>>>(keys_in, vals_in, keys_out, vals_out, n, g.per_cta, xor_mask,
                                                                       shift, bits, hist, bin_total);
  TFS_CUDA(cudaGetLastError());
  return 4;
}

}  // namespace

size_t radix_scratch_bytes(const DeviceInfo& dev) {
  return (size_t(kMaxBins) * size_t(dev.sm_count * kCtasPerSm) + kMaxBins + 64) * sizeof(unsigned int);
}

int radix_passes(int bits) { return bits <= 0 ? 0 : (bits + kMaxBits - 1) / kMaxBits; }

int launch_radix_pass(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_out, void* vals_out,
                      int64_t n, uint64_t xor_mask, int shift, int bits, void* scratch, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  TFS_CHECK(bits >= 1 && bits <= kMaxBits, TFS_ERR_INVALID, "internal: radix pass of " + std::to_string(bits) + " bits");
  TFS_CHECK(n < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "internal: radix pass over 2^32 or more pairs");
  using u32 = unsigned int;
  using u64 = unsigned long long;
  if (key_bytes == 4 && val_bytes == 4) {
    if (vals_in == nullptr)
      return launch_pass_t<u32, u32, true>(static_cast<const u32*>(keys_in), nullptr, static_cast<u32*>(keys_out), static_cast<u32*>(vals_out),
                                           n, u32(xor_mask), shift, bits, scratch, dev, stream);
    return launch_pass_t<u32, u32, false>(static_cast<const u32*>(keys_in), static_cast<const u32*>(vals_in), static_cast<u32*>(keys_out),
                                          static_cast<u32*>(vals_out), n, u32(xor_mask), shift, bits, scratch, dev, stream);
  }
  TFS_CHECK(vals_in != nullptr, TFS_ERR_INVALID, "internal: index values are 4 bytes wide");
  if (key_bytes == 4 && val_bytes == 8)
    return launch_pass_t<u32, u64, false>(static_cast<const u32*>(keys_in), static_cast<const u64*>(vals_in), static_cast<u32*>(keys_out),
                                          static_cast<u64*>(vals_out), n, u32(xor_mask), shift, bits, scratch, dev, stream);
  if (key_bytes == 8 && val_bytes == 4)
    return launch_pass_t<u64, u32, false>(static_cast<const u64*>(keys_in), static_cast<const u32*>(vals_in), static_cast<u64*>(keys_out),
                                          static_cast<u32*>(vals_out), n, u64(xor_mask), shift, bits, scratch, dev, stream);
  fail(TFS_ERR_INVALID, "internal: radix pass for this key / value width is not built");
}

// Stable sort of n pairs on key bits [0, bits): passes from the low bits up, reading `keys_in / vals_in` (vals_in null:
// the values are the element indices) and alternating between the two buffer pairs.  The last pass writes its keys only
// when `want_keys`.  Returns the index (0 / 1) of the buffer pair that holds the result; *launches is advanced.
int launch_radix_sort(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_buf[2], void* vals_buf[2],
                      int64_t n, uint64_t xor_mask, int bits, bool want_keys, void* scratch, const DeviceInfo& dev, cudaStream_t stream,
                      int* launches) {
  const int passes = radix_passes(bits);
  TFS_CHECK(passes >= 1, TFS_ERR_INVALID, "internal: radix sort over no bits");
  int shift = 0, into = 0;
  const void* kin = keys_in;
  const void* vin = vals_in;
  for (int p = 0; p < passes; ++p) {
    const int left = bits - shift;
    const int b = (left + (passes - p) - 1) / (passes - p);         // spread the bits evenly over the remaining passes
    const bool last = p == passes - 1;
    *launches += launch_radix_pass(key_bytes, val_bytes, kin, vin, (last && !want_keys) ? nullptr : keys_buf[into], vals_buf[into], n,
                                   xor_mask, shift, b, scratch, dev, stream);
    kin = keys_buf[into];
    vin = vals_buf[into];
    shift += b;
    into ^= 1;
  }
  return into ^ 1;
}

}  // namespace tfs
