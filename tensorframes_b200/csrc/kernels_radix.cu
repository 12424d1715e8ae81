// Stable radix partition passes (sm_100a) -- the row-by-slot and group-by-key orderings of the keyed reduction (K3)
// and the by-owner partition of its multi-GPU exchange.
//
// One pass orders n (key, value) pairs by up to 11 bits of the key, stably; a sort is a sequence of passes from the
// low bits up (LSD).  What the callers need is narrow -- 32-bit slot numbers carrying a row index or a 4/8-byte value,
// 64-bit group keys carrying a slot number -- so the pass is written for exactly that:
//
//   * every CTA owns one contiguous range of the input.  A histogram kernel counts the digits of each range, two small
//     scan kernels turn the [bin][cta] table into global write positions, and the scatter kernel walks its range tile
//     by tile with its bin cursors in shared memory.  No atomics decide an output position, so the pass is stable and
//     its output is a pure function of its input: equal keys keep their input order, run after run.
//   * the first pass takes its values from the element index (no row-id column is read), the last pass writes no keys.
//   * inside a tile (4096 pairs) a pair's rank is its warp's running count of the digit plus the number of lower lanes
//     that hold the same digit in the same round.  That lane set is collected in shared memory in the 256-bin
//     instantiation (every lane ORs its bit into its warp's word for the digit and reads the word back) and by one
//     ballot per digit bit in the 2048-bin one; MATCH.ANY, the one-instruction way, retires once per 60 cycles per SM
//     here.  One pass over the per-warp counters then gives every warp's offset inside every bin.  Pairs are staged in
//     shared memory in output order, so the global writes of a tile are runs of consecutive addresses; the next tile's
//     keys are loaded while the current one is written out.
//   * passes are 8 bits wide by default (three for 23 slot bits): with 2048 bins a tile holds two pairs per bin, every
//     write is a partial sector, and the open lines of 2 columns x 2048 bins x 296 CTAs do not fit L2 -- measured
//     13 GB of DRAM traffic for 4 GB of data (profiles/r02_experiments.md).  Passes of 9 to 11 bits remain available
//     (TFS_RADIX_BITS) and are used where a partition has that many bins by nature.
#include <cstdlib>
#include <type_traits>

#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;                       // pairs per thread per tile
constexpr int kTile = kThreads * kItems;         // 4096
constexpr int kMaxBits = 11;
constexpr int kMaxBins = 1 << kMaxBits;
constexpr int kCtasPerSm = 2;          // passes of 9 to 11 bits
constexpr int kSmallBits = 8;          // passes of at most 8 bits: 256 bins, small tables
constexpr int kCtasPerSmSmall = 2;   // (three were measured: no faster, and the packed modes spill at 80 registers)
constexpr int ctas_for(int mb) { return mb <= kSmallBits ? kCtasPerSmSmall : kCtasPerSm; }

template <class K>
__device__ __forceinline__ unsigned int digit_of(K key, K xor_mask, int shift, unsigned int mask) {
  return (unsigned int)((key ^ xor_mask) >> shift) & mask;
}

// hist[bin * grid + cta] = number of keys of CTA `cta`'s range whose digit is `bin`.  Same ranges as the scatter kernel
// (two CTAs per SM), so the CTAs are 512 threads wide: 256 threads x 8 loads in flight per thread left the read at
// 5.1 TB/s.
constexpr int kHistThreads = 512;
template <class K>
__global__ void __launch_bounds__(kHistThreads)
radix_hist_kernel(const K* __restrict__ keys, const int64_t n, const int64_t per_cta, const K xor_mask, const int shift,
                  const int bits, unsigned int* __restrict__ hist) {
  __shared__ unsigned int sm[kMaxBins];
  const int bins = 1 << bits;
  const unsigned int mask = (unsigned int)bins - 1u;
  for (int b = threadIdx.x; b < bins; b += kHistThreads) sm[b] = 0u;
  __syncthreads();
  const int64_t lo = int64_t(blockIdx.x) * per_cta;
  const int64_t hi = lo + per_cta < n ? lo + per_cta : n;
  int64_t i = lo + threadIdx.x;
  for (; i + 7 * kHistThreads < hi; i += 8 * kHistThreads) {      // eight loads in flight per thread
    K k[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) k[u] = __ldcs(keys + i + u * kHistThreads);
#pragma unroll
    for (int u = 0; u < 8; ++u) atomicAdd(&sm[digit_of<K>(k[u], xor_mask, shift, mask)], 1u);
  }
  for (; i < hi; i += kHistThreads) atomicAdd(&sm[digit_of<K>(__ldcs(keys + i), xor_mask, shift, mask)], 1u);
  __syncthreads();
  for (int b = threadIdx.x; b < bins; b += kHistThreads) hist[int64_t(b) * gridDim.x + blockIdx.x] = sm[b];
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
//   cursor[bins] u32 | bin_start[bins + 1] u32 | shift_of[bins] u32 | cnt[kWarps][bins] u16 | staged keys | staged values
//   | (256-bin instantiation) lane-mask tables [2][kWarps][bins] u32
template <class K, class V, int MB>
constexpr size_t scatter_smem_bytes() {
  return size_t(1 << MB) * 4 + size_t((1 << MB) + 8) * 4 + size_t(1 << MB) * 4 + size_t(kWarps) * (1 << MB) * 2 +
         size_t(kTile) * (sizeof(K) + sizeof(V)) + (MB <= kSmallBits ? size_t(2) * kWarps * (1 << MB) * 4 : 0);
}

// MODE: M_PLAIN = values loaded from vals_in; M_IOTA = the values are the element indices (4-byte values).
// The packed modes (4-byte keys and values) keep a pair in ONE 8-byte word, key << 32 | value, between the passes of a
// sort -- one load and one store per pair instead of two, and one write stream per bin instead of two:
// M_IOTA_PACK = as M_IOTA, written packed through keys_out; M_PACKED = packed in (through keys_in), packed out;
// M_UNPACK = packed in, values out; M_UNPACK_BOUNDS = as M_UNPACK, and for a sort's LAST pass on whole keys: the first
// and one-past-the-last output position of every key's run go to seg_lo[key] / seg_hi[key] (atomicMin / atomicMax at the
// run boundaries of a tile: the pairs of a bin reach the last pass ordered by their low bits, so equal keys are
// neighbours in the staged tile).
enum : int { M_PLAIN = 0, M_IOTA = 1, M_IOTA_PACK = 2, M_PACKED = 3, M_UNPACK = 4, M_UNPACK_BOUNDS = 5 };

template <class K, class V, int MODE, int MB>
__global__ void __launch_bounds__(kThreads, ctas_for(MB))
radix_scatter_kernel(const K* __restrict__ keys_in, const V* __restrict__ vals_in, K* __restrict__ keys_out, V* __restrict__ vals_out,
                     const int64_t n, const int64_t per_cta, const K xor_mask, const int shift, const int bits,
                     const unsigned int* __restrict__ hist, const unsigned int* __restrict__ bin_base,
                     unsigned int* __restrict__ seg_lo, unsigned int* __restrict__ seg_hi) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int bins = 1 << bits;
  const unsigned int mask = (unsigned int)bins - 1u;
  constexpr int kMB = 1 << MB;                                        // bins this instantiation has room for
  constexpr int kPT = kMB / kThreads >= 1 ? kMB / kThreads : 1;       // bins per thread in the per-bin phase
  unsigned int* cursor = reinterpret_cast<unsigned int*>(smem_raw);
  unsigned int* bin_start = cursor + kMB;
  unsigned int* shift_of = bin_start + kMB + 8;
  unsigned short* cnt = reinterpret_cast<unsigned short*>(shift_of + kMB);
  K* st_keys = reinterpret_cast<K*>(cnt + kWarps * kMB);
  V* st_vals = reinterpret_cast<V*>(st_keys + kTile);
  constexpr bool kPair64 = sizeof(K) == 4 && sizeof(V) == 4;      // staged as one word: key << 32 | value
  [[maybe_unused]] unsigned long long* st_pairs = reinterpret_cast<unsigned long long*>(st_keys);
  [[maybe_unused]] unsigned int* peer_tab = reinterpret_cast<unsigned int*>(st_vals + kTile);
  if constexpr (MB <= kSmallBits) {
    for (int i = threadIdx.x; i < 2 * kWarps * kMB; i += kThreads) peer_tab[i] = 0u;      // (ordered by the first tile's barrier)
  }

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int lt_mask = (1u << lane) - 1u;
  for (int b = threadIdx.x; b < bins; b += kThreads)
    cursor[b] = bin_base[b] + hist[int64_t(b) * gridDim.x + blockIdx.x];
  const int64_t lo = int64_t(blockIdx.x) * per_cta;
  const int64_t hi = lo + per_cta < n ? lo + per_cta : n;
  unsigned short* my_cnt = cnt + warp * bins;

  // Every phase below first issues all of its loads and only then its stores: with 8 warps per CTA the kernel lives on
  // instruction-level parallelism, and a store between two loads of arrays the compiler cannot tell apart would chain
  // them (measured: 30 000 cycles per tile before, latency-bound on shared memory and on the tile's global loads).
  constexpr bool kPackedIn = MODE == M_PACKED || MODE == M_UNPACK || MODE == M_UNPACK_BOUNDS;
  constexpr bool kPackedOut = MODE == M_IOTA_PACK || MODE == M_PACKED;
  using Raw = typename std::conditional<kPackedIn, unsigned long long, K>::type;
  const Raw* raw_in = reinterpret_cast<const Raw*>(keys_in);
  Raw raw[kItems];
  // a thread's items of a tile sit 32 apart from its first one: constant offsets from one pointer; `left` = items of
  // the range from that first one on (only a range's last tile can be short)
  auto load_raw = [&](int64_t tile_base, bool whole) {
    const int64_t first = tile_base + warp * (32 * kItems) + lane;
    const Raw* p = raw_in + first;
    if (whole) {
#pragma unroll
      for (int r = 0; r < kItems; ++r) raw[r] = __ldcs(p + r * 32);
    } else {
      const int64_t left64 = hi - first;
      const int left = left64 > 32 * kItems ? 32 * kItems : (left64 < 0 ? 0 : int(left64));
#pragma unroll
      for (int r = 0; r < kItems; ++r) raw[r] = r * 32 < left ? __ldcs(p + r * 32) : Raw(0);
    }
  };
  if (lo < hi) load_raw(lo, lo + kTile <= hi);
  // The tile body exists twice: for whole tiles (all but possibly the last of a range) nothing is predicated.
  auto tile_body = [&](const int64_t tile, auto whole_tag) {
    constexpr bool WHOLE = decltype(whole_tag)::value;
    // (a) counters to zero
    {
      unsigned int* z = reinterpret_cast<unsigned int*>(cnt);
      for (int i = threadIdx.x; i < kWarps * bins / 2; i += kThreads) z[i] = 0u;
    }
    const int64_t wbase = tile + warp * (32 * kItems) + lane;
    int left = 32 * kItems;
    if constexpr (!WHOLE) {
      const int64_t left64 = hi - wbase;
      left = left64 > 32 * kItems ? 32 * kItems : (left64 < 0 ? 0 : int(left64));
    }
    // the values are not needed before the staging phase: their loads travel while the tile is ranked
    K key[kItems];
    V val[kItems];
    [[maybe_unused]] const unsigned int iota0 = (unsigned int)wbase;
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      if constexpr (kPackedIn) {
        key[r] = K(raw[r] >> 32);
        val[r] = V(raw[r]);
      } else {
        key[r] = K(raw[r]);
        if constexpr (MODE == M_PLAIN) val[r] = (WHOLE || r * 32 < left) ? __ldcs(vals_in + wbase + r * 32) : V(0);
        else val[r] = V(iota0 + (unsigned int)(r * 32));
      }
    }
    __syncthreads();       // also: the previous tile's write-out and cursor update are done
    // (b) rank inside the warp.  Warp w owns items [w * 512, w * 512 + 512) of the tile, round r = 32 of them.
    unsigned short lrank[kItems];
    unsigned int dig[kItems];
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const bool valid = WHOLE || r * 32 < left;
      dig[r] = valid ? digit_of<K>(key[r], xor_mask, shift, mask) : 0xffffffffu;
    }
    // peers = the lanes that hold the same digit.  __match_any_sync is one instruction, but MATCH.ANY retires once per
    // 60 cycles per SM on this GPU whatever the occupancy -- 1.6 of the 1.9 ms an 8-bit pass over 2.5e8 pairs took with
    // it.  One ballot per digit bit with its mask: 29 cycles for 8 bits.  Lane masks collected in shared memory (every
    // lane ORs its bit into its warp's word for the digit, then reads the word back): 10 cycles
    // (profiles/experiments/match_rate.cu).  The 256-bin instantiation has room for the tables: two per warp, used in
    // turn, so that clearing the words of a round waits one round and costs no extra warp barrier.
    if constexpr (MB <= kSmallBits) {
      unsigned int* const tab0 = peer_tab + warp * kMB;
      unsigned int* const tab1 = peer_tab + (kWarps + warp) * kMB;
      const unsigned int lane_bit = 1u << lane;
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        unsigned int* const cur = (r & 1) ? tab1 : tab0;
        unsigned int* const old = (r & 1) ? tab0 : tab1;
        const bool valid = WHOLE || dig[r] != 0xffffffffu;
        unsigned int base = 0;
        if (valid) {
          atomicOr(cur + dig[r], lane_bit);
          base = my_cnt[dig[r]];
        }
        __syncwarp();
        unsigned int peers = 0;
        if (valid) peers = cur[dig[r]];
        if (r > 0 && (WHOLE || dig[r > 0 ? r - 1 : 0] != 0xffffffffu)) old[dig[r > 0 ? r - 1 : 0]] = 0u;
        const unsigned int before = __popc(peers & lt_mask);
        if (valid && before == 0u) my_cnt[dig[r]] = (unsigned short)(base + __popc(peers));
        __syncwarp();
        lrank[r] = (unsigned short)(base + before);
      }
      if (WHOLE || dig[kItems - 1] != 0xffffffffu) (((kItems - 1) & 1) ? tab1 : tab0)[dig[kItems - 1]] = 0u;
    } else {
      unsigned int peers[kItems];
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        unsigned int same = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < MB; ++b) {      // (bits of the digit above `bits` are zero in every lane: their ballots change nothing)
          asm volatile(
              "{\n\t.reg .pred p;\n\t.reg .b32 t, v;\n\t"
              "and.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
              "vote.sync.ballot.b32 v, p, 0xffffffff;\n\t"
              "selp.b32 t, -1, 0, p;\n\tlop3.b32 %0, %0, v, t, 0x90;\n\t}"      // same &= ~(v ^ t)
              : "+r"(same)
              : "r"(dig[r]), "r"(1u << b));
        }
        if constexpr (!WHOLE) {      // items past the end of the range only match one another
          const bool valid = dig[r] != 0xffffffffu;
          const unsigned int vote = __ballot_sync(0xffffffffu, valid);
          same &= valid ? vote : ~vote;
        }
        peers[r] = same;
      }
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        const bool valid = WHOLE || dig[r] != 0xffffffffu;
        const unsigned int before = __popc(peers[r] & lt_mask);
        unsigned int base = 0;
        if (valid) base = my_cnt[dig[r]];
        __syncwarp();
        if (valid && before == 0u) my_cnt[dig[r]] = (unsigned short)(base + __popc(peers[r]));
        __syncwarp();
        lrank[r] = (unsigned short)(base + before);
      }
    }
    __syncthreads();
    // (c) per bin: each warp's offset inside the bin (in place) and the bin's start inside the tile
    constexpr bool kFold = MB <= kSmallBits;      // one bin per thread
    {
      const int per_thread = (bins + kThreads - 1) / kThreads;       // 8 at 2048 bins
      const int b0 = threadIdx.x * per_thread;
      unsigned int tot[kPT];
      unsigned int mine = 0;
      [[maybe_unused]] unsigned int within[kWarps];      // (256-bin instantiation) this thread's bin: every warp's offset in it
      if constexpr (kFold) {
        tot[0] = 0;
        if (b0 < bins) {
#pragma unroll
          for (int w = 0; w < kWarps; ++w) within[w] = cnt[w * bins + b0];
          unsigned int run = 0;
#pragma unroll
          for (int w = 0; w < kWarps; ++w) {
            const unsigned int v = within[w];
            within[w] = run;
            run += v;
          }
          tot[0] = run;
          mine = run;
        }
      } else if (kPT == 8 && per_thread == 8) {
        // 8 neighbouring 16-bit counters of one warp row are one 16-byte word: 8 vector loads, then 8 vector stores
        uint4 c[kWarps];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) c[w] = *reinterpret_cast<const uint4*>(cnt + w * bins + b0);
        unsigned int run[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
          const unsigned int in[4] = {c[w].x, c[w].y, c[w].z, c[w].w};
          unsigned int out[4];
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const unsigned int lo16 = in[h] & 0xffffu, hi16 = in[h] >> 16;
            out[h] = run[2 * h] | (run[2 * h + 1] << 16);
            run[2 * h] += lo16;
            run[2 * h + 1] += hi16;
          }
          c[w] = make_uint4(out[0], out[1], out[2], out[3]);
        }
#pragma unroll
        for (int w = 0; w < kWarps; ++w) *reinterpret_cast<uint4*>(cnt + w * bins + b0) = c[w];
#pragma unroll
        for (int j = 0; j < kPT; ++j) { tot[j] = run[j % 8]; mine += run[j % 8]; }
      } else {
#pragma unroll
        for (int j = 0; j < kPT; ++j) {
          tot[j] = 0;
          const int b = b0 + j;
          if (j < per_thread && b < bins) {
            unsigned int c[kWarps];
#pragma unroll
            for (int w = 0; w < kWarps; ++w) c[w] = cnt[w * bins + b];
            unsigned int run = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
              const unsigned int v = c[w];
              c[w] = run;
              run += v;
            }
#pragma unroll
            for (int w = 0; w < kWarps; ++w) cnt[w * bins + b] = (unsigned short)c[w];
            tot[j] = run;
            mine += run;
          }
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
      // bin_start[b] = first position of bin b inside the tile; shift[b] = where that position goes in the output
      unsigned int cur[kPT];
#pragma unroll
      for (int j = 0; j < kPT; ++j) cur[j] = (j < per_thread && b0 + j < bins) ? cursor[b0 + j] : 0u;
#pragma unroll
      for (int j = 0; j < kPT; ++j) {
        const int b = b0 + j;
        if (j < per_thread && b < bins) {
          if constexpr (kFold) {
            // the bin's start goes into the per-warp offsets themselves (at most 4096: fits the 16-bit counters), so
            // the staging phase looks up one word per pair instead of two
#pragma unroll
            for (int w = 0; w < kWarps; ++w) cnt[w * bins + b] = (unsigned short)(run + within[w]);
          } else {
            bin_start[b] = run;
          }
          shift_of[b] = cur[j] - run;
          cursor[b] = cur[j] + tot[j];      // (read again only in the next tile's phase (c))
          run += tot[j];
        }
      }
      if (threadIdx.x == kThreads - 1) bin_start[bins] = run;      // = pairs in this tile
    }
    __syncthreads();
    // (d) stage the pairs in output order: all positions first, then the stores
    {
      unsigned int pos[kItems];
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        const unsigned int d = (WHOLE || dig[r] != 0xffffffffu) ? dig[r] : 0u;
        if constexpr (kFold) pos[r] = my_cnt[d] + lrank[r];
        else pos[r] = bin_start[d] + my_cnt[d] + lrank[r];
      }
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        if (WHOLE || dig[r] != 0xffffffffu) {
          if constexpr (kPair64) {      // one 8-byte store per pair instead of two scattered 4-byte ones
            st_pairs[pos[r]] = ((unsigned long long)key[r] << 32) | (unsigned long long)val[r];
          } else {
            st_keys[pos[r]] = key[r];
            st_vals[pos[r]] = val[r];
          }
        }
      }
    }
    // the next tile's keys start their trip now: the write-out below and the next tile's first phase cover the latency
    if (tile + kTile < hi) load_raw(tile + kTile, tile + 2 * kTile <= hi);
    __syncthreads();
    // (e) write out: position p of the tile goes to shift_of[digit] + p
    {
      const unsigned int in_tile = bin_start[bins];
      K k[kItems];
      V v[kItems];
#pragma unroll
      for (int j = 0; j < kItems; ++j) {
        const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
        if constexpr (kPair64) {
          const unsigned long long w = (WHOLE || p < in_tile) ? st_pairs[p] : 0ULL;
          k[j] = K(w >> 32);
          v[j] = V(w);
        } else {
          k[j] = (WHOLE || p < in_tile) ? st_keys[p] : K(0);
          v[j] = (WHOLE || p < in_tile) ? st_vals[p] : V(0);
        }
      }
      unsigned int g[kItems];
#pragma unroll
      for (int j = 0; j < kItems; ++j) {
        const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
        g[j] = shift_of[digit_of<K>(k[j], xor_mask, shift, mask)] + p;
      }
      if constexpr (MODE == M_UNPACK_BOUNDS) {
        // run boundaries inside the staged tile (a key whose run crosses tiles or CTAs gets several candidates)
        K prev[kItems], next[kItems];
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
          const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
          // (M_UNPACK_BOUNDS is a mode of 4-byte keys and values: the staged key is the high half of the pair's word)
          const unsigned int* st_hi = reinterpret_cast<const unsigned int*>(st_pairs) + 1;
          prev[j] = (p > 0u && (WHOLE || p < in_tile)) ? K(st_hi[2 * (p - 1u)]) : ~k[j];
          next[j] = p + 1u < in_tile ? K(st_hi[2 * (p + 1u)]) : ~k[j];
        }
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
          const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
          if (WHOLE || p < in_tile) {
            if (prev[j] != k[j]) atomicMin(seg_lo + k[j], g[j]);
            if (next[j] != k[j]) atomicMax(seg_hi + k[j], g[j] + 1u);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kItems; ++j) {
        const unsigned int p = (unsigned int)(j * kThreads + threadIdx.x);
        if (WHOLE || p < in_tile) {
          if constexpr (kPackedOut) {
            reinterpret_cast<unsigned long long*>(keys_out)[g[j]] = ((unsigned long long)k[j] << 32) | (unsigned long long)v[j];
          } else {
            if (keys_out != nullptr) keys_out[g[j]] = k[j];
            vals_out[g[j]] = v[j];
          }
        }
      }
    }
    // (the next tile's first __syncthreads orders the staged pairs and shift_of against their next use)
  };
  for (int64_t tile = lo; tile < hi; tile += kTile) {
    if (tile + kTile <= hi) tile_body(tile, std::true_type{});
    else tile_body(tile, std::false_type{});
  }
}

struct PassGeom { int grid; int64_t per_cta; };
PassGeom pass_geometry(int64_t n, int bits, const DeviceInfo& dev) {
  PassGeom g;
  g.grid = dev.sm_count * ctas_for(bits);
  int64_t per = (n + g.grid - 1) / g.grid;
  per = (per + kTile - 1) / kTile * kTile;
  g.per_cta = per < kTile ? kTile : per;
  return g;
}

template <class K, class V, int MODE>
int launch_pass_t(const K* keys_in, const V* vals_in, K* keys_out, V* vals_out, int64_t n, K xor_mask, int shift, int bits,
                  void* scratch, const DeviceInfo& dev, cudaStream_t stream, unsigned int* seg_lo = nullptr, unsigned int* seg_hi = nullptr) {
  const PassGeom g = pass_geometry(n, bits, dev);
  const int bins = 1 << bits;
  unsigned int* hist = static_cast<unsigned int*>(scratch);
  unsigned int* bin_total = hist + size_t(bins) * size_t(g.grid);
  if constexpr (MODE == M_PACKED || MODE == M_UNPACK || MODE == M_UNPACK_BOUNDS)
    radix_hist_kernel<unsigned long long><<<g.grid, kHistThreads, 0, stream>>>(reinterpret_cast<const unsigned long long*>(keys_in), n, g.per_cta,
                                                                          0ULL, shift + 32, bits, hist);
  else
    radix_hist_kernel<K><<<g.grid, kHistThreads, 0, stream>>>(keys_in, n, g.per_cta, xor_mask, shift, bits, hist);
  TFS_CUDA(cudaGetLastError());
  radix_scan_bins_kernel<<<(bins * 32 + kThreads - 1) / kThreads, kThreads, 0, stream>>>(hist, g.grid, bins, bin_total);
  TFS_CUDA(cudaGetLastError());
  radix_scan_totals_kernel<<<1, 1024, 0, stream>>>(bin_total, bins);
  TFS_CUDA(cudaGetLastError());
  if (bits <= kSmallBits) {
    constexpr size_t smem = scatter_smem_bytes<K, V, kSmallBits>();
    static PerDeviceOnce attr;
    attr.run(dev.device, [] {
      TFS_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<K, V, MODE, kSmallBits>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    });
    radix_scatter_kernel<K, V, MODE, kSmallBits><<<g.grid, kThreads, smem, stream>>>(keys_in, vals_in, keys_out, vals_out, n, g.per_cta,
                                                                                      xor_mask, shift, bits, hist, bin_total, seg_lo, seg_hi);
  } else {
    constexpr size_t smem = scatter_smem_bytes<K, V, kMaxBits>();
    static PerDeviceOnce attr;
    attr.run(dev.device, [] {
      TFS_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<K, V, MODE, kMaxBits>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    });
    radix_scatter_kernel<K, V, MODE, kMaxBits><<<g.grid, kThreads, smem, stream>>>(keys_in, vals_in, keys_out, vals_out, n, g.per_cta,
                                                                                    xor_mask, shift, bits, hist, bin_total, seg_lo, seg_hi);
  }
  TFS_CUDA(cudaGetLastError());
  return 4;
}

}  // namespace

size_t radix_scratch_bytes(const DeviceInfo& dev) {
  return (size_t(kMaxBins) * size_t(dev.sm_count * kCtasPerSmSmall) + kMaxBins + 64) * sizeof(unsigned int);
}

namespace {
// Bits per pass of the generic (two-column) passes: 8 unless TFS_RADIX_BITS says otherwise.  Measured on 2.5e8
// (slot, row) pairs: an 11-bit pass writing two columns moves 13 GB through DRAM for 4 GB of data (its 2 x 2048 x 296
// open cache lines do not fit L2) and takes 8.8 ms; an 8-bit pass stays at its 4 GB and takes 2.1 ms.  The packed
// two-pass path of launch_radix_rows_by_slot does not go through here.
int pass_bits() {
  static const int v = [] {
    const char* e = std::getenv("TFS_RADIX_BITS");
    const int b = e ? std::atoi(e) : kSmallBits;
    return b >= 1 && b <= kMaxBits ? b : kSmallBits;
  }();
  return v;
}
}  // namespace

// Small inputs (up to 4M pairs: the group sort of K3, the merge of partial groups) stay in L2 whatever the fan-out, so they
// take 11-bit passes: six instead of eight for 64-bit keys, and their cost is launches, not bytes.
int bits_per_pass(int64_t n) { return n <= (int64_t(1) << 22) ? kMaxBits : pass_bits(); }

int radix_passes(int64_t n, int bits) { return bits <= 0 ? 0 : (bits + bits_per_pass(n) - 1) / bits_per_pass(n); }

int launch_radix_pass(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_out, void* vals_out,
                      int64_t n, uint64_t xor_mask, int shift, int bits, void* scratch, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  TFS_CHECK(bits >= 1 && bits <= kMaxBits, TFS_ERR_INVALID, "internal: radix pass of " + std::to_string(bits) + " bits");
  TFS_CHECK(n < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "internal: radix pass over 2^32 or more pairs");
  using u32 = unsigned int;
  using u64 = unsigned long long;
  if (key_bytes == 4 && val_bytes == 4) {
    if (vals_in == nullptr)
      return launch_pass_t<u32, u32, M_IOTA>(static_cast<const u32*>(keys_in), nullptr, static_cast<u32*>(keys_out), static_cast<u32*>(vals_out),
                                           n, u32(xor_mask), shift, bits, scratch, dev, stream);
    return launch_pass_t<u32, u32, M_PLAIN>(static_cast<const u32*>(keys_in), static_cast<const u32*>(vals_in), static_cast<u32*>(keys_out),
                                          static_cast<u32*>(vals_out), n, u32(xor_mask), shift, bits, scratch, dev, stream);
  }
  TFS_CHECK(vals_in != nullptr, TFS_ERR_INVALID, "internal: index values are 4 bytes wide");
  if (key_bytes == 4 && val_bytes == 8)
    return launch_pass_t<u32, u64, M_PLAIN>(static_cast<const u32*>(keys_in), static_cast<const u64*>(vals_in), static_cast<u32*>(keys_out),
                                          static_cast<u64*>(vals_out), n, u32(xor_mask), shift, bits, scratch, dev, stream);
  if (key_bytes == 8 && val_bytes == 4)
    return launch_pass_t<u64, u32, M_PLAIN>(static_cast<const u64*>(keys_in), static_cast<const u32*>(vals_in), static_cast<u64*>(keys_out),
                                          static_cast<u32*>(vals_out), n, u64(xor_mask), shift, bits, scratch, dev, stream);
  fail(TFS_ERR_INVALID, "internal: radix pass for this key / value width is not built");
}

// Stable sort of n pairs on key bits [0, bits): passes from the low bits up, reading `keys_in / vals_in` (vals_in null:
// the values are the element indices) and alternating between the two buffer pairs.  The last pass writes its keys only
// when `want_keys`.  Returns the index (0 / 1) of the buffer pair that holds the result; *launches is advanced.
int launch_radix_sort(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_buf[2], void* vals_buf[2],
                      int64_t n, uint64_t xor_mask, int bits, bool want_keys, void* scratch, const DeviceInfo& dev, cudaStream_t stream,
                      int* launches) {
  const int passes = radix_passes(n, bits);
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

// The row-by-slot ordering of K3: slots[n] -> rows_out[n] = row indices ordered by slot, stably.  Between the passes the
// pairs travel packed (slot << 32 | row) through the two 8-byte scratch columns; the first pass reads the row index off
// the element's position, the last one writes row indices only.
int launch_radix_rows_by_slot(const unsigned int* slots, int64_t n, int bits, unsigned long long* pairs_a, unsigned long long* pairs_b,
                              unsigned int* rows_out, unsigned int* seg_lo, unsigned int* seg_hi, void* scratch, const DeviceInfo& dev,
                              cudaStream_t stream) {
  if (n <= 0) return 0;
  TFS_CHECK((seg_lo == nullptr) == (seg_hi == nullptr), TFS_ERR_INVALID, "internal: segment bounds come as a pair");
  TFS_CHECK(n < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "internal: radix pass over 2^32 or more pairs");
  using u32 = unsigned int;
  const int passes = radix_passes(n, bits);
  int launches = 0, shift = 0;
  TFS_CHECK(seg_lo == nullptr || passes >= 2, TFS_ERR_INVALID, "internal: segment bounds need a partition of two passes or more");
  if (passes == 1)
    return launch_pass_t<u32, u32, M_IOTA>(slots, nullptr, nullptr, rows_out, n, 0u, 0, bits, scratch, dev, stream);
  unsigned long long* cur = nullptr;
  for (int p = 0; p < passes; ++p) {
    const int b = (bits - shift + (passes - p) - 1) / (passes - p);
    unsigned long long* dst = (cur == pairs_a) ? pairs_b : pairs_a;
    if (p == 0)
      launches += launch_pass_t<u32, u32, M_IOTA_PACK>(slots, nullptr, reinterpret_cast<u32*>(dst), nullptr, n, 0u, shift, b, scratch, dev, stream);
    else if (p < passes - 1)
      launches += launch_pass_t<u32, u32, M_PACKED>(reinterpret_cast<const u32*>(cur), nullptr, reinterpret_cast<u32*>(dst), nullptr, n, 0u,
                                                    shift, b, scratch, dev, stream);
    else if (seg_lo == nullptr)
      launches += launch_pass_t<u32, u32, M_UNPACK>(reinterpret_cast<const u32*>(cur), nullptr, nullptr, rows_out, n, 0u, shift, b, scratch,
                                                    dev, stream);
    else
      launches += launch_pass_t<u32, u32, M_UNPACK_BOUNDS>(reinterpret_cast<const u32*>(cur), nullptr, nullptr, rows_out, n, 0u, shift, b,
                                                           scratch, dev, stream, seg_lo, seg_hi);
    cur = dst;
    shift += b;
  }
  return launches;
}

}  // namespace tfs
