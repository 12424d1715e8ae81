// K4 (tensor-core path) -- out[M, N] = x[M, K] * w[K, N] + bias[N] in float32 on tcgen05 (sm_100a).
//
// The DataFrame-shaped MatMul of BASELINE config 5: a tall block of rows times a small constant
// matrix.  Float32 results are required to 1e-5, which one TF32 pass (10-bit mantissa) cannot
// give, so every operand is split into two TF32-exact parts, v = hi + lo with hi = v with its low
// 13 mantissa bits cleared and lo = v - hi (exact in fp32), and three MMAs accumulate
// hi*hi + lo*hi + hi*lo in the fp32 TMEM accumulator (the dropped lo*lo term is ~2^-22 relative).
//
// Structure (one persistent CTA per SM, 128-row tiles, the whole N <= 256 per tile):
//   warp 0      TMA producer   cp.async.bulk.tensor loads of the x chunk [128 x 32] and the two
//                              weight chunks [N x 32] (hi, lo) per K step, 128B-swizzled, 2 stages
//   warps 6-9   splitter       compute lo(x) = x - trunc_tf32(x) of the landed x chunk into a second buffer
//                              (the raw chunk serves as hi(x): the tensor core ignores the low 13 bits)
//   warp 1      MMA issuer     one elected thread: 12 tcgen05.mma.kind::tf32 (M128 x N x K8) per K step
//                              into one of two TMEM accumulators (2 x 256 columns)
//   warps 2-5   epilogue       tcgen05.ld the finished accumulator, add bias, store rows -- while the
//                              MMA warp already works on the next tile in the other accumulator
// All hand-offs are mbarriers (TMA complete_tx, tcgen05.commit, thread arrivals).
// The weights are pre-split once into two [N, K] K-major buffers (split_weights_kernel).
//
// Roofline: tensor pipe (3 TF32 MMAs per product) for N = K = 256; HBM moves 2 KB per row.
#include <cuda.h>

#include <cstdlib>
#include <string>

#include "kernels.h"

// timeline hook for experiments (scratch builds define it to record clock64 per role and iteration)
#ifndef TFS_MM_TRACE
#define TFS_MM_TRACE(ev, idx)
#endif

namespace tfs {
namespace {

constexpr int kBlockM = 128;
constexpr int kChunkK = 32;            // floats per K step = one 128-byte swizzle row
constexpr int kStages = 2;
constexpr int kMaxN = 256;
constexpr int kThreads = 320;          // 10 warps, roles above
constexpr uint32_t kTmemCols = 512;

constexpr uint32_t kABytes = kBlockM * kChunkK * 4;      // 16 KB
constexpr uint32_t kBBytesMax = kMaxN * kChunkK * 4;     // 32 KB
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytesMax;   // x hi, x lo, w hi, w lo = 96 KB
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024;    // + slack to align the base to 1024

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// K-major, 128-byte-swizzled operand tile: rows of 128 bytes, 8-row atoms 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc(const void* smem_tile) {
  return uint64_t((smem_u32(smem_tile) >> 4) & 0x3FFF) | (uint64_t(1) << 16) | (uint64_t(1024 >> 4) << 32) |
         (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}

struct Shared {
  uint64_t full[kStages];        // TMA bytes of the stage have landed
  uint64_t split[kStages];       // splitter finished the stage
  uint64_t empty[kStages];       // MMAs reading the stage have completed
  uint64_t tmem_full[2];         // accumulator finished
  uint64_t tmem_empty[2];        // accumulator drained by the epilogue
  uint32_t tmem_base;
  float bias[kMaxN];
};

__global__ void __launch_bounds__(kThreads, 1)
matmul_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                 const __grid_constant__ CUtensorMap map_wlo, const float* __restrict__ bias, float* __restrict__ out,
                 const int64_t M, const int N, const int K, const uint32_t idesc) {
  extern __shared__ unsigned char dyn_smem_raw[];
  __shared__ Shared sh;
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(dyn_smem_raw) + 1023) & ~uintptr_t(1023));
  auto stage_x_hi = [&](int s) { return smem + size_t(s) * kStageBytes; };
  auto stage_x_lo = [&](int s) { return smem + size_t(s) * kStageBytes + kABytes; };
  auto stage_w_hi = [&](int s) { return smem + size_t(s) * kStageBytes + 2 * kABytes; };
  auto stage_w_lo = [&](int s) { return smem + size_t(s) * kStageBytes + 2 * kABytes + kBBytesMax; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_tiles = (M + kBlockM - 1) / kBlockM;
  const int k_chunks = K / kChunkK;
  const uint32_t w_bytes = uint32_t(N) * kChunkK * 4;

  for (int i = threadIdx.x; i < N; i += kThreads) sh.bias[i] = bias ? bias[i] : 0.f;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&sh.full[s], 1); mbar_init(&sh.split[s], 4); mbar_init(&sh.empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&sh.tmem_full[b], 1); mbar_init(&sh.tmem_empty[b], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // TMEM allocation is a whole-warp operation; this warp also frees it
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh.tmem_base)), "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem_base = sh.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kc = 0; kc < k_chunks; ++kc, ++it) {
          const int s = it % kStages;
          mbar_wait(&sh.empty[s], ((it / kStages) & 1) ^ 1);
          mbar_expect_tx(&sh.full[s], kABytes + 2 * w_bytes);
          tma_load_2d(stage_x_hi(s), &map_x, kc * kChunkK, int(tile * kBlockM), &sh.full[s]);
          tma_load_2d(stage_w_hi(s), &map_whi, kc * kChunkK, 0, &sh.full[s]);
          tma_load_2d(stage_w_lo(s), &map_wlo, kc * kChunkK, 0, &sh.full[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    uint32_t it = 0, tile_it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tile_it) {
      const int buf = tile_it & 1;
      mbar_wait(&sh.tmem_empty[buf], ((tile_it >> 1) & 1) ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;");
      const uint32_t d = tmem_base + uint32_t(buf) * 256u;
      for (int kc = 0; kc < k_chunks; ++kc, ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        const uint64_t a_hi = umma_desc(stage_x_hi(s)), a_lo = umma_desc(stage_x_lo(s));
        const uint64_t b_hi = umma_desc(stage_w_hi(s)), b_lo = umma_desc(stage_w_lo(s));
        // The tensor core reads only the top 19 bits of each fp32 operand, so the landed x chunk already
        // acts as its own hi part: the two products that need nothing else start as soon as TMA is done ...
        mbar_wait(&sh.full[s], ph);
        asm volatile("tcgen05.fence::after_thread_sync;");
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < kChunkK / 8; ++k) {          // UMMA_K = 8 tf32 values = 32 bytes: +2 in 16-byte units
            const uint64_t o = uint64_t(k * 2);
            umma_tf32(d, a_hi + o, b_hi + o, idesc, (kc | k) != 0);
            umma_tf32(d, a_hi + o, b_lo + o, idesc, 1);
          }
        }
        __syncwarp();
        // ... and the lo(x) * hi(w) product follows once the splitter has written lo(x)
        mbar_wait(&sh.split[s], ph);
        asm volatile("tcgen05.fence::after_thread_sync;");
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < kChunkK / 8; ++k) umma_tf32(d, a_lo + uint64_t(k * 2), b_hi + uint64_t(k * 2), idesc, 1);
          umma_commit(&sh.empty[s]);                       // stage reusable once these MMAs have read it
          if (kc == k_chunks - 1) umma_commit(&sh.tmem_full[buf]);
        }
        __syncwarp();
      }
    }
  } else if (warp < 6) {
    // ------------------------------------------------ epilogue (warps 2..5 <-> TMEM lane quarters 2,3,0,1)
    const int quarter = warp & 3;
    uint32_t tile_it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tile_it) {
      const int buf = tile_it & 1;
      mbar_wait(&sh.tmem_full[buf], (tile_it >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;");
      const int64_t row = tile * kBlockM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + uint32_t(buf) * 256u;
      for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + uint32_t(c0), r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (row < M) {
          float* dst = out + row * N + c0;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (c0 + j < N) {   // N is a multiple of 16, so groups of 4 never straddle the edge
              float4 v;
              v.x = __fadd_rn(__uint_as_float(r[j + 0]), sh.bias[c0 + j + 0]);
              v.y = __fadd_rn(__uint_as_float(r[j + 1]), sh.bias[c0 + j + 1]);
              v.z = __fadd_rn(__uint_as_float(r[j + 2]), sh.bias[c0 + j + 2]);
              v.w = __fadd_rn(__uint_as_float(r[j + 3]), sh.bias[c0 + j + 3]);
              __stcs(reinterpret_cast<float4*>(dst + j), v);
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;");
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh.tmem_empty[buf]);
    }
  } else {
    // ------------------------------------------------ splitter (warps 6..9): x -> (hi, lo)
    const int t = threadIdx.x - 6 * 32;                    // 0..127
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      for (int kc = 0; kc < k_chunks; ++kc, ++it) {
        const int s = it % kStages;
        mbar_wait(&sh.full[s], (it / kStages) & 1);
        const uint4* hi = reinterpret_cast<const uint4*>(stage_x_hi(s));
        uint4* lo = reinterpret_cast<uint4*>(stage_x_lo(s));
#pragma unroll
        for (int j = 0; j < int(kABytes / 16 / 128); ++j) {     // 1024 16-byte vectors over 128 threads
          const int idx = j * 128 + t;
          uint4 v = hi[idx], h, l;
          h.x = v.x & 0xFFFFE000u; h.y = v.y & 0xFFFFE000u; h.z = v.z & 0xFFFFE000u; h.w = v.w & 0xFFFFE000u;
          l.x = __float_as_uint(__fsub_rn(__uint_as_float(v.x), __uint_as_float(h.x)));
          l.y = __float_as_uint(__fsub_rn(__uint_as_float(v.y), __uint_as_float(h.y)));
          l.z = __float_as_uint(__fsub_rn(__uint_as_float(v.z), __uint_as_float(h.z)));
          l.w = __float_as_uint(__fsub_rn(__uint_as_float(v.w), __uint_as_float(h.w)));
          lo[idx] = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA unit
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.split[s]);
      }
    }
  }

  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}


// ---------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2): two CTAs of one cluster -- the two SMs of a TPC -- work on
// one 256-row tile.  Each CTA stages its own 128 rows of x (raw + lo part) and HALF of the weight
// chunk (N/2 rows of hi and of lo); the leader CTA issues M256 x N x K8 MMAs that read both shared
// memories, and each CTA's TMEM receives the accumulator of its own 128 rows.  Per CTA and K step
// this moves 48 KB into shared memory instead of 80 KB and the tensor core reads 8 KB per MMA instead
// of 12 KB: the single-CTA kernel above needs more shared-memory bandwidth (~256 KB per K step) than
// an SM has in the time its 12 MMAs take (1536 clk x 128 B/clk = 192 KB).
//
// A per-role clock64 timeline of the first version (profiles/) showed what else stalls the tensor
// pipe, and this kernel is laid out against it:
//   * the x stream comes from HBM (~2000 clk from issue to landed) while w comes from L2: they get
//     separate rings, 4 deep for x and 2 deep for w, filled by separate producer warps;
//   * the two products that need only the raw chunk (hi(x) = the raw bits, truncated by the tensor
//     core) start as soon as both CTAs have their x chunk, the lo(x) product when the splitters
//     are done;
//   * per-thread row stores in the epilogue (32 lines per warp instruction) kept the LSU busy and
//     stretched the splitter 4x while a tile drained: the epilogue now writes 32 x 32 boxes into
//     swizzled shared memory and lets TMA store them.
//
// Barriers (same offsets in both CTAs; "leader" = cluster rank 0):
//   xfull[s]       local     this CTA's x chunk has landed (TMA complete_tx)        -> local splitter
//   xboth[s]       leader    2 arrivals: each CTA saw its xfull[s]                  -> MMA issuer
//   split[s]       leader    8 arrivals: the 4 splitter warps of each CTA (remote mbarrier.arrive from
//                            the peer)                                               -> MMA issuer
//   wfull[s]       leader    both halves of w hi/lo have landed (the peer's TMA signals the leader's
//                            barrier: cp.async.bulk.tensor .cta_group::2)            -> MMA issuer
//   xempty[s], wempty[s]  both   tcgen05.commit multicast: the MMAs reading the stage are done -> producers
//   tmem_full[b]   both      tcgen05.commit multicast: accumulator b is complete     -> epilogue warps
//   tmem_empty[b]  leader    8 arrivals: the 4 epilogue warps of each CTA            -> MMA issuer
constexpr int kXStages = 4;
constexpr int kWStages = 2;
constexpr int kThreads2 = 352;         // warps: 0 x producer, 1 MMA, 2-5 epilogue, 6-9 splitter, 10 w producer
constexpr uint32_t kBHalfBytesMax = (kMaxN / 2) * kChunkK * 4;                 // 16 KB
constexpr uint32_t kXStageBytes = 2 * kABytes;                                 // raw + lo: 32 KB
constexpr uint32_t kWStageBytes = 2 * kBHalfBytesMax;                          // hi + lo halves: 32 KB
constexpr uint32_t kOutBoxBytes = 32 * 32 * 4;                                 // one epilogue warp's 32 x 32 box
constexpr uint32_t kOutBytes = 4 * 2 * kOutBoxBytes;                           // 4 warps, double-buffered: 32 KB
constexpr uint32_t kSmemBytes2 = kXStages * kXStageBytes + kWStages * kWStageBytes + kOutBytes + 1024;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (an object in this CTA's shared memory) as seen in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
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
// TMA load into this CTA's shared memory whose completion bytes are credited to a barrier that may live in
// the peer CTA (cluster address)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint32_t bar_cluster_addr) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {   // arrives on `bar` in both CTAs of the pair
  const uint16_t mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}

struct SharedPair {
  uint64_t xfull[kXStages];
  uint64_t xboth[kXStages];
  uint64_t split[kXStages];
  uint64_t xempty[kXStages];
  uint64_t wfull[kWStages];
  uint64_t wempty[kWStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  float bias[kMaxN];
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
matmul_tc_pair_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                      const __grid_constant__ CUtensorMap map_wlo, const __grid_constant__ CUtensorMap map_out,
                      const float* __restrict__ bias, const int64_t M, const int N, const int K, const uint32_t idesc) {
  extern __shared__ unsigned char dyn_smem_raw[];
  __shared__ SharedPair sh;
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(dyn_smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* w_ring = smem + size_t(kXStages) * kXStageBytes;
  unsigned char* out_stage = w_ring + size_t(kWStages) * kWStageBytes;
  auto stage_x_hi = [&](int s) { return smem + size_t(s) * kXStageBytes; };
  auto stage_x_lo = [&](int s) { return smem + size_t(s) * kXStageBytes + kABytes; };
  auto stage_w_hi = [&](int s) { return w_ring + size_t(s) * kWStageBytes; };
  auto stage_w_lo = [&](int s) { return w_ring + size_t(s) * kWStageBytes + kBHalfBytesMax; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int64_t pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int64_t n_tiles = (M + 2 * kBlockM - 1) / (2 * kBlockM);     // 256-row tiles
  const int k_chunks = K / kChunkK;
  const int half_n = N / 2;
  const uint32_t w_half_bytes = uint32_t(half_n) * kChunkK * 4;

  for (int i = threadIdx.x; i < N; i += kThreads2) sh.bias[i] = bias ? bias[i] : 0.f;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kXStages; ++s) {
      mbar_init(&sh.xfull[s], 1); mbar_init(&sh.xboth[s], 2); mbar_init(&sh.split[s], 8); mbar_init(&sh.xempty[s], 1);
    }
    for (int s = 0; s < kWStages; ++s) { mbar_init(&sh.wfull[s], 1); mbar_init(&sh.wempty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&sh.tmem_full[b], 1); mbar_init(&sh.tmem_empty[b], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // one warp of each CTA of the pair takes part in the allocation (and later frees it)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh.tmem_base)), "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  cluster_sync_all();   // barriers of both CTAs are initialised before anyone signals across the pair
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem_base = sh.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------ x producer: this CTA's 128 rows, chunk by chunk
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = pair_id; tile < n_tiles; tile += n_pairs) {
        const int row0 = int(tile * 2 * kBlockM + int64_t(rank) * kBlockM);
        for (int kc = 0; kc < k_chunks; ++kc, ++it) {
          const int s = it % kXStages;
          mbar_wait(&sh.xempty[s], ((it / kXStages) & 1) ^ 1);
          TFS_MM_TRACE(0, it);
          mbar_expect_tx(&sh.xfull[s], kABytes);
          tma_load_2d(stage_x_hi(s), &map_x, kc * kChunkK, row0, &sh.xfull[s]);
        }
      }
    }
  } else if (warp == 10) {
    // ------------------------------------------------ w producer: this CTA's half of the hi and lo chunks
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = pair_id; tile < n_tiles; tile += n_pairs) {
        for (int kc = 0; kc < k_chunks; ++kc, ++it) {
          const int s = it % kWStages;
          mbar_wait(&sh.wempty[s], ((it / kWStages) & 1) ^ 1);
          const uint32_t wfull_leader = mapa_u32(&sh.wfull[s], 0);
          if (leader) mbar_expect_tx(&sh.wfull[s], 4 * w_half_bytes);     // hi + lo halves of both CTAs
          tma_load_2d_pair(stage_w_hi(s), &map_whi, kc * kChunkK, int(rank) * half_n, wfull_leader);
          tma_load_2d_pair(stage_w_lo(s), &map_wlo, kc * kChunkK, int(rank) * half_n, wfull_leader);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (leader CTA only)
    if (leader) {
      uint32_t it = 0, tile_it = 0;
      for (int64_t tile = pair_id; tile < n_tiles; tile += n_pairs, ++tile_it) {
        const int buf = tile_it & 1;
        mbar_wait_cluster(&sh.tmem_empty[buf], ((tile_it >> 1) & 1) ^ 1);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t d = tmem_base + uint32_t(buf) * 256u;
        for (int kc = 0; kc < k_chunks; ++kc, ++it) {
          const int xs = it % kXStages, ws = it % kWStages;
          const uint32_t xph = (it / kXStages) & 1, wph = (it / kWStages) & 1;
          const uint64_t a_hi = umma_desc(stage_x_hi(xs)), a_lo = umma_desc(stage_x_lo(xs));
          const uint64_t b_hi = umma_desc(stage_w_hi(ws)), b_lo = umma_desc(stage_w_lo(ws));
          // raw x acts as hi(x) (the tensor core reads only the top 19 bits): these two products need
          // nothing from the splitters
          mbar_wait_cluster(&sh.wfull[ws], wph);
          mbar_wait_cluster(&sh.xboth[xs], xph);
          asm volatile("tcgen05.fence::after_thread_sync;");
          if (lane == 0) {
            TFS_MM_TRACE(1, it);
#pragma unroll
            for (int k = 0; k < kChunkK / 8; ++k) {          // UMMA_K = 8 tf32 values = 32 bytes: +2 in 16-byte units
              const uint64_t o = uint64_t(k * 2);
              umma_tf32_pair(d, a_hi + o, b_hi + o, idesc, (kc | k) != 0);
              umma_tf32_pair(d, a_hi + o, b_lo + o, idesc, 1);
            }
          }
          __syncwarp();
          mbar_wait_cluster(&sh.split[xs], xph);
          asm volatile("tcgen05.fence::after_thread_sync;");
          if (lane == 0) {
            TFS_MM_TRACE(2, it);
#pragma unroll
            for (int k = 0; k < kChunkK / 8; ++k) umma_tf32_pair(d, a_lo + uint64_t(k * 2), b_hi + uint64_t(k * 2), idesc, 1);
            umma_commit_pair(&sh.xempty[xs]);
            umma_commit_pair(&sh.wempty[ws]);
            if (kc == k_chunks - 1) umma_commit_pair(&sh.tmem_full[buf]);
            TFS_MM_TRACE(3, it);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp < 6) {
    // ------------------------------------------------ epilogue (warps 2..5 <-> TMEM lane quarters 2,3,0,1)
    // Each warp owns 32 accumulator rows: 32-column slices go TMEM -> registers (+ bias) -> a 32 x 32 box in
    // shared memory (128-byte rows, 16-byte chunks XOR-swizzled by row like the TMA map expects) -> TMA store.
    // Rows past M are clipped by TMA.
    const int quarter = warp & 3;
    const uint32_t tmem_empty_leader0 = mapa_u32(&sh.tmem_empty[0], 0), tmem_empty_leader1 = mapa_u32(&sh.tmem_empty[1], 0);
    unsigned char* my_out = out_stage + size_t(warp - 2) * 2 * kOutBoxBytes;
    uint32_t tile_it = 0, box_it = 0;
    for (int64_t tile = pair_id; tile < n_tiles; tile += n_pairs, ++tile_it) {
      const int buf = tile_it & 1;
      mbar_wait(&sh.tmem_full[buf], (tile_it >> 1) & 1);
      if (warp == 2 && lane == 0) { TFS_MM_TRACE(6, tile_it); }
      asm volatile("tcgen05.fence::after_thread_sync;");
      const int row0 = int(tile * 2 * kBlockM + int64_t(rank) * kBlockM + quarter * 32);
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + uint32_t(buf) * 256u;
      for (int c0 = 0; c0 < N; c0 += 32, ++box_it) {
        uint32_t r[32];
        tmem_ld32(taddr + uint32_t(c0), r);
        unsigned char* box = my_out + size_t(box_it & 1) * kOutBoxBytes;
        // the TMA store that last read this buffer (two boxes ago) must have finished reading it
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int ncols = N - c0 < 32 ? N - c0 : 32;     // 16 or 32 (N is a multiple of 16)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j * 4 < ncols) {
            float4 v;
            v.x = __fadd_rn(__uint_as_float(r[4 * j + 0]), sh.bias[c0 + 4 * j + 0]);
            v.y = __fadd_rn(__uint_as_float(r[4 * j + 1]), sh.bias[c0 + 4 * j + 1]);
            v.z = __fadd_rn(__uint_as_float(r[4 * j + 2]), sh.bias[c0 + 4 * j + 2]);
            v.w = __fadd_rn(__uint_as_float(r[4 * j + 3]), sh.bias[c0 + 4 * j + 3]);
            *reinterpret_cast<float4*>(box + lane * 128 + ((j ^ (lane & 7)) << 4)) = v;
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&map_out, box, c0, row0);       // columns past N are clipped as well
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;");
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(buf ? tmem_empty_leader1 : tmem_empty_leader0);
      if (warp == 2 && lane == 0) { TFS_MM_TRACE(7, tile_it); }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // all boxes written before the CTA leaves
  } else {
    // ------------------------------------------------ splitter (warps 6..9): lo(x) = x - trunc_tf32(x)
    const int t = threadIdx.x - 6 * 32;                    // 0..127
    uint32_t it = 0;
    for (int64_t tile = pair_id; tile < n_tiles; tile += n_pairs) {
      for (int kc = 0; kc < k_chunks; ++kc, ++it) {
        const int s = it % kXStages;
        mbar_wait(&sh.xfull[s], (it / kXStages) & 1);
        if (t == 0) {
          mbar_arrive_cluster(mapa_u32(&sh.xboth[s], 0));   // tell the MMA issuer this CTA's chunk is in
          TFS_MM_TRACE(4, it);
        }
        const uint4* hi = reinterpret_cast<const uint4*>(stage_x_hi(s));
        uint4* lo = reinterpret_cast<uint4*>(stage_x_lo(s));
#pragma unroll
        for (int j = 0; j < int(kABytes / 16 / 128); ++j) {     // 1024 16-byte vectors over 128 threads
          const int idx = j * 128 + t;
          uint4 v = hi[idx], h, l;
          h.x = v.x & 0xFFFFE000u; h.y = v.y & 0xFFFFE000u; h.z = v.z & 0xFFFFE000u; h.w = v.w & 0xFFFFE000u;
          l.x = __float_as_uint(__fsub_rn(__uint_as_float(v.x), __uint_as_float(h.x)));
          l.y = __float_as_uint(__fsub_rn(__uint_as_float(v.y), __uint_as_float(h.y)));
          l.z = __float_as_uint(__fsub_rn(__uint_as_float(v.z), __uint_as_float(h.z)));
          l.w = __float_as_uint(__fsub_rn(__uint_as_float(v.w), __uint_as_float(h.w)));
          lo[idx] = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA unit
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_cluster(mapa_u32(&sh.split[s], 0));
          if (t == 0) { TFS_MM_TRACE(5, it); }
        }
      }
    }
  }

  asm volatile("tcgen05.fence::before_thread_sync;");
  cluster_sync_all();   // neither CTA may leave (or free TMEM) while the pair's MMAs can still touch it
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

// w: row-major [K, N] (or [N, K] when already transposed) -> hi/lo parts as [N, K] K-major
__global__ void split_weights_kernel(const float* __restrict__ w, const bool transposed, const int K, const int N,
                                     float* __restrict__ hi, float* __restrict__ lo) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= K * N) return;
  const int n = idx / K, k = idx % K;
  const float v = transposed ? w[n * K + k] : w[k * N + n];
  const float h = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
  hi[idx] = h;
  lo[idx] = __fsub_rn(v, h);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static const EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    TFS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    TFS_CHECK(p != nullptr && q == cudaDriverEntryPointSuccess, TFS_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// rows x cols float32, row-major; box = box_rows x 32 floats, 128-byte swizzle, out-of-bounds rows read as zero
CUtensorMap make_map(const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 4};
  cuuint32_t box[2] = {kChunkK, box_rows};
  cuuint32_t elem[2] = {1, 1};
  CUresult r = encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, elem,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFS_CHECK(r == CUDA_SUCCESS, TFS_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string(int(r)));
  return m;
}

}  // namespace

bool matmul_tc_supported(const void* a, const void* out, int64_t M, int64_t N, int64_t K, bool ta) {
  auto aligned = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  return !ta && M >= 1 && K >= kChunkK && K % kChunkK == 0 && N >= 16 && N <= kMaxN && N % 16 == 0 && aligned(a) &&
         aligned(out) && M < (int64_t(1) << 31);
}

size_t matmul_tc_weight_bytes(int64_t N, int64_t K) { return size_t(N) * size_t(K) * 4; }

int launch_split_weights(const float* w, bool transposed, int64_t K, int64_t N, float* hi, float* lo, cudaStream_t stream) {
  const int total = int(K * N);
  split_weights_kernel<<<(total + 255) / 256, 256, 0, stream>>>(w, transposed, int(K), int(N), hi, lo);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

int launch_matmul_tc(const float* x, const float* w_hi, const float* w_lo, const float* bias, float* out, int64_t M,
                     int64_t N, int64_t K, const DeviceInfo& dev, cudaStream_t stream) {
  static PerDeviceOnce attr;
  attr.run(dev.device, [] {
    TFS_CUDA(cudaFuncSetAttribute(matmul_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmemBytes)));
    TFS_CUDA(cudaFuncSetAttribute(matmul_tc_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmemBytes2)));
  });
  static const bool single_cta = [] {
    const char* impl = std::getenv("TFS_MATMUL_IMPL");   // "tc1": the single-CTA kernel, kept for A/B measurements
    return impl != nullptr && std::string(impl) == "tc1";
  }();
  const CUtensorMap map_x = make_map(x, uint64_t(M), uint64_t(K), kBlockM);
  if (single_cta) {
    const CUtensorMap map_whi = make_map(w_hi, uint64_t(N), uint64_t(K), uint32_t(N));
    const CUtensorMap map_wlo = make_map(w_lo, uint64_t(N), uint64_t(K), uint32_t(N));
    // instruction descriptor: D = f32, A = B = tf32, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(kBlockM >> 4) << 24);
    const int64_t n_tiles = (M + kBlockM - 1) / kBlockM;
    const int grid = int(std::min<int64_t>(dev.sm_count, n_tiles));
    matmul_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>(map_x, map_whi, map_wlo, bias, out, M, int(N), int(K), idesc);
  } else {
    const CUtensorMap map_whi = make_map(w_hi, uint64_t(N), uint64_t(K), uint32_t(N / 2));
    const CUtensorMap map_wlo = make_map(w_lo, uint64_t(N), uint64_t(K), uint32_t(N / 2));
    const CUtensorMap map_out = make_map(out, uint64_t(M), uint64_t(N), 32);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t((2 * kBlockM) >> 4) << 24);
    const int64_t n_tiles = (M + 2 * kBlockM - 1) / (2 * kBlockM);
    const int pairs = int(std::min<int64_t>(dev.sm_count / 2, n_tiles));
    matmul_tc_pair_kernel<<<2 * pairs, kThreads2, kSmemBytes2, stream>>>(map_x, map_whi, map_wlo, map_out, bias, M, int(N), int(K), idesc);
  }
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace tfs
