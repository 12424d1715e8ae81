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

#include "kernels.h"

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
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    TFS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    TFS_CHECK(p != nullptr && q == cudaDriverEntryPointSuccess, TFS_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
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
  static bool attr_set = false;
  if (!attr_set) {
    TFS_CUDA(cudaFuncSetAttribute(matmul_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmemBytes)));
    attr_set = true;
  }
  const CUtensorMap map_x = make_map(x, uint64_t(M), uint64_t(K), kBlockM);
  const CUtensorMap map_whi = make_map(w_hi, uint64_t(N), uint64_t(K), uint32_t(N));
  const CUtensorMap map_wlo = make_map(w_lo, uint64_t(N), uint64_t(K), uint32_t(N));
  // instruction descriptor: D = f32, A = B = tf32, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(kBlockM >> 4) << 24);
  const int64_t n_tiles = (M + kBlockM - 1) / kBlockM;
  const int grid = int(std::min<int64_t>(dev.sm_count, n_tiles));
  matmul_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>(map_x, map_whi, map_wlo, bias, out, M, int(N), int(K), idesc);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace tfs
