// Experiment: issue-to-completion rate of back-to-back tcgen05.mma kind::tf32 / kind::f16 (operands in smem).
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory"); } while (!done);
}
__device__ __forceinline__ uint64_t umma_desc(const void* smem_tile) {
  return uint64_t((smem_u32(smem_tile) >> 4) & 0x3FFF) | (uint64_t(1) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
// KIND 0: tf32, 1: f16 (bf16 inputs).  PAIR: cta_group::2
template <int KIND, bool PAIR>
__global__ void __launch_bounds__(128, 1) rate_kernel(int n_mma, int N, int distinct, long long* out, const __grid_constant__ CUtensorMap mapx, int n_chunks, int split_iters) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar; __shared__ uint32_t tmem_base; __shared__ uint64_t full[4], empty[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank = 0;
  if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); } mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    if (PAIR) { asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512u)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;"); }
    else { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512u)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;"); }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  if (PAIR) { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); } else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t d = tmem_base;
  const int M = PAIR ? 256 : 128;
  const uint32_t idesc = KIND == 0 ? ((1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24))
                                   : ((1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24));
  if (warp == 1 && lane == 0 && n_chunks > 0) {          // TMA producer: x chunks [128 x 32] from HBM into a 4-deep ring at 160 KB
    unsigned char* ring = smem + 160 * 1024;
    for (int it = 0; it < n_chunks; ++it) {
      const int s = it & 3;
      mbar_wait(&empty[s], ((it >> 2) & 1) ^ 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(16384u) : "memory");
      const int64_t tile = blockIdx.x + int64_t(it >> 3) * gridDim.x;
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                   ::"r"(smem_u32(ring + s * 16384)), "l"(&mapx), "r"(smem_u32(&full[s])), "r"((it & 7) * 32), "r"(int(tile * 128)) : "memory");
    }
  } else if (warp == 2 && lane == 0 && n_chunks > 0) {   // consumer
    const long long t0 = clock64();
    for (int it = 0; it < n_chunks; ++it) {
      const int s = it & 3;
      mbar_wait(&full[s], (it >> 2) & 1);
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
    }
    if (blockIdx.x == 0) out[1] = clock64() - t0;
  } else if (warp == 3 && split_iters > 0) {             // splitter-like LDS/STS traffic: 16 KB read + 16 KB written per iteration by one warp
    const long long t0 = clock64();
    uint4* src = reinterpret_cast<uint4*>(smem); uint4* dst = reinterpret_cast<uint4*>(smem + 16384);
    for (int it = 0; it < split_iters; ++it) {
#pragma unroll 8
      for (int j = 0; j < 32; ++j) { uint4 v = src[j * 32 + lane]; v.x &= 0xFFFFE000u; dst[j * 32 + lane] = v; }
    }
    if (blockIdx.x == 0 && lane == 0) out[2] = clock64() - t0;
  }
  if (warp == 0 && rank == 0) {
    long long t0 = 0, t1 = 0;
    if (lane == 0) {
      t0 = clock64();
      for (int i = 0; i < n_mma; ++i) {
        // rotate through `distinct` different operand tiles (A: 16 KB tiles from offset 0, B: 32 KB tiles from 64 KB) and the 4 k-steps of each
        const int tile = (i >> 2) % distinct, k = i & 3;
        const uint64_t a = umma_desc(smem + size_t(tile) * 16384) + uint64_t(k * 2);
        const uint64_t b = umma_desc(smem + 65536 + size_t(tile % 3) * 32768) + uint64_t(k * 2);
        if (KIND == 0) {
          if (PAIR) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(i) : "memory");
          else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(i) : "memory");
        } else {
          if (PAIR) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(i) : "memory");
          else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(i) : "memory");
        }
      }
      if (PAIR) { const uint16_t mask = 1; asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"(mask) : "memory"); }
      else asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    if (lane == 0) { t1 = clock64(); if (blockIdx.x == 0) out[0] = t1 - t0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  if (PAIR) { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); } else __syncthreads();
  if (warp == 0) {
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}
CUtensorMap g_mapx;
template <int KIND, bool PAIR>
int run(const char* name, int N, int distinct, int grid, long long* out, int n_mma = 4096, int n_chunks = 0, int split_iters = 0) {
  const int smem = 226 * 1024, n = n_mma;
  CK(cudaFuncSetAttribute(rate_kernel<KIND, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = PAIR ? 2 : 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  for (int rep = 0; rep < 2; ++rep) { CK(cudaMemset(out, 0, 32)); CK(cudaLaunchKernelEx(&cfg, rate_kernel<KIND, PAIR>, n, N, distinct, out, g_mapx, n_chunks, split_iters)); CK(cudaDeviceSynchronize()); }
  long long hh[4]; CK(cudaMemcpy(hh, out, 32, cudaMemcpyDeviceToHost));
  const long long h = hh[0]; const double clk = n ? double(h) / n : 0;
  const int M = PAIR ? 256 : 128, Kk = KIND == 0 ? 8 : 16;
  printf("%-40s N=%3d grid=%3d: %6.1f clk/MMA (%.0f FLOP/clk/SM)", name, N, grid, clk, n ? 2.0 * M * N * Kk / clk / (PAIR ? 2 : 1) : 0.0);
  if (n_chunks) printf("  | x stream %.0f clk/chunk = %.1f B/clk", double(hh[1]) / n_chunks, 16384.0 * n_chunks / double(hh[1]));
  if (split_iters) printf("  | split warp %.0f clk/32KB", double(hh[2]) / split_iters);
  printf("\n");
  return 0;
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  long long* out; CK(cudaMalloc(&out, 32));
  const int64_t M = 4194304; float* x; CK(cudaMalloc(&x, M * 1024)); CK(cudaMemset(x, 0, M * 1024));
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  cuuint32_t ones[2] = {1, 1}; cuuint64_t d[2] = {256, (cuuint64_t)M}; cuuint64_t st[1] = {1024}; cuuint32_t bx[2] = {32, 128};
  printf("enc %d\n", ((EncodeTiledFn)p)(&g_mapx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, x, d, st, bx, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
  const int grid = 148;
  run<0, false>("tf32 1cta: MMA only", 256, 4, grid, out, 8192, 0, 0);
  run<0, false>("tf32 1cta: TMA x stream only", 256, 4, grid, out, 0, 1024, 0);
  run<0, false>("tf32 1cta: split warp only", 256, 4, grid, out, 0, 0, 512);
  run<0, false>("tf32 1cta: MMA + x stream", 256, 4, grid, out, 8192, 1024, 0);
  run<0, false>("tf32 1cta: MMA + split warp", 256, 4, grid, out, 8192, 0, 512);
  run<0, false>("tf32 1cta: MMA + x stream + split", 256, 4, grid, out, 8192, 1024, 512);
  run<0, true>("tf32 pair: MMA only", 256, 4, grid, out, 8192, 0, 0);
  run<0, true>("tf32 pair: MMA + x stream", 256, 4, grid, out, 8192, 1024, 0);
  run<0, true>("tf32 pair: MMA + split warp", 256, 4, grid, out, 8192, 0, 512);
  run<0, true>("tf32 pair: MMA + x stream + split", 256, 4, grid, out, 8192, 1024, 512);
  return 0;
}
