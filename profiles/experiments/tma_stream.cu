// Experiment: HBM read throughput of TMA tile loads of x[M, 256] f32 in different box shapes.
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory"); } while (!done);
}
// MODE 0: 2-D map (32 floats, rows): box 32 x 128, one load per (tile, chunk)            16 KB per op
// MODE 1: 3-D map (32 floats, rows, chunks): box 32 x 32 x 8, one load per 32-row block   32 KB per op
// MODE 2: 4-D map (32 floats, 8 rows, chunks, groups): box 32 x 8 x 4 x 16: half tile     64 KB per op
// MODE 3: 4-D map, box 32 x 8 x 8 x 4: 32 rows x full K in UMMA-friendly layout           32 KB per op
template <int MODE>
__global__ void __launch_bounds__(64, 1) stream_kernel(const __grid_constant__ CUtensorMap map, int64_t M, int depth, uint32_t op_bytes, unsigned long long* sink) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[16], empty[16];
  if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  const int64_t n_tiles = M / 128;
  const int ops_per_tile = MODE == 0 ? 8 : (MODE == 1 ? 4 : (MODE == 2 ? 2 : 4));
  if (threadIdx.x == 0) {
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int op = 0; op < ops_per_tile; ++op, ++it) {
        const int s = it % depth;
        mbar_wait(&empty[s], ((it / depth) & 1) ^ 1);
        mbar_expect_tx(&full[s], op_bytes);
        void* dst = smem + size_t(s) * op_bytes;
        if (MODE == 0)
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(smem_u32(dst)), "l"(&map), "r"(smem_u32(&full[s])), "r"(op * 32), "r"(int(tile * 128)) : "memory");
        else if (MODE == 1)
          asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                       ::"r"(smem_u32(dst)), "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"(int(tile * 128 + op * 32)), "r"(0) : "memory");
        else if (MODE == 2)
          asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                       ::"r"(smem_u32(dst)), "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"(0), "r"(op * 4), "r"(int(tile * 16)) : "memory");
        else
          asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                       ::"r"(smem_u32(dst)), "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"(0), "r"(0), "r"(int(tile * 16 + op * 4)) : "memory");
      }
  } else if (threadIdx.x == 32) {
    uint32_t it = 0; unsigned long long acc = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int op = 0; op < ops_per_tile; ++op, ++it) {
        const int s = it % depth;
        mbar_wait(&full[s], (it / depth) & 1);
        acc += *reinterpret_cast<unsigned long long*>(smem + size_t(s) * op_bytes);
        mbar_arrive(&empty[s]);
      }
    if (acc == 0x1234567) *sink = acc;
  }
}

// x chunk boxes (16 KB, from HBM) plus two 16 KB boxes of a small hot matrix (L2) per step, like the matmul kernel
__global__ void __launch_bounds__(96, 1) stream_xw_kernel(const __grid_constant__ CUtensorMap mapx, const __grid_constant__ CUtensorMap mapw, int64_t M, int depth, int wdepth, int w_rows, unsigned long long* sink) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  unsigned char* wsm = smem + 8 * 16384;
  __shared__ uint64_t full[16], empty[16], wfull[4], wempty[4];
  if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); } for (int i = 0; i < 4; ++i) { mbar_init(&wfull[i], 1); mbar_init(&wempty[i], 1); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  const int64_t n_tiles = M / 128;
  const uint32_t wbytes = uint32_t(w_rows) * 128;
  if (threadIdx.x == 0) {
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int op = 0; op < 8; ++op, ++it) {
        const int s = it % depth;
        mbar_wait(&empty[s], ((it / depth) & 1) ^ 1);
        mbar_expect_tx(&full[s], 16384);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(smem_u32(smem + size_t(s) * 16384)), "l"(&mapx), "r"(smem_u32(&full[s])), "r"(op * 32), "r"(int(tile * 128)) : "memory");
      }
  } else if (threadIdx.x == 64) {
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int op = 0; op < 8; ++op, ++it) {
        const int s = it % wdepth;
        mbar_wait(&wempty[s], ((it / wdepth) & 1) ^ 1);
        mbar_expect_tx(&wfull[s], 2 * wbytes);
        for (int h = 0; h < 2; ++h)
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(smem_u32(wsm + size_t(s) * 32768 + h * 16384)), "l"(&mapw), "r"(smem_u32(&wfull[s])), "r"(op * 32), "r"(h * 256 + (blockIdx.x & 1) * 128) : "memory");
      }
  } else if (threadIdx.x == 32) {
    uint32_t it = 0; unsigned long long acc = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
      for (int op = 0; op < 8; ++op, ++it) {
        const int s = it % depth, ws = it % wdepth;
        mbar_wait(&full[s], (it / depth) & 1);
        mbar_wait(&wfull[ws], (it / wdepth) & 1);
        acc += *reinterpret_cast<unsigned long long*>(smem + size_t(s) * 16384);
        mbar_arrive(&empty[s]);
        mbar_arrive(&wempty[ws]);
      }
    if (acc == 0x1234567) *sink = acc;
  }
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  const int64_t M = 16777216 / 2; const int K = 256;
  float* x; unsigned long long* sink;
  CK(cudaMalloc(&x, M * K * 4)); CK(cudaMalloc(&sink, 8)); CK(cudaMemset(x, 0, M * K * 4));
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  EncodeTiledFn enc = (EncodeTiledFn)p;
  cuuint32_t ones[4] = {1, 1, 1, 1};
  CUtensorMap m0, m1, m2, m3;
  { cuuint64_t d[2] = {256, (cuuint64_t)M}; cuuint64_t st[1] = {1024}; cuuint32_t b[2] = {32, 128};
    printf("enc0 %d\n", enc(&m0, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, x, d, st, b, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
  { cuuint64_t d[3] = {32, (cuuint64_t)M, 8}; cuuint64_t st[2] = {1024, 128}; cuuint32_t b[3] = {32, 32, 8};
    printf("enc1 %d\n", enc(&m1, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, x, d, st, b, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
  { cuuint64_t d[4] = {32, 8, 8, (cuuint64_t)M / 8}; cuuint64_t st[3] = {1024, 128, 8192}; cuuint32_t b[4] = {32, 8, 4, 16};
    printf("enc2 %d\n", enc(&m2, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, d, st, b, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
  { cuuint64_t d[4] = {32, 8, 8, (cuuint64_t)M / 8}; cuuint64_t st[3] = {1024, 128, 8192}; cuuint32_t b[4] = {32, 8, 8, 4};
    printf("enc3 %d\n", enc(&m3, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, d, st, b, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int smem = 200 * 1024;
  CK(cudaFuncSetAttribute(stream_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(stream_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(stream_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  auto report = [&](const char* name, float ms) { printf("%-44s %.3f ms  %.0f GB/s\n", name, ms, M * K * 4.0 / ms / 1e6); };
  for (int rep = 0; rep < 1; ++rep) {
    float ms;
    for (int depth : {2, 4, 8, 12}) {
      cudaEventRecord(a); stream_kernel<0><<<148, 64, smem>>>(m0, M, depth, 16384, sink); cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
      char n[64]; snprintf(n, 64, "2D 128x32 chunk boxes (16 KB), depth %d", depth); report(n, ms);
    }
    for (int depth : {2, 4, 6}) {
      cudaEventRecord(a); stream_kernel<1><<<148, 64, smem>>>(m1, M, depth, 32768, sink); cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
      char n[64]; snprintf(n, 64, "3D 32 rows x full K (32 KB), depth %d", depth); report(n, ms);
    }
    for (int depth : {2, 3}) {
      cudaEventRecord(a); stream_kernel<2><<<148, 64, smem>>>(m2, M, depth, 65536, sink); cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
      char n[64]; snprintf(n, 64, "4D 128 rows x half K (64 KB), depth %d", depth); report(n, ms);
    }
    for (int depth : {2, 4, 6}) {
      cudaEventRecord(a); stream_kernel<3><<<148, 64, smem>>>(m3, M, depth, 32768, sink); cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
      char n[64]; snprintf(n, 64, "4D 32 rows x full K, umma layout, depth %d", depth); report(n, ms);
    }
  }
  {
    float* w; CK(cudaMalloc(&w, 512 * 256 * 4)); CK(cudaMemset(w, 0, 512 * 256 * 4));
    CK(cudaFuncSetAttribute(stream_xw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    for (int w_rows : {128, 64, 16}) {
      CUtensorMap mw; cuuint64_t d[2] = {256, 512}; cuuint64_t st[1] = {1024}; cuuint32_t bx[2] = {32, (cuuint32_t)w_rows};
      printf("encw %d\n", enc(&mw, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, w, d, st, bx, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
      for (int rep = 0; rep < 2; ++rep) {
        float ms;
        cudaEventRecord(a); stream_xw_kernel<<<148, 96, smem>>>(m0, mw, M, 4, 2, w_rows, sink); cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
        char n[80]; snprintf(n, 80, "x chunks depth 4 + 2 x %d-row w boxes depth 2", w_rows); report(n, ms);
      }
    }
  }
  CK(cudaDeviceSynchronize());
  return 0;
}
