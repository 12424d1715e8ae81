// Experiment: per-role timeline of the CTA-pair matmul kernel (CTA 0 and CTA 1 of pair 0).
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
constexpr int kTraceN = 128;
__device__ long long g_trace[2][8][kTraceN];
#define TFS_MM_TRACE(ev, idx) do { if (blockIdx.x < 2 && (idx) < kTraceN) g_trace[blockIdx.x][ev][idx] = clock64(); } while (0)
#include "../../tensorframes_b200/csrc/kernels_matmul_tc.cu"
int main() {
  const int64_t M = 16777216 / 4; const int N = 256, K = 256;
  float *x, *w, *hi, *lo, *bias, *out;
  cudaMalloc(&x, M * K * 4); cudaMalloc(&w, K * N * 4); cudaMalloc(&hi, K * N * 4); cudaMalloc(&lo, K * N * 4); cudaMalloc(&bias, N * 4); cudaMalloc(&out, M * N * 4);
  cudaMemset(x, 0x3c, M * K * 4); cudaMemset(w, 0x3c, K * N * 4); cudaMemset(bias, 0, N * 4);
  tfs::DeviceInfo dev; dev.sm_count = 148;
  tfs::launch_split_weights(w, false, K, N, hi, lo, 0);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int it = 0; it < 3; ++it) {
    cudaEventRecord(a);
    tfs::launch_matmul_tc(x, hi, lo, bias, out, M, N, K, dev, 0);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); printf("run %d: %.3f ms (%s)\n", it, ms, cudaGetErrorString(cudaGetLastError()));
  }
  static long long h[2][8][kTraceN];
  cudaMemcpyFromSymbol(h, g_trace, sizeof(h));
  const char* names[8] = {"P:xempty-ok", "M:x+w-ok", "M:split-ok", "M:issued", "S:xfull-ok", "S:arrived", "E:full-ok", "E:drained"};
  for (int c = 0; c < 2; ++c) {
    long long t0 = h[c][0][0];
    printf("CTA %d (clk relative to first producer event)\n it  ", c);
    for (int e = 0; e < 6; ++e) printf("%12s", names[e]);
    printf("\n");
    for (int i = 0; i < 40; ++i) {
      printf("%3d  ", i);
      for (int e = 0; e < 6; ++e) printf("%12lld", h[c][e][i] ? h[c][e][i] - t0 : -1);
      printf("\n");
    }
    printf(" tile %12s%12s\n", names[6], names[7]);
    for (int i = 0; i < 6; ++i) printf("%3d   %12lld%12lld\n", i, h[c][6][i] - t0, h[c][7][i] - t0);
  }
  return 0;
}
