// Throughput of __match_any_sync against an 8-ballot equivalent (which lanes share my 8-bit digit), per SM.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/match_rate match_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned mix(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned* out, int iters) {
  unsigned d[16], acc = 0;
  for (int r = 0; r < 16; ++r) d[r] = mix(threadIdx.x * 16 + r + blockIdx.x * 4096);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const unsigned dig = (d[r] >> (it & 15)) & 255u;
      unsigned peers;
      if (MODE == 0) peers = __match_any_sync(0xffffffffu, dig);
      else {
        peers = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          const unsigned bal = __ballot_sync(0xffffffffu, (dig >> b) & 1u);
          peers &= ((dig >> b) & 1u) ? bal : ~bal;
        }
      }
      acc += __popc(peers) + (peers & 7u);
    }
  }
  out[blockIdx.x * 256 + threadIdx.x] = acc;
}
int main() {
  unsigned* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int iters = 2000;
  for (int ctas = 2; ctas <= 8; ctas *= 2)
    for (int mode = 0; mode < 2; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<148 * ctas, 256>>>(out, iters); else k<1><<<148 * ctas, 256>>>(out, iters);
        cudaEventRecord(b); cudaEventSynchronize(b);
      }
      float ms; cudaEventElapsedTime(&ms, a, b);
      const double warp_ops = double(ctas) * 8 * iters * 16;      // per SM
      printf("%s ctas/SM=%d: %.3f ms, %.1f cycles per warp-op per SM at 1.965 GHz\n", mode == 0 ? "match.any" : "8 ballots", ctas, ms,
             ms * 1e-3 * 1.965e9 / warp_ops);
    }
  return 0;
}
