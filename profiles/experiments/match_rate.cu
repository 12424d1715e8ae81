// Throughput of __match_any_sync against an 8-ballot equivalent (which lanes share my 8-bit digit), per SM.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/match_rate match_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned mix(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned* out, int iters) {
  __shared__ unsigned table[2][8][256];      // MODE 4: per warp, lane masks by digit, two copies used in turn
  for (int i = threadIdx.x; i < 2 * 8 * 256; i += 256) (&table[0][0][0])[i] = 0u;
  __syncthreads();
  const unsigned lane_bit = 1u << (threadIdx.x & 31);
  unsigned* const t0 = table[0][threadIdx.x >> 5];
  unsigned* const t1 = table[1][threadIdx.x >> 5];
  unsigned prev_dig = 0;
  unsigned d[16], acc = 0;
  for (int r = 0; r < 16; ++r) d[r] = mix(threadIdx.x * 16 + r + blockIdx.x * 4096);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const unsigned dig = (d[r] >> (it & 15)) & 255u;
      unsigned peers;
      if (MODE == 0 || (MODE == 2 && r < 5)) peers = __match_any_sync(0xffffffffu, dig);
      else if (MODE == 4) {      // lane masks collected in shared memory
        unsigned* cur = (r & 1) ? t1 : t0;
        unsigned* old = (r & 1) ? t0 : t1;
        atomicOr(cur + dig, lane_bit);
        __syncwarp();
        peers = cur[dig];
        old[prev_dig] = 0u;
        prev_dig = dig;
        __syncwarp();
      } else if (MODE == 3) {      // the votes alone
        peers = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) peers ^= __ballot_sync(0xffffffffu, (dig >> b) & 1u);
      } else {
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
  for (int ctas = 2; ctas <= 2; ctas *= 2)
    for (int mode = 0; mode < 5; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<148 * ctas, 256>>>(out, iters);
        else if (mode == 1) k<1><<<148 * ctas, 256>>>(out, iters);
        else if (mode == 2) k<2><<<148 * ctas, 256>>>(out, iters);
        else if (mode == 3) k<3><<<148 * ctas, 256>>>(out, iters);
        else k<4><<<148 * ctas, 256>>>(out, iters);
        cudaEventRecord(b); cudaEventSynchronize(b);
      }
      float ms; cudaEventElapsedTime(&ms, a, b);
      const double warp_ops = double(ctas) * 8 * iters * 16;      // per SM
      printf("%s ctas/SM=%d: %.3f ms, %.1f cycles per warp-op per SM at 1.965 GHz\n", mode == 0 ? "match.any" : mode == 1 ? "8 ballots" : mode == 2 ? "5 of 16 match.any, 11 ballots" : mode == 3 ? "8 votes, no masks" : "shared-memory lane masks", ctas, ms,
             ms * 1e-3 * 1.965e9 / warp_ops);
    }
  return 0;
}
