// Experiment (not part of the product): why does the hash-insert pass miss L2?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
constexpr unsigned long long kEmpty = 0x8000000000000000ULL;
__host__ __device__ inline unsigned long long mix64(unsigned long long x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31; return x;
}
__global__ void gen_keys(unsigned long long* keys, int64_t n, unsigned long long universe) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    keys[i] = mix64(mix64(i + 12345) % universe + 777) >> 1;
}
__global__ void fill(unsigned long long* p, int64_t n, unsigned long long v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__device__ __forceinline__ unsigned long long pol_keep() { unsigned long long p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ unsigned long long pol_stream() { unsigned long long p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ unsigned long long ld_hint(const unsigned long long* p, unsigned long long pol) { unsigned long long v; asm volatile("ld.global.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol) : "memory"); return v; }

// MODE bits: 1 = hints, 2 = count red, 4 = tag store, 8 = 32-bit tags
template <int MODE>
__global__ void __launch_bounds__(256) insert(const unsigned long long* __restrict__ keys, int64_t n, unsigned long long* slot_key,
                                              unsigned int* counts, unsigned long long slots, unsigned long long* tag, unsigned int* sink) {
  const unsigned long long mask = slots - 1;
  const unsigned long long keep = pol_keep(), stream = pol_stream();
  unsigned int acc = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long k = (MODE & 1) ? ld_hint(keys + i, stream) : __ldcs(keys + i);
    unsigned long long h = mix64(k) & mask;
    while (true) {
      unsigned long long cur = (MODE & 1) ? ld_hint(slot_key + h, keep) : *(volatile unsigned long long*)(slot_key + h);
      if (cur == k) break;
      if (cur == kEmpty) {
        unsigned long long prev = atomicCAS(slot_key + h, kEmpty, k);
        if (prev == kEmpty || prev == k) break;
      }
      h = (h + 1) & mask;
    }
    acc += (unsigned)h;
    if (MODE & 4) {
      if (MODE & 8) __stcs(reinterpret_cast<unsigned int*>(tag) + i, (unsigned)h);
      else __stcs(tag + i, (h << 32) | (unsigned long long)i);
    }
    if (MODE & 2) atomicAdd(counts + h, 1u);
  }
  if (acc == 0xdeadbeef) *sink = acc;
}
template <int MODE>
float run(const unsigned long long* keys, int64_t n, unsigned long long* slot_key, unsigned int* counts, unsigned long long slots,
          unsigned long long* tag, unsigned int* sink, int grid) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9;
  for (int it = 0; it < 3; ++it) {
    fill<<<1184, 256>>>(slot_key, slots, kEmpty);
    cudaMemset(counts, 0, slots * 4);
    cudaEventRecord(a);
    insert<MODE><<<grid, 256>>>(keys, n, slot_key, counts, slots, tag, sink);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  return best;
}

// variant: U keys per thread, all first probes in flight together, lanes reconverge before the coalesced store
template <int MODE, int U>
__global__ void __launch_bounds__(256) insert2(const unsigned long long* __restrict__ keys, int64_t n, unsigned long long* slot_key,
                                               unsigned int* counts, unsigned long long slots, unsigned long long* tag, unsigned int* sink) {
  const unsigned long long mask = slots - 1;
  const int64_t tile = int64_t(blockDim.x) * U;
  const int64_t tiles = (n + tile - 1) / tile;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t base = t * tile + threadIdx.x;
    unsigned long long k[U], h[U], cur[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int64_t i = base + int64_t(u) * blockDim.x; k[u] = i < n ? __ldcs(keys + i) : 0ULL; }
#pragma unroll
    for (int u = 0; u < U; ++u) { h[u] = mix64(k[u]) & mask; cur[u] = *(volatile unsigned long long*)(slot_key + h[u]); }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + int64_t(u) * blockDim.x;
      if (i < n) {
        unsigned long long c = cur[u];
        while (true) {
          if (c == k[u]) break;
          if (c == kEmpty) {
            unsigned long long prev = atomicCAS(slot_key + h[u], kEmpty, k[u]);
            if (prev == kEmpty || prev == k[u]) break;
          }
          h[u] = (h[u] + 1) & mask;
          c = *(volatile unsigned long long*)(slot_key + h[u]);
        }
      }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + int64_t(u) * blockDim.x;
      if (i < n) {
        if (MODE & 4) {
          if (MODE & 8) __stcs(reinterpret_cast<unsigned int*>(tag) + i, (unsigned)h[u]);
          else __stcs(tag + i, (h[u] << 32) | (unsigned long long)i);
        }
        if (MODE & 2) atomicAdd(counts + h[u], 1u);
      }
    }
  }
}
template <int MODE, int U>
float run2(const unsigned long long* keys, int64_t n, unsigned long long* slot_key, unsigned int* counts, unsigned long long slots,
          unsigned long long* tag, unsigned int* sink, int grid) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9;
  for (int it = 0; it < 3; ++it) {
    fill<<<1184, 256>>>(slot_key, slots, kEmpty);
    cudaMemset(counts, 0, slots * 4);
    cudaEventRecord(a);
    insert2<MODE, U><<<grid, 256>>>(keys, n, slot_key, counts, slots, tag, sink);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  return best;
}
int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? atoll(argv[1]) : 250000000LL;
  unsigned long long *keys, *slot_key, *tag; unsigned int *counts, *sink;
  const unsigned long long max_slots = 1ULL << 24;
  CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&tag, n * 8)); CK(cudaMalloc(&slot_key, max_slots * 8)); CK(cudaMalloc(&counts, max_slots * 4)); CK(cudaMalloc(&sink, 4));
  for (unsigned long long universe : {1000ULL, 1000000ULL, 4000000ULL}) {
    unsigned long long slots = 1024; while (slots < 2 * universe) slots <<= 1;
    gen_keys<<<1184, 256>>>(keys, n, universe); CK(cudaDeviceSynchronize());
    printf("universe %llu slots %llu (table %.1f MB + counts %.1f MB)\n", universe, slots, slots * 8 / 1e6, slots * 4 / 1e6);
    printf("  probe only, no hints      : %.3f ms\n", run<0>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  probe only, hints         : %.3f ms\n", run<1>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  probe + count, no hints   : %.3f ms\n", run<2>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  probe + tag64, no hints   : %.3f ms\n", run<4>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  probe + tag32, no hints   : %.3f ms\n", run<12>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  all, no hints             : %.3f ms\n", run<6>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  all, hints(ld only)       : %.3f ms\n", run<7>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  v2 U=1 all                : %.3f ms\n", run2<6, 1>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=2 all                : %.3f ms\n", run2<6, 2>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=4 all                : %.3f ms\n", run2<6, 4>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=8 all                : %.3f ms\n", run2<6, 8>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=4 all grid 1184      : %.3f ms\n", run2<6, 4>(keys, n, slot_key, counts, slots, tag, sink, 1184));
    printf("  v2 U=4 all grid 4736      : %.3f ms\n", run2<6, 4>(keys, n, slot_key, counts, slots, tag, sink, 4736));
    printf("  v2 U=4 probe only         : %.3f ms\n", run2<0, 4>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=4 probe+tag64        : %.3f ms\n", run2<4, 4>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  v2 U=4 probe+tag32+count  : %.3f ms\n", run2<14, 4>(keys, n, slot_key, counts, slots, tag, sink, 2368));
    printf("  all, no hints, grid 592   : %.3f ms\n", run<6>(keys, n, slot_key, counts, slots, tag, sink, 592));
    printf("  all, no hints, grid 4736  : %.3f ms\n", run<6>(keys, n, slot_key, counts, slots, tag, sink, 4736));
  }
  CK(cudaDeviceSynchronize());
  return 0;
}
