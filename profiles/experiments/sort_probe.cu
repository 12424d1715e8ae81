// Experiment: cost of the stable slot sort in different encodings.
#include <cstdio>
#include <cstdint>
#include <cub/device/device_radix_sort.cuh>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__host__ __device__ inline unsigned long long mix64(unsigned long long x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31; return x;
}
__global__ void gen(unsigned long long* tag, unsigned int* slot, unsigned int* row, int64_t n, unsigned int mask) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned int s = (unsigned int)mix64(i) & mask;
    tag[i] = ((unsigned long long)s << 32) | (unsigned long long)i; slot[i] = s; row[i] = (unsigned int)i;
  }
}
int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? atoll(argv[1]) : 250000000LL;
  unsigned long long *tag, *tag2; unsigned int *slot, *slot2, *row, *row2; void* tmp; size_t tmp_bytes = 3ULL << 30;
  CK(cudaMalloc(&tag, n * 8)); CK(cudaMalloc(&tag2, n * 8)); CK(cudaMalloc(&slot, n * 4)); CK(cudaMalloc(&slot2, n * 4));
  CK(cudaMalloc(&row, n * 4)); CK(cudaMalloc(&row2, n * 4)); CK(cudaMalloc(&tmp, tmp_bytes));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int bits : {8, 16, 20, 22, 24}) {
    gen<<<1184, 256>>>(tag, slot, row, n, (1u << bits) - 1); CK(cudaDeviceSynchronize());
    float ms;
    for (int it = 0; it < 2; ++it) {
      size_t need = 0; cub::DeviceRadixSort::SortKeys(nullptr, need, tag, tag2, n, 32, 32 + bits);
      if (need > tmp_bytes) { printf("tmp too small\n"); return 1; }
      cudaEventRecord(a); cub::DeviceRadixSort::SortKeys(tmp, need, tag, tag2, n, 32, 32 + bits); cudaEventRecord(b); CK(cudaEventSynchronize(b));
      cudaEventElapsedTime(&ms, a, b);
    }
    printf("bits %2d  SortKeys u64        : %.3f ms\n", bits, ms);
    for (int it = 0; it < 2; ++it) {
      size_t need = 0; cub::DeviceRadixSort::SortPairs(nullptr, need, slot, slot2, row, row2, n, 0, bits);
      cudaEventRecord(a); cub::DeviceRadixSort::SortPairs(tmp, need, slot, slot2, row, row2, n, 0, bits); cudaEventRecord(b); CK(cudaEventSynchronize(b));
      cudaEventElapsedTime(&ms, a, b);
    }
    printf("bits %2d  SortPairs u32,u32   : %.3f ms\n", bits, ms);
    for (int it = 0; it < 2; ++it) {
      size_t need = 0; cub::DeviceRadixSort::SortKeys(nullptr, need, slot, slot2, n, 0, bits);
      cudaEventRecord(a); cub::DeviceRadixSort::SortKeys(tmp, need, slot, slot2, n, 0, bits); cudaEventRecord(b); CK(cudaEventSynchronize(b));
      cudaEventElapsedTime(&ms, a, b);
    }
    printf("bits %2d  SortKeys u32 (ref)  : %.3f ms\n", bits, ms);
  }
  return 0;
}
