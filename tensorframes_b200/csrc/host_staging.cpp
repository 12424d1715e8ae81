// K0, pageable side: large copies between PAGEABLE host memory and the device.
//
// A JVM caller hands over direct ByteBuffers, numpy hands over malloc'ed arrays: pageable memory, which the driver
// copies through its own bounce buffer at ~11 GB/s on the B200 box (measured; pinned memory moves at 55 GB/s, and
// page-locking the caller's buffer for one call costs as much as the slow copy).  Here a few host threads each run
// their own small pipeline instead: memcpy a chunk into a pinned slot of a process-wide ring, queue the DMA on a
// private stream, move on to the next chunk while it flies.  The caller's stream is ordered after the helpers.
// Pinned callers (tfs_host_alloc) never come here; small copies do not either.
//
// Stands in for the copies inside Tensor.create(shape, buffer) / Tensor.writeTo(buffer)
// (reference: impl/datatypes.scala:344-347, 365-370).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

#include "kernels.h"

namespace tfs {
namespace {

constexpr size_t kChunk = size_t(8) << 20;       // bytes per pinned slot
constexpr int kSlotsPerWorker = 2;
constexpr int kMaxWorkers = 6;

struct Ring {
  std::mutex mu;                  // one staged copy at a time per device
  char* pinned = nullptr;         // kMaxWorkers * kSlotsPerWorker slots
  cudaStream_t stream[kMaxWorkers] = {};
  cudaEvent_t slot_free[kMaxWorkers][kSlotsPerWorker] = {};
  cudaEvent_t fork = nullptr, join[kMaxWorkers] = {};
  bool ready = false;
};
Ring g_rings[64];

int worker_count() {
  const unsigned hw = std::thread::hardware_concurrency();
  return int(std::max(2u, std::min<unsigned>(kMaxWorkers, hw ? hw / 2 : 2)));
}

void ensure_ring(Ring& r) {
  if (r.ready) return;
  TFS_CUDA(cudaMallocHost(reinterpret_cast<void**>(&r.pinned), size_t(kMaxWorkers) * kSlotsPerWorker * kChunk));
  for (int w = 0; w < kMaxWorkers; ++w) {
    TFS_CUDA(cudaStreamCreateWithFlags(&r.stream[w], cudaStreamNonBlocking));
    TFS_CUDA(cudaEventCreateWithFlags(&r.join[w], cudaEventDisableTiming));
    for (int s = 0; s < kSlotsPerWorker; ++s) TFS_CUDA(cudaEventCreateWithFlags(&r.slot_free[w][s], cudaEventDisableTiming));
  }
  TFS_CUDA(cudaEventCreateWithFlags(&r.fork, cudaEventDisableTiming));
  r.ready = true;
}

}  // namespace

bool staged_copy_pays(size_t bytes) {
  static const bool off = [] { const char* e = std::getenv("TFS_STAGED_COPY"); return e && std::string(e) == "off"; }();   // A/B switch
  return !off && bytes >= (size_t(64) << 20);
}

// dst/src: one is device memory on `device`, the other PAGEABLE host memory.  On return the copy has been handed
// to the GPU (host -> device: the host buffer may be reused; device -> host: the data is in dst), and `stream` is
// ordered after it.
void staged_copy(int device, void* dst, const void* src, size_t bytes, bool to_device, cudaStream_t stream) {
  Ring& r = g_rings[device & 63];
  std::lock_guard<std::mutex> lock(r.mu);
  ensure_ring(r);
  const int workers = worker_count();
  const size_t n_chunks = (bytes + kChunk - 1) / kChunk;
  // the helpers start after whatever the caller's stream has queued (it may still be using the destination)
  TFS_CUDA(cudaEventRecord(r.fork, stream));
  std::vector<cudaError_t> status(size_t(workers), cudaSuccess);
  auto body = [&](int w) {
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(r.stream[w], r.fork, 0);
    size_t k = 0;
    for (size_t c = size_t(w); c < n_chunks && e == cudaSuccess; c += size_t(workers), ++k) {
      const int s = int(k % kSlotsPerWorker);
      char* slot = r.pinned + (size_t(w) * kSlotsPerWorker + size_t(s)) * kChunk;
      const size_t off = c * kChunk, len = std::min(kChunk, bytes - off);
      if (to_device) {
        // the DMA that last read this slot -- earlier in this call or at the end of the previous one -- must be done
        // (an event that was never recorded is complete)
        e = cudaEventSynchronize(r.slot_free[w][s]);
        if (e != cudaSuccess) break;
        std::memcpy(slot, static_cast<const char*>(src) + off, len);
        e = cudaMemcpyAsync(static_cast<char*>(dst) + off, slot, len, cudaMemcpyHostToDevice, r.stream[w]);
        if (e == cudaSuccess) e = cudaEventRecord(r.slot_free[w][s], r.stream[w]);
      } else {
        // device -> host: keep one DMA in flight while the previous slot is drained
        e = cudaMemcpyAsync(slot, static_cast<const char*>(src) + off, len, cudaMemcpyDeviceToHost, r.stream[w]);
        if (e == cudaSuccess) e = cudaEventRecord(r.slot_free[w][s], r.stream[w]);
        if (k >= 1 && e == cudaSuccess) {
          const size_t pc = c - size_t(workers);
          const int ps = int((k - 1) % kSlotsPerWorker);
          char* pslot = r.pinned + (size_t(w) * kSlotsPerWorker + size_t(ps)) * kChunk;
          e = cudaEventSynchronize(r.slot_free[w][ps]);
          if (e == cudaSuccess) std::memcpy(static_cast<char*>(dst) + pc * kChunk, pslot, std::min(kChunk, bytes - pc * kChunk));
        }
      }
    }
    if (!to_device && e == cudaSuccess && k >= 1) {      // drain the last slot
      const size_t lc = size_t(w) + (k - 1) * size_t(workers);
      const int ls = int((k - 1) % kSlotsPerWorker);
      char* lslot = r.pinned + (size_t(w) * kSlotsPerWorker + size_t(ls)) * kChunk;
      e = cudaEventSynchronize(r.slot_free[w][ls]);
      if (e == cudaSuccess) std::memcpy(static_cast<char*>(dst) + lc * kChunk, lslot, std::min(kChunk, bytes - lc * kChunk));
    }
    if (e == cudaSuccess) e = cudaEventRecord(r.join[w], r.stream[w]);
    status[size_t(w)] = e;
  };
  std::vector<std::thread> threads;
  for (int w = 1; w < workers; ++w) threads.emplace_back(body, w);
  body(0);
  for (std::thread& t : threads) t.join();
  for (int w = 0; w < workers; ++w) TFS_CUDA(status[size_t(w)]);
  for (int w = 0; w < workers; ++w) TFS_CUDA(cudaStreamWaitEvent(stream, r.join[w], 0));
}

}  // namespace tfs
