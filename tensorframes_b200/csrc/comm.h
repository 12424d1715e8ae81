// Comm: this rank's end of the exchange steps of reduceBlocks and aggregate.
//
// The reference hands those steps to Spark: RDD.reduce pulls every partition's result row to the driver and merges
// pairwise (impl/DebugRowOps.scala:524-525, 741-750); the UDAF's partial buffers are hash-shuffled to their key's
// reducer (impl/DebugRowOps.scala:573-576).  With one partition per GPU of a box the same two exchanges run over
// NVLink: NCCL is loaded at run time (dlopen of libnccl.so.2 -- the library has no link-time dependency on it, a
// single-GPU caller never touches it) and driven from C++, so that a step costs no Python and no allocation.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "common.h"

namespace tfs {

constexpr int kCommIdBytes = 128;     // an ncclUniqueId

class Comm {
 public:
  // rank 0 of a job draws the id; the caller carries it to the other ranks (Spark broadcast, a file, torch.distributed)
  static void unique_id(void* out128);
  // one rank per process (or per thread): collective over all ranks of the job, like ncclCommInitRank
  static Comm* create(int device, int world, int rank, const void* id128);
  // one process driving n devices: all n communicators at once (ncclCommInitAll); exchanges must then be issued
  // from one thread per device
  static void create_all(int n, const int* devices, Comm** out);
  ~Comm();
  Comm(const Comm&) = delete;
  Comm& operator=(const Comm&) = delete;

  int world() const { return world_; }
  int rank() const { return rank_; }
  int device() const { return device_; }
  cudaStream_t side_stream() const { return side_; }

  // collectives, enqueued on `stream` (dtype = TFS_* code, red_op = R_SUM / R_MIN / R_MAX)
  void group_start();
  void group_end();
  void all_reduce(const void* send, void* recv, size_t count, int dtype, int red_op, cudaStream_t stream);
  void all_gather(const void* send, void* recv, size_t count, int dtype, cudaStream_t stream);
  void all_gather_bytes(const void* send, void* recv, size_t bytes, cudaStream_t stream);
  void send_bytes(const void* buf, size_t bytes, int peer, cudaStream_t stream);
  void recv_bytes(void* buf, size_t bytes, int peer, cudaStream_t stream);

  // ---- overlapped exchanges: a ring of small device slots so that an exchange on the side stream can still read a
  // step's partial rows while the launch stream is already producing the next step's
  static constexpr int kRing = 4;
  static constexpr size_t kSlotBytes = size_t(64) << 10;
  // next slot; `launch` first waits until the exchange that last read that slot is done
  char* ring_acquire(cudaStream_t launch, int* slot);
  // the slot has been filled on `launch`: the side stream may start
  void ring_produced(int slot, cudaStream_t launch);
  // the exchange enqueued on the side stream is the last reader of the slot
  void ring_consumed(int slot);
  // `stream` waits for every exchange enqueued on the side stream so far
  void wait_side(cudaStream_t stream);

  // small pinned + device scratch for the per-peer counts of the aggregate exchange: [world * world] int64 each
  int64_t* counts_dev() const { return counts_dev_; }
  int64_t* counts_host() const { return counts_host_; }

 private:
  Comm() = default;
  void init_resources();
  int world_ = 1, rank_ = 0, device_ = 0;
  void* nccl_ = nullptr;
  cudaStream_t side_ = nullptr;
  char* ring_ = nullptr;
  cudaEvent_t produced_[kRing] = {}, consumed_[kRing] = {};
  bool used_[kRing] = {};
  int next_ = 0;
  cudaEvent_t side_done_ = nullptr;
  bool side_pending_ = false;
  int64_t* counts_dev_ = nullptr;
  int64_t* counts_host_ = nullptr;
};

}  // namespace tfs
