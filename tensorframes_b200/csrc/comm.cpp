// Comm: NCCL loaded at run time and the small amount of state the exchanges need.  See comm.h.
#include "comm.h"

#include <dlfcn.h>

#include <cstdlib>
#include <memory>
#include <mutex>
#include <vector>

#include "graph.h"
#include "kernels.h"

namespace tfs {
namespace {

// The handful of NCCL entry points the exchanges use, declared here so that building the library needs no nccl.h.
struct NcclId { char internal[kCommIdBytes]; };
using NcclComm = void*;
enum : int { kNcclInt8 = 0, kNcclInt32 = 2, kNcclInt64 = 4, kNcclFloat32 = 7, kNcclFloat64 = 8 };
enum : int { kNcclSum = 0, kNcclMax = 2, kNcclMin = 3 };

struct NcclApi {
  void* handle = nullptr;
  std::string error;
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
  int (*CommInitAll)(NcclComm*, int, const int*) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
};

template <class F>
bool bind(void* h, const char* name, F& fn, std::string& err) {
  fn = reinterpret_cast<F>(dlsym(h, name));
  if (fn == nullptr) err = std::string("NCCL symbol ") + name + " not found";
  return fn != nullptr;
}

const NcclApi& nccl() {
  static const NcclApi api = [] {
    NcclApi a;
    // a process that already holds an NCCL (torch ships its own) gets that one: same soname, same handle
    std::vector<std::string> names;
    if (const char* e = std::getenv("TFS_NCCL_LIB")) names.push_back(e);
    names.push_back("libnccl.so.2");
    names.push_back("libnccl.so");
    for (const std::string& n : names) {
      a.handle = dlopen(n.c_str(), RTLD_NOW | RTLD_LOCAL);
      if (a.handle) break;
      const char* why = dlerror();
      a.error += n + ": " + (why ? why : "not found") + "; ";
    }
    if (!a.handle) return a;
    a.error.clear();
    void* h = a.handle;
    bool ok = bind(h, "ncclGetUniqueId", a.GetUniqueId, a.error) && bind(h, "ncclCommInitRank", a.CommInitRank, a.error) &&
              bind(h, "ncclCommInitAll", a.CommInitAll, a.error) && bind(h, "ncclCommDestroy", a.CommDestroy, a.error) &&
              bind(h, "ncclGetErrorString", a.GetErrorString, a.error) && bind(h, "ncclGroupStart", a.GroupStart, a.error) &&
              bind(h, "ncclGroupEnd", a.GroupEnd, a.error) && bind(h, "ncclAllReduce", a.AllReduce, a.error) &&
              bind(h, "ncclAllGather", a.AllGather, a.error) && bind(h, "ncclSend", a.Send, a.error) &&
              bind(h, "ncclRecv", a.Recv, a.error);
    if (!ok) a.handle = nullptr;
    return a;
  }();
  TFS_CHECK(api.handle != nullptr, TFS_ERR_UNSUPPORTED,
            "the multi-GPU exchanges need NCCL (libnccl.so.2; TFS_NCCL_LIB overrides the name): " + api.error);
  return api;
}

void nccl_check(int r, const char* what) {
  if (r != 0) {
    const char* s = nccl().GetErrorString ? nccl().GetErrorString(r) : "?";
    fail(TFS_ERR_RUNTIME, std::string(what) + ": NCCL error " + std::to_string(r) + " (" + s + ")");
  }
}
#define TFS_NCCL(call) nccl_check((call), #call)

int nccl_dtype(int dtype) {
  switch (dtype) {
    case TFS_FLOAT: return kNcclFloat32;
    case TFS_DOUBLE: return kNcclFloat64;
    case TFS_INT32: return kNcclInt32;
    case TFS_INT64: return kNcclInt64;
  }
  fail(TFS_ERR_UNSUPPORTED, "exchange: unsupported dtype");
}

struct DeviceScope {
  int prev = -1;
  explicit DeviceScope(int device) {
    TFS_CUDA(cudaGetDevice(&prev));
    if (prev != device) TFS_CUDA(cudaSetDevice(device));
    else prev = -1;
  }
  ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

void Comm::unique_id(void* out128) {
  TFS_CHECK(out128 != nullptr, TFS_ERR_INVALID, "null id buffer");
  NcclId id{};
  TFS_NCCL(nccl().GetUniqueId(&id));
  memcpy(out128, id.internal, kCommIdBytes);
}

void Comm::init_resources() {
  TFS_CUDA(cudaStreamCreateWithFlags(&side_, cudaStreamNonBlocking));
  TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&ring_), kRing * kSlotBytes));
  for (int i = 0; i < kRing; ++i) {
    TFS_CUDA(cudaEventCreateWithFlags(&produced_[i], cudaEventDisableTiming));
    TFS_CUDA(cudaEventCreateWithFlags(&consumed_[i], cudaEventDisableTiming));
  }
  TFS_CUDA(cudaEventCreateWithFlags(&side_done_, cudaEventDisableTiming));
  const size_t n = size_t(world_) * size_t(world_) * sizeof(int64_t);
  TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&counts_dev_), n));
  TFS_CUDA(cudaMallocHost(reinterpret_cast<void**>(&counts_host_), n));
}

Comm* Comm::create(int device, int world, int rank, const void* id128) {
  TFS_CHECK(world >= 1 && rank >= 0 && rank < world, TFS_ERR_INVALID, "communicator: rank must be in [0, world)");
  TFS_CHECK(id128 != nullptr, TFS_ERR_INVALID, "communicator: null id");
  int count = 0;
  TFS_CHECK(cudaGetDeviceCount(&count) == cudaSuccess && device >= 0 && device < count, TFS_ERR_CUDA,
            "communicator: device ordinal " + std::to_string(device) + " is not available");
  DeviceScope scope(device);
  std::unique_ptr<Comm> c(new Comm());
  c->world_ = world;
  c->rank_ = rank;
  c->device_ = device;
  NcclId id{};
  memcpy(id.internal, id128, kCommIdBytes);
  NcclComm nc = nullptr;
  TFS_NCCL(nccl().CommInitRank(&nc, world, id, rank));
  c->nccl_ = nc;
  c->init_resources();
  return c.release();
}

void Comm::create_all(int n, const int* devices, Comm** out) {
  TFS_CHECK(n >= 1 && devices != nullptr && out != nullptr, TFS_ERR_INVALID, "communicators: bad arguments");
  std::vector<NcclComm> ncs(size_t(n), nullptr);
  TFS_NCCL(nccl().CommInitAll(ncs.data(), n, devices));
  std::vector<std::unique_ptr<Comm>> cs;
  for (int i = 0; i < n; ++i) {
    DeviceScope scope(devices[i]);
    std::unique_ptr<Comm> c(new Comm());
    c->world_ = n;
    c->rank_ = i;
    c->device_ = devices[i];
    c->nccl_ = ncs[size_t(i)];
    c->init_resources();
    cs.push_back(std::move(c));
  }
  for (int i = 0; i < n; ++i) out[i] = cs[size_t(i)].release();
}

Comm::~Comm() {
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(device_);
  if (side_) cudaStreamSynchronize(side_);
  if (nccl_) nccl().CommDestroy(nccl_);
  for (int i = 0; i < kRing; ++i) {
    if (produced_[i]) cudaEventDestroy(produced_[i]);
    if (consumed_[i]) cudaEventDestroy(consumed_[i]);
  }
  if (side_done_) cudaEventDestroy(side_done_);
  if (ring_) cudaFree(ring_);
  if (counts_dev_) cudaFree(counts_dev_);
  if (counts_host_) cudaFreeHost(counts_host_);
  if (side_) cudaStreamDestroy(side_);
  if (prev >= 0) cudaSetDevice(prev);
}

void Comm::group_start() { TFS_NCCL(nccl().GroupStart()); }
void Comm::group_end() { TFS_NCCL(nccl().GroupEnd()); }

void Comm::all_reduce(const void* send, void* recv, size_t count, int dtype, int red_op, cudaStream_t stream) {
  const int op = red_op == R_SUM ? kNcclSum : (red_op == R_MIN ? kNcclMin : kNcclMax);
  TFS_NCCL(nccl().AllReduce(send, recv, count, nccl_dtype(dtype), op, nccl_, stream));
}
void Comm::all_gather(const void* send, void* recv, size_t count, int dtype, cudaStream_t stream) {
  TFS_NCCL(nccl().AllGather(send, recv, count, nccl_dtype(dtype), nccl_, stream));
}
void Comm::all_gather_bytes(const void* send, void* recv, size_t bytes, cudaStream_t stream) {
  TFS_NCCL(nccl().AllGather(send, recv, bytes, kNcclInt8, nccl_, stream));
}
void Comm::send_bytes(const void* buf, size_t bytes, int peer, cudaStream_t stream) {
  TFS_NCCL(nccl().Send(buf, bytes, kNcclInt8, peer, nccl_, stream));
}
void Comm::recv_bytes(void* buf, size_t bytes, int peer, cudaStream_t stream) {
  TFS_NCCL(nccl().Recv(buf, bytes, kNcclInt8, peer, nccl_, stream));
}

char* Comm::ring_acquire(cudaStream_t launch, int* slot) {
  const int s = next_;
  next_ = (next_ + 1) % kRing;
  if (used_[s]) TFS_CUDA(cudaStreamWaitEvent(launch, consumed_[s], 0));
  used_[s] = true;
  *slot = s;
  return ring_ + size_t(s) * kSlotBytes;
}
void Comm::ring_produced(int slot, cudaStream_t launch) {
  TFS_CUDA(cudaEventRecord(produced_[slot], launch));
  TFS_CUDA(cudaStreamWaitEvent(side_, produced_[slot], 0));
}
void Comm::ring_consumed(int slot) {
  TFS_CUDA(cudaEventRecord(consumed_[slot], side_));
  side_pending_ = true;
}
void Comm::wait_side(cudaStream_t stream) {
  if (!side_pending_) return;
  TFS_CUDA(cudaEventRecord(side_done_, side_));
  TFS_CUDA(cudaStreamWaitEvent(stream, side_done_, 0));
}

}  // namespace tfs
