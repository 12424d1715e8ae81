// Session: planning and execution. See session.h.
#include "session.h"

#include <mutex>

#include <algorithm>
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>

namespace tfs {
namespace {

constexpr size_t kAlign = 256;
size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

bool is_elt(const Node& n) {
  return n.cls == OpClass::Unary || n.cls == OpClass::Binary || n.cls == OpClass::Fill;
}
// nodes whose value is a small integer vector computed on the host from shapes (no kernel unless someone reads it as data)
bool is_host_value(const Node& n) {
  return n.cls == OpClass::ShapeOf || n.cls == OpClass::StridedSlice || n.cls == OpClass::Pack;
}
bool is_like_fill(const Node& n) { return n.cls == OpClass::Unary && (n.elt_op == E_ONES_LIKE || n.elt_op == E_ZEROS_LIKE); }
bool is_compute_barrier(const Node& n) {
  return n.cls == OpClass::Reduce || n.cls == OpClass::MatMul || n.cls == OpClass::ExpandDims || n.cls == OpClass::Tile ||
         n.cls == OpClass::ArgMinMax || n.cls == OpClass::SegmentSum || is_host_value(n);
}

// data operands of an element-wise node (Fill's dims input is a static attribute, not data)
std::vector<int> elt_operands(const Node& n) {
  if (n.cls == OpClass::Fill) return {n.in[1]};
  if (is_like_fill(n)) return {};            // OnesLike / ZerosLike take only the operand's shape
  return n.in;
}

bool matmul_force_simt() {
  static const bool v = [] { const char* e = std::getenv("TFS_MATMUL_IMPL"); return e && std::string(e) == "simt"; }();
  return v;
}

bool is_pinned_host_pointer(const void* p) {
  if (p == nullptr) return false;
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

struct Step {
  enum Kind { ELT, REDUCE_AXIS0, REDUCE_ROWS, REDUCE_MID, COPY, MATMUL, HOSTVAL, TILE, ARGMINMAX, SEGSUM } kind = ELT;
  std::vector<int> nodes;   // ELT: the group (topological order); others: one node
  Shape shape;              // ELT: iteration shape
  int64_t outer = 0, inner = 0;  // reductions, ARGMINMAX
  int64_t len = 0;               // ARGMINMAX: length of the searched axis; SEGSUM: number of segments
  int out_node = -1, bias_node = -1;  // MATMUL with a fused "+ bias[N]" epilogue: the Add node and its bias operand
};

}  // namespace

struct Session::Plan {
  std::vector<int> needed;            // topological
  std::vector<Shape> shape;           // by node index
  std::vector<int> feed_of;           // by node index: feed position or -1
  std::vector<int> fetch_nodes;
  std::vector<Step> steps;            // execution order
  std::vector<int> step_of;           // by node index, -1 = not computed by a step
  std::vector<char> materialize;      // by node index: output needs a device buffer
  // row mode (mapRows): the graph describes ONE row; tensors that depend on a placeholder carry an extra
  // leading dimension of n_batch rows at execution time
  bool rows = false;
  int64_t n_batch = 0;
  std::vector<char> batched;          // by node index
  std::vector<Shape> cshape;          // by node index: per-row (cell) shape
  std::vector<std::vector<int64_t>> ival;   // by node index: integer value of shape-derived nodes (has_ival)
  std::vector<char> has_ival;
  bool may_raise = false;             // some step can set the device error flag (integer division, segment ids)
};

// ---------------------------------------------------------------------------
// Per-device pools.  The reference opens a Graph + Session for every partition (impl/DebugRowOps.scala:783-802), so
// a drop-in caller creates and destroys sessions at that rate: creation must not cost device allocations.  What a
// session needs on every run -- an error flag on the device with its pinned host mirror, and a modest workspace -- is
// borrowed from these pools and handed back by the destructor; the device properties are read once.
// ---------------------------------------------------------------------------
namespace {
constexpr int kMaxDevices = 64;
constexpr size_t kPooledWorkspaceMax = size_t(64) << 20;   // larger workspaces are freed, not parked
constexpr size_t kPooledWorkspaces = 4;

struct ErrSlot { int32_t* dev; int32_t* host; };
struct DevicePool {
  std::mutex mu;
  bool probed = false;
  int cc_major = 0, sm_count = 0;
  std::vector<ErrSlot> free_slots;
  std::vector<std::pair<char*, size_t>> free_ws;
};
DevicePool g_pools[kMaxDevices];

ErrSlot borrow_err_slot(DevicePool& p) {
  std::lock_guard<std::mutex> lock(p.mu);
  if (p.free_slots.empty()) {   // a batch of 32 slots, 256 bytes apart on the device, one cache line apart on the host
    constexpr int kBatch = 32;
    char* d = nullptr;
    char* h = nullptr;
    TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&d), kBatch * kAlign));
    TFS_CUDA(cudaMemset(d, 0, kBatch * kAlign));
    TFS_CUDA(cudaMallocHost(reinterpret_cast<void**>(&h), kBatch * 64));
    memset(h, 0, kBatch * 64);
    for (int i = 0; i < kBatch; ++i)
      p.free_slots.push_back(ErrSlot{reinterpret_cast<int32_t*>(d + size_t(i) * kAlign), reinterpret_cast<int32_t*>(h + size_t(i) * 64)});
  }
  ErrSlot s = p.free_slots.back();
  p.free_slots.pop_back();
  return s;
}
}  // namespace

Session::Session(const Graph& g, int device) : g_(g) { dev_.device = device; }

Session::~Session() {
  if (!device_ready_) return;
  cudaSetDevice(dev_.device);
  // everything this session enqueued, on whichever stream: the event outlives a caller-provided stream that has
  // been destroyed in the meantime, the stream handle itself would not
  if (work_event_) cudaEventSynchronize(work_event_);
  cudaStreamSynchronize(own_stream_);
  for (auto& kv : const_dev_) cudaFree(kv.second);
  for (auto& kv : tc_weights_) { cudaFree(kv.second.first); cudaFree(kv.second.second); }
  DevicePool& pool = g_pools[dev_.device];
  {
    std::lock_guard<std::mutex> lock(pool.mu);
    if (ws_ && ws_cap_ <= kPooledWorkspaceMax && pool.free_ws.size() < kPooledWorkspaces) {
      pool.free_ws.emplace_back(ws_, ws_cap_);
      ws_ = nullptr;
    }
    if (err_flag_) {
      *err_host_ = 0;
      cudaMemsetAsync(err_flag_, 0, sizeof(int32_t), own_stream_);     // the next borrower finds a clean flag
      cudaStreamSynchronize(own_stream_);
      pool.free_slots.push_back(ErrSlot{err_flag_, err_host_});
    }
  }
  if (ws_) cudaFree(ws_);
  if (red_scratch_) cudaFree(red_scratch_);
  if (seg_scratch_) cudaFree(seg_scratch_);
  for (int k = 0; k < 4; ++k)
    if (xchg_[k]) cudaFree(xchg_[k]);
  for (int k = 0; k < 2; ++k)
    if (pipe_stream_[k]) cudaStreamDestroy(pipe_stream_[k]);
  if (pipe_event_) cudaEventDestroy(pipe_event_);
  if (work_event_) cudaEventDestroy(work_event_);
  if (own_stream_) cudaStreamDestroy(own_stream_);
}

void Session::ensure_device() {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    fail(TFS_ERR_CUDA, "no CUDA device is available: libtfsb200 has no CPU execution path");
  }
  TFS_CHECK(dev_.device >= 0 && dev_.device < count, TFS_ERR_INVALID,
            "device ordinal " + std::to_string(dev_.device) + " out of range (" + std::to_string(count) + " devices)");
  TFS_CUDA(cudaSetDevice(dev_.device));
  if (device_ready_) return;
  TFS_CHECK(dev_.device < kMaxDevices, TFS_ERR_UNSUPPORTED, "device ordinal above the supported maximum");
  DevicePool& pool = g_pools[dev_.device];
  {
    std::lock_guard<std::mutex> lock(pool.mu);
    if (!pool.probed) {       // (cudaGetDeviceProperties costs milliseconds; two attributes are all that is needed)
      TFS_CUDA(cudaDeviceGetAttribute(&pool.cc_major, cudaDevAttrComputeCapabilityMajor, dev_.device));
      TFS_CUDA(cudaDeviceGetAttribute(&pool.sm_count, cudaDevAttrMultiProcessorCount, dev_.device));
      pool.probed = true;
    }
  }
  TFS_CHECK(pool.cc_major >= 10, TFS_ERR_CUDA,
            "device " + std::to_string(dev_.device) + " has compute capability " + std::to_string(pool.cc_major) +
                ".x; this library is built for sm_100a (B200) only");
  dev_.sm_count = pool.sm_count;
  TFS_CUDA(cudaStreamCreateWithFlags(&own_stream_, cudaStreamNonBlocking));
  TFS_CUDA(cudaEventCreateWithFlags(&work_event_, cudaEventDisableTiming));
  if (!stream_) stream_ = own_stream_;
  const ErrSlot slot = borrow_err_slot(pool);
  err_flag_ = slot.dev;
  err_host_ = slot.host;
  device_ready_ = true;
}

// Runs that only enqueue (device-resident feeds and fetches) return with work pending, and every run shares the
// workspace, the reduction scratch and the error flag: a new stream first waits for what the old one was given.
void Session::set_stream(cudaStream_t s) {
  cudaStream_t next = s ? s : own_stream_;
  if (device_ready_ && next != nullptr && next != stream_ && work_event_) TFS_CUDA(cudaStreamWaitEvent(next, work_event_, 0));
  stream_ = next;
}

void Session::mark_work() {
  if (work_event_) TFS_CUDA(cudaEventRecord(work_event_, stream_));
}

void Session::sync() {
  if (!device_ready_) return;
  TFS_CUDA(cudaSetDevice(dev_.device));
  check_device_error();
}

void Session::check_device_error() {
  TFS_CUDA(cudaMemcpyAsync(err_host_, err_flag_, sizeof(int32_t), cudaMemcpyDeviceToHost, stream_));
  TFS_CUDA(cudaStreamSynchronize(stream_));
  if (*err_host_ != 0) {
    const int32_t code = *err_host_;
    *err_host_ = 0;
    TFS_CUDA(cudaMemsetAsync(err_flag_, 0, sizeof(int32_t), stream_));
    TFS_CUDA(cudaStreamSynchronize(stream_));
    fail(TFS_ERR_RUNTIME, code == 2 ? "UnsortedSegmentSum: segment ids out of range" : "Integer division by zero");
  }
}

void Session::reserve_workspace(size_t bytes) {
  if (bytes <= ws_cap_) return;
  if (!ws_ && bytes <= kPooledWorkspaceMax) {     // a parked workspace of a finished session
    DevicePool& pool = g_pools[dev_.device];
    std::lock_guard<std::mutex> lock(pool.mu);
    for (size_t i = 0; i < pool.free_ws.size(); ++i) {
      if (pool.free_ws[i].second >= bytes) {
        ws_ = pool.free_ws[i].first;
        ws_cap_ = pool.free_ws[i].second;
        pool.free_ws.erase(pool.free_ws.begin() + long(i));
        return;
      }
    }
  }
  TFS_CUDA(cudaStreamSynchronize(stream_));
  if (ws_) TFS_CUDA(cudaFree(ws_));
  ws_ = nullptr;
  ws_cap_ = 0;
  size_t cap = std::max(align_up(bytes + bytes / 8), size_t(1) << 20);   // at least 1 MiB, so a parked one is reusable
  TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&ws_), cap));
  ws_cap_ = cap;
}

// Host <-> device copies of caller data.  Pinned buffers and small blocks are plain async copies; large pageable
// blocks (a JVM direct buffer, a numpy array) take the staged path of host_staging.cpp, ~3x the driver's own.
void Session::copy_in(void* dst_dev, const void* src_host, size_t bytes) {
  if (bytes == 0) return;
  if (staged_copy_pays(bytes) && !is_pinned_host_pointer(src_host)) staged_copy(dev_.device, dst_dev, src_host, bytes, true, stream_);
  else TFS_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, stream_));
}
void Session::copy_out(void* dst_host, const void* src_dev, size_t bytes) {
  if (bytes == 0) return;
  if (staged_copy_pays(bytes) && !is_pinned_host_pointer(dst_host)) staged_copy(dev_.device, dst_host, src_dev, bytes, false, stream_);
  else TFS_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, stream_));
}

void* Session::const_device(int node) {
  auto it = const_dev_.find(node);
  if (it != const_dev_.end()) return it->second;
  const TensorValue& v = g_.nodes[size_t(node)].value;
  void* p = nullptr;
  size_t bytes = std::max<size_t>(v.bytes.size(), 16);
  TFS_CUDA(cudaMalloc(&p, bytes));
  if (!v.bytes.empty())
    TFS_CUDA(cudaMemcpyAsync(p, v.bytes.data(), v.bytes.size(), cudaMemcpyHostToDevice, stream_));
  // v.bytes is pageable: the copy above has been staged when the call returns.
  const_dev_[node] = p;
  return p;
}

// ---------------------------------------------------------------------------
// Planning
// ---------------------------------------------------------------------------
namespace {
Shape drop_lead(const Shape& s) {
  Shape r;
  r.rank = s.rank - 1;
  for (int i = 0; i < r.rank; ++i) r.d[i] = s.d[i + 1];
  return r;
}
Shape with_lead(int64_t n, const Shape& s) {
  TFS_CHECK(s.rank < TFS_MAX_RANK, TFS_ERR_UNSUPPORTED, "tensor rank above the supported maximum");
  Shape r;
  r.rank = s.rank + 1;
  r.d[0] = n;
  for (int i = 0; i < s.rank; ++i) r.d[i + 1] = s.d[i];
  return r;
}
}  // namespace

void Session::build_plan(const std::vector<TensorArg>& feeds, const std::vector<std::string>& fetch_names,
                         Plan& plan, bool rows) {
  const size_t N = g_.nodes.size();
  plan.rows = rows;
  plan.batched.assign(N, 0);
  plan.cshape.assign(N, Shape());
  plan.shape.assign(N, Shape());
  plan.feed_of.assign(N, -1);
  plan.step_of.assign(N, -1);
  plan.materialize.assign(N, 0);
  plan.ival.assign(N, {});
  plan.has_ival.assign(N, 0);

  for (size_t i = 0; i < feeds.size(); ++i) {
    int idx = g_.find(feeds[i].name);
    TFS_CHECK(idx >= 0, TFS_ERR_INVALID, "feed '" + feeds[i].name + "' is not a node of the graph");
    const Node& n = g_.nodes[size_t(idx)];
    TFS_CHECK(n.cls == OpClass::Placeholder || allow_internal_feeds_, TFS_ERR_INVALID,
              "Invalid type for input node " + n.name + ". It has to be a placeholder");
    TFS_CHECK(plan.feed_of[size_t(idx)] < 0, TFS_ERR_INVALID, "placeholder '" + n.name + "' is fed twice");
    // No implicit casting (reference: impl/DebugRowOps.scala:328-333).
    TFS_CHECK(feeds[i].dtype == n.dtype, TFS_ERR_INVALID,
              "The type of node '" + n.name + "' (" + dtype_name(n.dtype) +
                  ") is not compatible with the data type of the feed (" + dtype_name(feeds[i].dtype) + ")");
    TFS_CHECK(feeds[i].shape.fully_known(), TFS_ERR_INVALID,
              "feed '" + n.name + "' has unknown dimensions: " + feeds[i].shape.str());
    if (rows) {
      TFS_CHECK(feeds[i].shape.rank >= 1, TFS_ERR_INVALID, "row mode: feed '" + n.name + "' needs a leading row dimension");
      if (i == 0) plan.n_batch = feeds[i].shape.d[0];
      TFS_CHECK(feeds[i].shape.d[0] == plan.n_batch, TFS_ERR_INVALID, "row mode: feeds differ in their number of rows");
    }
    // in row mode the placeholder describes one cell (reference: impl/DebugRowOps.scala:424-428)
    const Shape fs = rows ? drop_lead(feeds[i].shape) : feeds[i].shape;
    // Shape.checkMorePreciseThan (reference: Shape.scala:54-59)
    if (n.sshape.known_rank()) {
      bool ok = n.sshape.rank == fs.rank;
      for (int d = 0; ok && d < fs.rank; ++d) ok = n.sshape.d[d] < 0 || n.sshape.d[d] == fs.d[d];
      TFS_CHECK(ok, TFS_ERR_INVALID,
                "The data fed to '" + n.name + "' has shape " + fs.str() + " (not compatible) with shape " +
                    n.sshape.str() + " requested by the TF graph");
    }
    plan.feed_of[size_t(idx)] = int(i);
  }

  for (const std::string& f : fetch_names) {
    int idx = g_.find(f);
    TFS_CHECK(idx >= 0, TFS_ERR_INVALID, "fetch '" + f + "' is not a node of the graph");
    plan.fetch_nodes.push_back(idx);
  }
  TFS_CHECK(!plan.fetch_nodes.empty(), TFS_ERR_INVALID, "Output cannot be empty: no fetch requested");

  // needed nodes, topological
  std::vector<char> state(N, 0);
  for (int root : plan.fetch_nodes) {
    if (state[size_t(root)]) continue;
    std::vector<std::pair<int, size_t>> stack{{root, 0}};
    state[size_t(root)] = 1;
    while (!stack.empty()) {
      auto& [u, k] = stack.back();
      const Node& nu = g_.nodes[size_t(u)];
      if (k < nu.in.size() && plan.feed_of[size_t(u)] < 0) {       // (a fed node is a leaf, whatever its op)
        int v = nu.in[k++];
        if (!state[size_t(v)]) { state[size_t(v)] = 1; stack.push_back({v, 0}); }
      } else {
        plan.needed.push_back(u);
        stack.pop_back();
      }
    }
  }

  // concrete shapes (row mode: per-row shapes first, then the row dimension in front of everything that
  // depends on a placeholder)
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    if (n.cls == OpClass::Placeholder || plan.feed_of[size_t(idx)] >= 0) {
      int f = plan.feed_of[size_t(idx)];
      TFS_CHECK(f >= 0, TFS_ERR_INVALID, "placeholder '" + n.name + "' is needed by the fetches but was not fed");
      plan.cshape[size_t(idx)] = rows ? drop_lead(feeds[size_t(f)].shape) : feeds[size_t(f)].shape;
      plan.batched[size_t(idx)] = rows ? 1 : 0;
    } else {
      std::vector<Shape> in;
      IntVals vals;
      for (int k : n.in) {
        in.push_back(plan.cshape[size_t(k)]);
        vals.push_back(plan.has_ival[size_t(k)] ? &plan.ival[size_t(k)] : nullptr);
        // a Shape node reads the operand's shape, not its rows
        if (plan.batched[size_t(k)] && !is_host_value(n)) plan.batched[size_t(idx)] = 1;
      }
      plan.has_ival[size_t(idx)] = infer_int_value(g_, n, in, vals, plan.ival[size_t(idx)]) ? 1 : 0;
      if (is_host_value(n))
        TFS_CHECK(plan.has_ival[size_t(idx)], TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (" + n.op + "): supported only on integer shape vectors (Shape, Const, Pack, StridedSlice)");
      if (n.cls == OpClass::Tile)
        TFS_CHECK(vals[1] != nullptr, TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (Tile): multiples must follow from shapes and constants");
      if (n.cls == OpClass::SegmentSum)
        TFS_CHECK(vals[2] != nullptr && vals[2]->size() == 1 && (*vals[2])[0] >= 0, TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (UnsortedSegmentSum): num_segments must be a non-negative scalar that follows from shapes and constants");
      Shape sh = infer_shape(g_, n, in, vals);
      TFS_CHECK(sh.fully_known(), TFS_ERR_INVALID, "could not resolve the shape of node '" + n.name + "': " + sh.str());
      plan.cshape[size_t(idx)] = sh;
      if (rows && plan.batched[size_t(idx)])
        TFS_CHECK(n.cls != OpClass::MatMul && n.cls != OpClass::Tile && n.cls != OpClass::ArgMinMax &&
                      n.cls != OpClass::SegmentSum && n.cls != OpClass::ExpandDims,
                  TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (" + n.op + ") is not supported in row mode; use map_blocks");
    }
    plan.shape[size_t(idx)] = plan.batched[size_t(idx)] ? with_lead(plan.n_batch, plan.cshape[size_t(idx)])
                                                        : plan.cshape[size_t(idx)];
    TFS_CHECK(plan.shape[size_t(idx)].rank <= TFS_MAX_RANK, TFS_ERR_UNSUPPORTED, "tensor rank above the supported maximum");
    // sizes are int64 throughout (the reference stops at Int.MaxValue elements, impl/datatypes.scala:74-78), but a
    // product of dimensions must still be a number: 2^62 bytes is far beyond any device
    int64_t elems = 1;
    bool fits = true;
    for (int d = 0; d < plan.shape[size_t(idx)].rank && fits; ++d) {
      if (plan.shape[size_t(idx)].d[d] == 0) { elems = 0; break; }
      fits = !__builtin_mul_overflow(elems, plan.shape[size_t(idx)].d[d], &elems) && elems < (int64_t(1) << 59);
    }
    TFS_CHECK(fits, TFS_ERR_INVALID, "node '" + n.name + "': " + plan.shape[size_t(idx)].str() + " has too many elements");
  }

  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    const bool int_div = n.cls == OpClass::Binary && n.elt_op == E_DIV && (n.dtype == TFS_INT32 || n.dtype == TFS_INT64);
    if (int_div || n.cls == OpClass::SegmentSum) plan.may_raise = true;
  }

  // levels: number of reduce/matmul nodes on the longest path below a node
  std::vector<int> level(N, 0);
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    int l = 0;
    for (int k : n.in) l = std::max(l, level[size_t(k)] + (is_compute_barrier(g_.nodes[size_t(k)]) ? 1 : 0));
    level[size_t(idx)] = l;
  }

  // MatMul followed only by "+ bias[N]" becomes one step: the bias is added in the MatMul epilogue
  std::vector<int> consumers(N, 0), fused_from(N, -1), fused_bias(N, -1);
  std::vector<char> fetched(N, 0);
  for (int idx : plan.needed)
    for (int k : g_.nodes[size_t(idx)].in) consumers[size_t(k)]++;
  for (int f : plan.fetch_nodes) fetched[size_t(f)] = 1;
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    if (n.cls != OpClass::Binary || n.elt_op != E_ADD) continue;
    for (int side = 0; side < 2; ++side) {
      const int mm = n.in[size_t(side)], b = n.in[size_t(1 - side)];
      const Node& nm = g_.nodes[size_t(mm)];
      if (nm.cls != OpClass::MatMul || consumers[size_t(mm)] != 1 || fetched[size_t(mm)] || mm == b) continue;
      const Shape& bs = plan.shape[size_t(b)];
      const int64_t ncols = plan.shape[size_t(mm)].d[1];
      int lead = 0;
      while (lead < bs.rank && bs.d[lead] == 1) ++lead;
      if (bs.rank - lead == 1 && bs.d[lead] == ncols && g_.nodes[size_t(b)].cls != OpClass::MatMul) {
        fused_from[size_t(idx)] = mm;
        fused_bias[size_t(mm)] = b;
        break;
      }
    }
  }

  // steps: element-wise nodes of equal (level, shape) fuse into one group
  std::map<std::string, int> group_of_key;
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    if (plan.feed_of[size_t(idx)] >= 0) continue;        // fed: nothing to compute (a placeholder, or an internal node the caller supplies)
    if (fused_from[size_t(idx)] >= 0) {
      const int st = plan.step_of[size_t(fused_from[size_t(idx)])];
      plan.steps[size_t(st)].out_node = idx;
      plan.steps[size_t(st)].bias_node = fused_bias[size_t(fused_from[size_t(idx)])];
      plan.step_of[size_t(idx)] = st;
    } else if (is_elt(n)) {
      std::string key = std::to_string(level[size_t(idx)]) + "|" + plan.shape[size_t(idx)].str();
      auto it = group_of_key.find(key);
      if (it == group_of_key.end()) {
        Step s;
        s.kind = Step::ELT;
        s.shape = plan.shape[size_t(idx)];
        plan.steps.push_back(s);
        it = group_of_key.emplace(key, int(plan.steps.size()) - 1).first;
      }
      plan.steps[size_t(it->second)].nodes.push_back(idx);
      plan.step_of[size_t(idx)] = it->second;
    } else if (n.cls == OpClass::Reduce) {
      Step s;
      s.nodes = {idx};
      const Shape& xs = plan.shape[size_t(n.in[0])];
      const bool in_batched = plan.batched[size_t(n.in[0])] != 0;
      std::vector<int> axes = normalize_axes(n, plan.cshape[size_t(n.in[0])].rank);
      if (in_batched)
        for (int& a : axes) a += 1;     // the row dimension sits in front of the cell
      const int k = int(axes.size());
      bool leading = true, trailing = true;
      for (int i = 0; i < k; ++i) {
        leading = leading && axes[size_t(i)] == i;
        trailing = trailing && axes[size_t(i)] == xs.rank - k + i;
      }
      if (k == 0) {
        s.kind = Step::COPY;
      } else if (leading) {
        s.kind = Step::REDUCE_AXIS0;
        s.outer = s.inner = 1;
        for (int i = 0; i < xs.rank; ++i) (i < k ? s.outer : s.inner) *= xs.d[i];
      } else if (trailing) {
        s.kind = Step::REDUCE_ROWS;
        s.outer = s.inner = 1;
        for (int i = 0; i < xs.rank; ++i) (i < xs.rank - k ? s.outer : s.inner) *= xs.d[i];
      } else {
        bool contiguous = true;
        for (int i = 1; i < k; ++i) contiguous = contiguous && axes[size_t(i)] == axes[size_t(i - 1)] + 1;
        TFS_CHECK(contiguous, TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (" + n.op + "): the reduced axes must be adjacent, input shape " + xs.str());
        s.kind = Step::REDUCE_MID;            // [outer, len, inner] along len
        s.outer = s.len = s.inner = 1;
        for (int i = 0; i < xs.rank; ++i) (i < axes[0] ? s.outer : (i <= axes[size_t(k - 1)] ? s.len : s.inner)) *= xs.d[i];
      }
      plan.steps.push_back(s);
      plan.step_of[size_t(idx)] = int(plan.steps.size()) - 1;
    } else if (n.cls == OpClass::MatMul) {
      Step s;
      s.kind = Step::MATMUL;
      s.nodes = {idx};
      plan.steps.push_back(s);
      plan.step_of[size_t(idx)] = int(plan.steps.size()) - 1;
    } else if (is_host_value(n) || n.cls == OpClass::ExpandDims || n.cls == OpClass::Tile || n.cls == OpClass::ArgMinMax ||
               n.cls == OpClass::SegmentSum) {
      Step s;
      s.nodes = {idx};
      if (is_host_value(n)) {
        s.kind = Step::HOSTVAL;
      } else if (n.cls == OpClass::ExpandDims) {
        s.kind = Step::COPY;                  // same bytes, one more unit dimension
      } else if (n.cls == OpClass::Tile) {
        s.kind = Step::TILE;
      } else if (n.cls == OpClass::ArgMinMax) {
        s.kind = Step::ARGMINMAX;
        const Shape& xs = plan.shape[size_t(n.in[0])];
        int ax = n.axis < 0 ? n.axis + xs.rank : n.axis;
        s.outer = s.inner = 1;
        for (int i = 0; i < xs.rank; ++i) (i < ax ? s.outer : s.inner) *= (i == ax ? 1 : xs.d[i]);
        s.len = xs.d[ax];
        TFS_CHECK(s.len >= 1 || xs.numel() == 0, TFS_ERR_INVALID, "node '" + n.name + "' (" + n.op + "): the searched axis is empty");
      } else {
        s.kind = Step::SEGSUM;
        s.len = plan.ival[size_t(n.in[2])][0];
      }
      plan.steps.push_back(s);
      plan.step_of[size_t(idx)] = int(plan.steps.size()) - 1;
    }
  }

  // order steps (Kahn) and mark which node outputs cross a step boundary
  const size_t S = plan.steps.size();
  std::vector<std::set<int>> deps(S);
  for (int idx : plan.needed) {
    int st = plan.step_of[size_t(idx)];
    if (st < 0) continue;
    const Node& n = g_.nodes[size_t(idx)];
    std::vector<int> operands = is_elt(n) ? elt_operands(n) : std::vector<int>{n.in[0]};
    if (n.cls == OpClass::MatMul) operands = n.in;
    if (n.cls == OpClass::SegmentSum) operands = {n.in[0], n.in[1]};
    if (is_host_value(n)) operands.clear();
    if (fused_from[size_t(idx)] >= 0) operands = {fused_from[size_t(idx)], fused_bias[size_t(fused_from[size_t(idx)])]};
    for (int k : operands) {
      int sk = plan.step_of[size_t(k)];
      if (sk != st) plan.materialize[size_t(k)] = 1;
      if (sk >= 0 && sk != st) deps[size_t(st)].insert(sk);
    }
  }
  for (int f : plan.fetch_nodes) plan.materialize[size_t(f)] = 1;
  std::vector<int> order;
  std::vector<char> placed(S, 0);
  while (order.size() < S) {
    bool progress = false;
    for (size_t s = 0; s < S; ++s) {
      if (placed[s]) continue;
      bool ready = true;
      for (int d : deps[s]) ready = ready && placed[size_t(d)];
      if (ready) { placed[s] = 1; order.push_back(int(s)); progress = true; }
    }
    TFS_CHECK(progress, TFS_ERR_UNSUPPORTED, "internal: cyclic dependency between fused groups");
  }
  std::vector<Step> sorted;
  std::vector<int> remap(S, 0);
  for (size_t i = 0; i < S; ++i) { remap[size_t(order[i])] = int(i); sorted.push_back(plan.steps[size_t(order[i])]); }
  plan.steps.swap(sorted);
  for (int idx : plan.needed)
    if (plan.step_of[size_t(idx)] >= 0) plan.step_of[size_t(idx)] = remap[size_t(plan.step_of[size_t(idx)])];
}

void Session::resolve(const std::vector<TensorArg>& feeds, std::vector<TensorArg>& fetches, bool rows) {
  Plan plan;
  std::vector<std::string> names;
  for (auto& f : fetches) names.push_back(f.name);
  build_plan(feeds, names, plan, rows);
  for (size_t i = 0; i < fetches.size(); ++i) {
    int idx = plan.fetch_nodes[i];
    fetches[i].dtype = g_.nodes[size_t(idx)].dtype;
    // row mode: one cell per row for every fetch, also for fetches that do not depend on the row
    fetches[i].shape = (rows && !plan.batched[size_t(idx)]) ? with_lead(plan.n_batch, plan.cshape[size_t(idx)])
                                                            : plan.shape[size_t(idx)];
  }
}

// ---------------------------------------------------------------------------
// Element-wise group -> accumulator-machine program
// ---------------------------------------------------------------------------
namespace {

struct GroupCompiler {
  const Graph& g;
  const std::vector<Shape>& shape;
  const std::vector<void*>& dev_ptr;
  const std::vector<char>& materialize;
  const Step& step;
  // row mode (mapRows): which nodes carry the row dimension in front, and how many rows there are
  const std::vector<char>* batched = nullptr;
  int64_t n_batch = 0;
  EltProgram prog{};
  bool wide = false;
  std::set<int> members;
  std::map<int, int> uses;        // remaining internal uses per member
  std::map<int, int> tmp_of;      // member -> persistent tmp slot
  std::set<int> done;
  bool tmp_busy[kMaxTmp] = {false, false, false, false};

  // where an operand comes from
  struct Source { enum Kind { BUF, PERIODIC, ROWBCAST, IMM, TMP } kind; int arg = 0; uint64_t imm = 0; };

  GroupCompiler(const Graph& g_, const std::vector<Shape>& sh, const std::vector<void*>& dp,
                const std::vector<char>& mat, const Step& st)
      : g(g_), shape(sh), dev_ptr(dp), materialize(mat), step(st) {
    for (int n : step.nodes) members.insert(n);
    for (int n : step.nodes) {
      std::vector<int> ops = elt_operands(g.nodes[size_t(n)]);
      if (ops.size() == 2 && ops[0] == ops[1]) ops.pop_back();
      for (int o : ops)
        if (members.count(o)) uses[o]++;
      if (dtype_size(g.nodes[size_t(n)].dtype) == 8) wide = true;
      for (int o : ops)
        if (dtype_size(g.nodes[size_t(o)].dtype) == 8) wide = true;
    }
  }

  const Node& node(int i) const { return g.nodes[size_t(i)]; }
  int di(int n) const { return dtype_index(node(n).dtype); }

  Instr& emit(int code, int arg = 0, uint64_t imm = 0) {
    TFS_CHECK(prog.n_instr < kMaxInstr, TFS_ERR_UNSUPPORTED,
              "element-wise group too large for one kernel (" + std::to_string(kMaxInstr) + " instructions)");
    Instr& in = prog.ins[prog.n_instr++];
    in = Instr{};
    in.code = uint8_t(code);
    in.arg = uint8_t(arg);
    in.imm = imm;
    return in;
  }
  int alloc_tmp() {
    for (int t = 0; t < kMaxTmp; ++t)
      if (!tmp_busy[t]) { tmp_busy[t] = true; return t; }
    fail(TFS_ERR_UNSUPPORTED, "element-wise expression needs more than " + std::to_string(kMaxTmp) + " live temporaries");
  }
  int buffer(int n, int64_t period, int64_t rowspan = 0) {
    void* p = dev_ptr[size_t(n)];
    TFS_CHECK(p != nullptr, TFS_ERR_INVALID, "internal: no device buffer for node '" + node(n).name + "'");
    for (int i = 0; i < prog.n_bufs; ++i)
      if (prog.bufs[i].ptr == p && prog.bufs[i].period == period && prog.bufs[i].rowspan == rowspan) return i;
    TFS_CHECK(prog.n_bufs < kMaxBufs, TFS_ERR_UNSUPPORTED,
              "element-wise group touches more than " + std::to_string(kMaxBufs) + " tensors");
    BufRef& b = prog.bufs[prog.n_bufs];
    b.ptr = p;
    b.period = period;
    b.rowspan = rowspan;
    b.dt = di(n);
    // periodic operands: 1 = a vector of the kernel's width never wraps and is aligned (see load_periodic)
    const int64_t lanes = wide ? 2 : 4;
    b.pad = (rowspan == 0 && period > 0 && period % lanes == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) ? 1 : 0;
    return prog.n_bufs++;
  }

  // Operand `o`: an external tensor, or a member already held in a temporary (whose use is consumed).
  Source source_of(int o) {
    Source src;
    if (members.count(o)) {
      auto it = tmp_of.find(o);
      TFS_CHECK(it != tmp_of.end(), TFS_ERR_INVALID, "internal: member operand without a temporary");
      src.kind = Source::TMP;
      src.arg = it->second;
      if (--uses[o] == 0) { tmp_busy[it->second] = false; tmp_of.erase(it); }
      return src;
    }
    const Shape& os = shape[size_t(o)];
    const Shape& it = step.shape;
    const int64_t on = os.numel(), n = it.numel();
    const Node& no = node(o);
    if (on == 1 && no.cls == OpClass::Const) {
      src.kind = Source::IMM;
      memcpy(&src.imm, no.value.bytes.data(), size_t(dtype_size(no.dtype)));
      return src;
    }
    if (on == n) {
      src.kind = Source::BUF;
      src.arg = buffer(o, 0);
      return src;
    }
    // Row mode: an operand that has its own value per row (a reduction of the cell, a smaller fed cell) broadcasts
    // INSIDE each row.  Its shape already carries the row dimension, so matching it against the trailing dimensions
    // of the block below would be wrong -- and would silently pass whenever the row count equals a cell dimension.
    if (batched != nullptr && (*batched)[size_t(o)]) {
      TFS_CHECK(n_batch > 0 && os.rank >= 1 && it.rank >= 1 && os.d[0] == n_batch && it.d[0] == n_batch, TFS_ERR_UNSUPPORTED,
                "row mode: operand '" + no.name + "' " + os.str() + " does not line up with the rows of " + it.str());
      int lead = 1;
      while (lead < os.rank && os.d[lead] == 1) ++lead;
      const int k = os.rank - lead;
      bool ok = k <= it.rank - 1;
      for (int i = 0; ok && i < k; ++i) ok = os.d[lead + i] == it.d[it.rank - k + i];
      TFS_CHECK(ok, TFS_ERR_UNSUPPORTED,
                "broadcasting the per-row value " + os.str() + " against " + it.str() + " (operand '" + no.name +
                    "') is not supported: only scalars and trailing-dimension broadcasts within a row are");
      src.kind = Source::ROWBCAST;
      src.arg = buffer(o, on / n_batch, n / n_batch);
      return src;
    }
    // row broadcast: operand dims (leading 1s dropped) must equal the trailing dims of the block
    int lead = 0;
    while (lead < os.rank && os.d[lead] == 1) ++lead;
    const int k = os.rank - lead;
    bool ok = k <= it.rank;
    for (int i = 0; ok && i < k; ++i) ok = os.d[lead + i] == it.d[it.rank - k + i];
    TFS_CHECK(ok, TFS_ERR_UNSUPPORTED,
              "broadcasting " + os.str() + " against " + it.str() + " (operand '" + no.name +
                  "') is not supported: only scalars and trailing-dimension broadcasts are");
    src.kind = Source::PERIODIC;
    src.arg = buffer(o, on);
    return src;
  }

  bool needs_compute(int o) const { return members.count(o) && !tmp_of.count(o); }

  void into_acc(int o) {
    if (needs_compute(o)) {
      compute(o, true);
      return;
    }
    Source s = source_of(o);
    switch (s.kind) {
      case Source::BUF: emit(C_LOAD_BUF, s.arg); break;
      case Source::PERIODIC: emit(C_LOAD_PER, s.arg); break;
      case Source::ROWBCAST: emit(C_LOAD_ROW, s.arg); break;
      case Source::IMM: emit(C_LOAD_IMM, 0, s.imm); break;
      case Source::TMP: emit(C_LOAD_TMP, s.arg); break;
    }
  }

  static int bin_sub(int elt_op) {
    switch (elt_op) {
      case E_ADD: return B_ADD;
      case E_SUB: return B_SUB;
      case E_MUL: return B_MUL;
      case E_DIV: return B_DIV;
      case E_MIN: return B_MIN;
      default: return B_MAX;
    }
  }
  static int un_sub(int elt_op) {
    switch (elt_op) {
      case E_NEG: return U_NEG;
      case E_RELU: return U_RELU;
      case E_SQUARE: return U_SQUARE;
      case E_ABS: return U_ABS;
      case E_SQRT: return U_SQRT;
      default: return U_RECIP;
    }
  }

  // acc <- op(acc, operand)  (swap: op(operand, acc)), operand being a leaf
  void binary_with(int n, int operand, bool swap) {
    const int sub = bin_sub(node(n).elt_op), dt = di(n), sw = swap ? 1 : 0;
    Source s = source_of(operand);
    if (s.kind == Source::IMM) {
      emit(C_BIN_IMM + sub * 8 + dt * 2 + sw, 0, s.imm);
      return;
    }
    switch (s.kind) {
      case Source::BUF: emit(C_FETCH_BUF, s.arg); break;
      case Source::PERIODIC: emit(C_FETCH_PER, s.arg); break;
      case Source::ROWBCAST: emit(C_FETCH_ROW, s.arg); break;
      default: emit(C_FETCH_TMP, s.arg); break;
    }
    emit(C_BIN_X + sub * 8 + dt * 2 + sw);
  }

  // Leaves the value of member n in the accumulator.
  void compute(int n, bool consumer_takes_acc) {
    TFS_CHECK(!done.count(n), TFS_ERR_INVALID, "internal: element-wise node computed twice");
    const Node& nd = node(n);
    std::vector<int> ops = elt_operands(nd);
    if (is_like_fill(nd)) {
      uint64_t imm = 0;
      if (nd.elt_op == E_ONES_LIKE) {
        switch (nd.dtype) {
          case TFS_FLOAT: { const float one = 1.f; memcpy(&imm, &one, 4); break; }
          case TFS_DOUBLE: { const double one = 1.0; memcpy(&imm, &one, 8); break; }
          default: imm = 1;
        }
      }
      emit(C_LOAD_IMM, 0, imm);
    } else if (nd.cls == OpClass::Fill || nd.cls == OpClass::Unary) {
      into_acc(ops[0]);
      if (nd.cls == OpClass::Unary && nd.elt_op == E_CAST) {
        if (di(ops[0]) != di(n)) emit(C_CAST + di(ops[0]) * 4 + di(n));
      } else if (nd.cls == OpClass::Unary && nd.elt_op != E_IDENTITY) {
        emit(C_UNARY + un_sub(nd.elt_op) * 4 + di(n));
      }
    } else {
      const int a = ops[0], b = ops[1];
      const int sub = bin_sub(nd.elt_op), dt = di(n);
      if (a == b) {
        into_acc(a);
        emit(C_FETCH_ACC);
        emit(C_BIN_X + sub * 8 + dt * 2);
      } else if (needs_compute(a) && needs_compute(b)) {
        compute(a, true);
        if (!needs_compute(b)) {
          // evaluating a already produced b (a shared sub-expression, now parked in a temporary): b is a leaf here
          binary_with(n, b, false);
        } else {
          int t = alloc_tmp();
          emit(C_STTMP, t);
          compute(b, true);
          emit(C_FETCH_TMP, t);                   // acc holds b, x holds a: result = op(x, acc)
          emit(C_BIN_X + sub * 8 + dt * 2 + 1);
          tmp_busy[t] = false;
        }
      } else if (needs_compute(b)) {
        compute(b, true);
        binary_with(n, a, true);
      } else {
        into_acc(a);
        binary_with(n, b, false);
      }
    }
    done.insert(n);
    if (materialize[size_t(n)]) emit(C_STORE, buffer(n, 0));
    int remaining = uses.count(n) ? uses[n] : 0;
    if (consumer_takes_acc) remaining -= 1;
    uses[n] = remaining;
    if (remaining > 0) {
      int t = alloc_tmp();
      emit(C_STTMP, t);
      tmp_of[n] = t;
    }
  }

  void build(int32_t* err) {
    prog.err = err;
    for (int n : step.nodes)
      if (materialize[size_t(n)] && !done.count(n)) compute(n, false);
  }
};

// Recognises a compiled group the straight-line kernel can run (see kernels_chain.cu) and rewrites it for that
// kernel: one input column, then slots whose operand is a constant, another column, a row vector, the single saved
// tile or the accumulator itself; reloads, one saved tile, extra output columns of the same dtype, an optional
// final cast.  Returns false when the program needs the general interpreter.
bool make_chain(const EltProgram& prog, ChainProgram& ch) {
  if (prog.n_instr < 2) return false;
  const Instr& first = prog.ins[0];
  const Instr& last = prog.ins[prog.n_instr - 1];
  if (first.code != C_LOAD_BUF || last.code != C_STORE) return false;
  const BufRef& in = prog.bufs[first.arg];
  const BufRef& out = prog.bufs[last.arg];
  auto aligned = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!aligned(in.ptr) || !aligned(out.ptr)) return false;
  ch = ChainProgram{};
  ch.in = in.ptr;
  ch.out = out.ptr;
  ch.err = prog.err;
  ch.in_dt = uint8_t(in.dt);
  int dt = in.dt;
  int first_pc = 1;
  if (prog.n_instr > 2) {           // a leading Cast: the chain computes in the cast's target dtype
    const int c1 = prog.ins[1].code;
    // (to float or double only: int targets and saved-tile programs stay with the interpreter, which keeps the build small)
    if (c1 >= C_CAST && c1 < C_CAST + 16 && (c1 - C_CAST) / 4 == in.dt && (c1 - C_CAST) % 4 != in.dt &&
        ((c1 - C_CAST) % 4 == DI_F32 || (c1 - C_CAST) % 4 == DI_F64)) {
      dt = (c1 - C_CAST) % 4;
      first_pc = 2;
    }
  }
  ch.dt = uint8_t(dt);
  ch.out_dt = ch.dt;
  // elements per register tile: 16 bytes of the narrower of the input and the chain dtype (ChainGeom in kernels_chain.inc);
  // the tile span is the same either way
  const int64_t lanes = 16 / (((dt & 1) && (in.dt & 1)) ? 8 : 4);
  const int64_t span = int64_t(kChainThreads) * kChainVecsPerThread * (16 / ((dt & 1) ? 8 : 4));
  static const uint8_t kFwd[B_COUNT] = {CH_ADD, CH_SUB, CH_MUL, CH_DIV, CH_MIN, CH_MAX};
  static const uint8_t kRev[B_COUNT] = {CH_ADD, CH_RSUB, CH_MUL, CH_RDIV, CH_RMIN, CH_RMAX};
  static const uint8_t kUn[U_COUNT] = {CH_NEG, CH_RELU, CH_SQUARE, CH_ABS, CH_SQRT, CH_RECIP};

  struct Operand { uint8_t kind = CK_IMM; const void* buf = nullptr; uint32_t period = 0; uint64_t imm = 0; };
  // a column / row-vector operand the kernel can read as whole vectors
  auto column_operand = [&](const Instr& ins, bool periodic, Operand& o) {
    const BufRef& b = prog.bufs[ins.arg];
    if (b.dt != dt || !aligned(b.ptr)) return false;
    o = Operand{};
    o.buf = b.ptr;
    if (periodic) {
      if (b.period <= 0 || b.period % lanes != 0 || span % b.period != 0) return false;
      o.kind = CK_ROWVEC;
      o.period = uint32_t(b.period);
    } else {
      o.kind = CK_COL;
    }
    return true;
  };
  auto push = [&](const ChainSlot& slot) {
    if (ch.n_slots >= kChainSlots) return false;
    ch.slots[ch.n_slots++] = slot;
    return true;
  };
  auto slot_with = [](uint8_t op, const Operand& o) {
    ChainSlot slot{};
    slot.op = op;
    slot.kind = o.kind;
    slot.operand_buf = o.buf;
    slot.period = o.period;
    slot.imm = o.imm;
    return slot;
  };
  // the slot a save / store attaches to: the last one, or a no-op placed right after the initial load
  auto last_slot = [&]() -> ChainSlot* {
    if (ch.n_slots == 0) {
      ChainSlot nop{};
      nop.op = CH_NOP;
      nop.kind = CK_IMM;
      if (!push(nop)) return nullptr;
    }
    return &ch.slots[ch.n_slots - 1];
  };

  Operand pending;
  bool have_pending = false, cast_seen = false;
  for (int pc = first_pc; pc < prog.n_instr - 1; ++pc) {
    const Instr& ins = prog.ins[pc];
    const int c = ins.code;
    if (cast_seen) return false;                       // a cast may only start or end the chain
    Operand o;
    switch (c) {
      case C_FETCH_BUF:
      case C_FETCH_PER:
        if (have_pending || !column_operand(ins, c == C_FETCH_PER, pending)) return false;
        have_pending = true;
        continue;
      case C_FETCH_TMP:
        if (have_pending || ins.arg != 0) return false;    // one saved tile
        pending = Operand{};
        pending.kind = CK_TMP;
        have_pending = true;
        continue;
      case C_FETCH_ACC:
        if (have_pending) return false;
        pending = Operand{};
        pending.kind = CK_SELF;
        have_pending = true;
        continue;
      case C_LOAD_BUF:
      case C_LOAD_PER:
        if (have_pending || !column_operand(ins, c == C_LOAD_PER, o) || !push(slot_with(CH_LOAD, o))) return false;
        continue;
      case C_LOAD_TMP:
        if (have_pending || ins.arg != 0) return false;
        o.kind = CK_TMP;
        if (!push(slot_with(CH_LOAD, o))) return false;
        continue;
      case C_LOAD_IMM:
        if (have_pending) return false;
        o.kind = CK_IMM;
        o.imm = ins.imm;
        if (!push(slot_with(CH_LOAD, o))) return false;
        continue;
      case C_STTMP: {
        if (have_pending || ins.arg != 0) return false;
        ChainSlot* sl = last_slot();
        if (!sl || sl->store_buf != nullptr) return false;   // (a slot saves before it stores; keep the program order)
        sl->save = 1;
        continue;
      }
      case C_STORE: {                                        // a further output column, in the chain's dtype
        const BufRef& b = prog.bufs[ins.arg];
        if (have_pending || b.dt != dt || !aligned(b.ptr)) return false;
        ChainSlot* sl = last_slot();
        if (!sl) return false;
        if (sl->store_buf != nullptr) {                      // two stores in a row: a no-op slot carries the second
          ChainSlot nop{};
          nop.op = CH_NOP;
          nop.kind = CK_IMM;
          if (!push(nop)) return false;
          sl = &ch.slots[ch.n_slots - 1];
        }
        sl->store_buf = b.ptr;
        continue;
      }
      default: break;
    }
    if (c >= C_BIN_X && c < C_BIN_X + B_COUNT * 8) {
      const int k = c - C_BIN_X;
      if (!have_pending || (k % 8) / 2 != dt) return false;
      if (!push(slot_with((k & 1) ? kRev[k / 8] : kFwd[k / 8], pending))) return false;
      have_pending = false;
    } else if (c >= C_BIN_IMM && c < C_BIN_IMM + B_COUNT * 8) {
      const int k = c - C_BIN_IMM;
      if (have_pending || (k % 8) / 2 != dt) return false;
      o.kind = CK_IMM;
      o.imm = ins.imm;
      if (!push(slot_with((k & 1) ? kRev[k / 8] : kFwd[k / 8], o))) return false;
    } else if (c >= C_UNARY && c < C_UNARY + U_COUNT * 4) {
      const int k = c - C_UNARY;
      if (have_pending || k % 4 != dt) return false;
      if (!push(slot_with(kUn[k / 4], Operand{}))) return false;
    } else if (c >= C_CAST && c < C_CAST + 16) {
      const int k = c - C_CAST;
      if (have_pending || k / 4 != dt) return false;
      ch.out_dt = uint8_t(k % 4);
      cast_seen = true;
    } else {
      return false;
    }
  }
  if (have_pending) return false;
  if (ch.in_dt != ch.dt) {
    if (ch.out_dt != ch.dt) return false;     // a cast at both ends: the interpreter takes it
    for (int k = 0; k < ch.n_slots; ++k)
      if (ch.slots[k].save || ch.slots[k].kind == CK_TMP) return false;
  }
  return out.dt == ch.out_dt;
}

}  // namespace

// The constant operand of a tensor-core MatMul, split once per session into tf32 hi / lo parts laid out [N, K].
std::unordered_map<int, std::pair<float*, float*>>::iterator Session::ensure_tc_weights(const Node& n, const float* b, int64_t Nc,
                                                                                         int64_t K, cudaStream_t stream) {
  const int key = n.in[1] * 2 + (n.transpose_b ? 1 : 0);
  auto it = tc_weights_.find(key);
  if (it == tc_weights_.end()) {
    float *hi = nullptr, *lo = nullptr;
    TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&hi), matmul_tc_weight_bytes(Nc, K)));
    TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&lo), matmul_tc_weight_bytes(Nc, K)));
    launches_ += launch_split_weights(b, n.transpose_b, K, Nc, hi, lo, stream);
    it = tc_weights_.emplace(key, std::make_pair(hi, lo)).first;
  }
  return it;
}

// One step of a plan on `stream`; dev_ptr holds the device address of every materialised node.
void Session::exec_step(const Plan& plan, size_t step_index, std::vector<void*>& dev_ptr, cudaStream_t stream) {
  const Step& s = plan.steps[step_index];
  // A node that is needed only for its SHAPE (the operand of OnesLike / ZerosLike / Shape) has no buffer: its step
  // has nothing to write and is skipped.
  if (s.kind != Step::ELT && s.kind != Step::HOSTVAL) {
    const int out_node = (s.kind == Step::MATMUL && s.out_node >= 0) ? s.out_node : s.nodes[0];
    if (!plan.materialize[size_t(out_node)]) return;
  }
  switch (s.kind) {
    case Step::ELT: {
      if (s.shape.numel() == 0) break;
      GroupCompiler gc(g_, plan.shape, dev_ptr, plan.materialize, s);
      if (plan.rows) { gc.batched = &plan.batched; gc.n_batch = plan.n_batch; }
      gc.build(err_flag_);
      if (gc.prog.n_instr == 0) break;          // every member is needed for its shape only
      ChainProgram chain;
      // TFS_ELT_IMPL=interp sends every group to the interpreter (A/B measurements of the two kernels)
      static const bool force_interp = [] { const char* e = std::getenv("TFS_ELT_IMPL"); return e && std::string(e) == "interp"; }();
      const bool straight = !force_interp && make_chain(gc.prog, chain);
      static const bool trace = [] { const char* e = std::getenv("TFS_TRACE"); return e && *e && *e != '0'; }();
      if (trace)
        fprintf(stderr, "tfs: element-wise group of %zu nodes, %d instructions -> %s\n", s.nodes.size(), gc.prog.n_instr,
                straight ? "straight-line kernel" : "interpreter");
      if (straight) launches_ += launch_chain(chain, s.shape.numel(), dev_, stream);
      else launches_ += launch_eltwise(gc.prog, s.shape.numel(), gc.wide, dev_, stream);
      break;
    }
    case Step::REDUCE_AXIS0: {
      const Node& n = g_.nodes[size_t(s.nodes[0])];
      if (s.inner > 0)
        launches_ += launch_reduce_axis0(n.dtype, n.red_op, dev_ptr[size_t(n.in[0])], s.outer, s.inner,
                                         dev_ptr[size_t(s.nodes[0])], red_scratch_, dev_, stream);
      break;
    }
    case Step::REDUCE_ROWS: {
      const Node& n = g_.nodes[size_t(s.nodes[0])];
      launches_ += launch_reduce_rows(n.dtype, n.red_op, dev_ptr[size_t(n.in[0])], s.outer, s.inner,
                                      dev_ptr[size_t(s.nodes[0])], dev_, stream);
      break;
    }
    case Step::REDUCE_MID: {
      const Node& n = g_.nodes[size_t(s.nodes[0])];
      launches_ += launch_reduce_mid(n.dtype, n.red_op, dev_ptr[size_t(n.in[0])], s.outer, s.len, s.inner,
                                     dev_ptr[size_t(s.nodes[0])], dev_, stream);
      break;
    }
    case Step::COPY: {
      const Node& n = g_.nodes[size_t(s.nodes[0])];
      const size_t bytes = size_t(plan.shape[size_t(s.nodes[0])].numel()) * size_t(dtype_size(n.dtype));
      if (bytes > 0)
        TFS_CUDA(cudaMemcpyAsync(dev_ptr[size_t(s.nodes[0])], dev_ptr[size_t(n.in[0])], bytes, cudaMemcpyDeviceToDevice, stream));
      break;
    }
    case Step::HOSTVAL: {
      const int idx = s.nodes[0];
      if (!plan.materialize[size_t(idx)] || plan.ival[size_t(idx)].empty()) break;   // read as a value by the planner only
      const Node& n = g_.nodes[size_t(idx)];
      const std::vector<int64_t>& v = plan.ival[size_t(idx)];
      if (n.dtype == TFS_INT64) {
        TFS_CUDA(cudaMemcpyAsync(dev_ptr[size_t(idx)], v.data(), v.size() * 8, cudaMemcpyHostToDevice, stream));
      } else {
        std::vector<int32_t> v32(v.begin(), v.end());
        TFS_CUDA(cudaMemcpyAsync(dev_ptr[size_t(idx)], v32.data(), v32.size() * 4, cudaMemcpyHostToDevice, stream));
      }
      // (pageable source: the copy has been staged when the call returns, so the vectors may go)
      break;
    }
    case Step::TILE: {
      const int idx = s.nodes[0];
      const Node& n = g_.nodes[size_t(idx)];
      launches_ += launch_tile(dtype_size(n.dtype), dev_ptr[size_t(n.in[0])], plan.shape[size_t(n.in[0])], dev_ptr[size_t(idx)],
                               plan.shape[size_t(idx)], dev_, stream);
      break;
    }
    case Step::ARGMINMAX: {
      const int idx = s.nodes[0];
      const Node& n = g_.nodes[size_t(idx)];
      if (s.outer * s.inner > 0)
        launches_ += launch_argminmax(g_.nodes[size_t(n.in[0])].dtype, n.dtype, n.arg_min, dev_ptr[size_t(n.in[0])], s.outer, s.len,
                                      s.inner, dev_ptr[size_t(idx)], dev_, stream);
      break;
    }
    case Step::SEGSUM: {
      // UnsortedSegmentSum = the keyed block reduction (K3) with the segment ids as keys, then a scatter of the
      // per-group sums to out[id, :]; segments without rows stay zero.
      const int idx = s.nodes[0];
      const Node& n = g_.nodes[size_t(idx)];
      const Shape& xs = plan.shape[size_t(n.in[0])];
      const int64_t n_rows = xs.rank ? xs.d[0] : 0, segments = s.len;
      const int64_t cell = n_rows > 0 ? xs.numel() / n_rows : plan.shape[size_t(idx)].numel() / std::max<int64_t>(segments, 1);
      const size_t es = size_t(dtype_size(n.dtype));
      const size_t out_bytes = size_t(segments) * size_t(cell) * es;
      if (out_bytes > 0) TFS_CUDA(cudaMemsetAsync(dev_ptr[size_t(idx)], 0, out_bytes, stream));
      if (n_rows == 0 || cell == 0 || segments == 0) {
        TFS_CHECK(n_rows == 0 || segments > 0, TFS_ERR_RUNTIME, "UnsortedSegmentSum: segment ids out of range (num_segments is 0)");
        break;
      }
      TFS_CHECK(n_rows < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "UnsortedSegmentSum: blocks of 2^32 rows or more are not supported");
      const int64_t max_groups = std::min(n_rows, segments);
      const AggLayout L0 = agg_layout(n_rows, max_groups, dev_);
      AggLayout L = L0;
      const size_t off_keys64 = align_up(L.bytes);
      const size_t off_out_keys = off_keys64 + align_up(size_t(n_rows) * 8);
      const size_t off_sums = off_out_keys + align_up(size_t(max_groups) * 8);
      const size_t off_split = off_sums + align_up(size_t(max_groups) * size_t(cell) * es);
      const size_t split_bytes = agg_split_scratch_bytes(cell, dev_);
      const size_t need = off_split + align_up(split_bytes);
      if (need > seg_scratch_cap_) {
        TFS_CUDA(cudaStreamSynchronize(stream));
        if (seg_scratch_) TFS_CUDA(cudaFree(seg_scratch_));
        seg_scratch_ = nullptr;
        seg_scratch_cap_ = 0;
        TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&seg_scratch_), need + need / 8));
        seg_scratch_cap_ = need + need / 8;
      }
      char* scratch = seg_scratch_;
      const int64_t* keys = static_cast<const int64_t*>(dev_ptr[size_t(n.in[1])]);
      if (g_.nodes[size_t(n.in[1])].dtype == TFS_INT32) {
        launches_ += launch_widen_ids(static_cast<const int*>(dev_ptr[size_t(n.in[1])]), n_rows,
                                      reinterpret_cast<int64_t*>(scratch + off_keys64), dev_, stream);
        keys = reinterpret_cast<const int64_t*>(scratch + off_keys64);
      }
      int64_t* d_out_keys = reinterpret_cast<int64_t*>(scratch + off_out_keys);
      launches_ += launch_agg_build(keys, n_rows, max_groups, scratch, L, dev_, stream);
      unsigned int header[4] = {0, 0, 0, 0};
      TFS_CUDA(cudaMemcpyAsync(header, scratch + L.off_header, sizeof(header), cudaMemcpyDeviceToHost, stream));
      TFS_CUDA(cudaStreamSynchronize(stream));
      TFS_CHECK(header[1] == 0 && int64_t(header[2]) <= max_groups, TFS_ERR_RUNTIME,
                "UnsortedSegmentSum: segment ids out of range [0, " + std::to_string(segments) + ")");
      const int64_t n_groups = header[2];
      launches_ += launch_agg_segments(keys, n_rows, n_groups, scratch, L, d_out_keys, kGroupsByKey, true, dev_, stream);
      launches_ += launch_agg_reduce(n.dtype, R_SUM, dev_ptr[size_t(n.in[0])], cell, n_groups, n_rows, scratch, L, scratch + off_sums,
                                     scratch + off_split, split_bytes, dev_, stream);
      launches_ += launch_segment_scatter(n.dtype, scratch + off_sums, d_out_keys, n_groups, cell, segments, dev_ptr[size_t(idx)],
                                          err_flag_, dev_, stream);
      break;
    }
    case Step::MATMUL: {
      const int mm = s.nodes[0];
      const Node& n = g_.nodes[size_t(mm)];
      const Shape& as = plan.shape[size_t(n.in[0])];
      const Shape& os = plan.shape[size_t(mm)];
      const int64_t M = os.d[0], Nc = os.d[1], K = n.transpose_a ? as.d[0] : as.d[1];
      const void* a = dev_ptr[size_t(n.in[0])];
      const void* b = dev_ptr[size_t(n.in[1])];
      const void* bias = s.bias_node >= 0 ? dev_ptr[size_t(s.bias_node)] : nullptr;
      void* out = dev_ptr[size_t(s.out_node >= 0 ? s.out_node : mm)];
      if (M * Nc == 0) break;
      const bool const_b = g_.nodes[size_t(n.in[1])].cls == OpClass::Const;
      // The tensor-core kernel reads x and writes y through TMA tensor maps (its w loads use the cta_group::2 form):
      // those maps cover memory of this GPU only, so anything else -- managed memory, whose pages may live elsewhere --
      // takes the SIMT kernel.  (Memory of another device never gets here: the run refuses it.)
      const bool local = pointer_location(a) == dev_.device && pointer_location(out) == dev_.device;
      if (n.dtype == TFS_FLOAT && !matmul_force_simt() && const_b && local && matmul_tc_supported(a, out, M, Nc, K, n.transpose_a)) {
        auto it = ensure_tc_weights(n, static_cast<const float*>(b), Nc, K, stream);
        launches_ += launch_matmul_tc(static_cast<const float*>(a), it->second.first, it->second.second, static_cast<const float*>(bias),
                                      static_cast<float*>(out), M, Nc, K, dev_, stream);
      } else {
        launches_ += launch_matmul_simt(n.dtype, a, b, bias, out, M, Nc, K, n.transpose_a, n.transpose_b, dev_, stream);
      }
      break;
    }
  }
}

// ---------------------------------------------------------------------------
// Host ingest pipeline (K0) for row-separable map graphs
// ---------------------------------------------------------------------------
// A map whose whole graph is ONE fused element-wise group over [n_rows, cell...] can be evaluated chunk by
// chunk.  When every feed and fetch is a page-locked host buffer (tfs_host_alloc), the block is cut into
// chunks of rows that go round two streams: while one chunk is being computed, the next one is on its way
// in over PCIe and the previous one on its way out, so a step costs about max(H2D, D2H) instead of
// H2D + kernel + D2H.  This replaces the reference's pack -> Tensor.create -> run -> writeTo sequence
// (impl/TFDataOps.scala:27-59, impl/datatypes.scala:344-347, 365-370), which is fully serial.
bool Session::run_pipelined(const std::vector<TensorArg>& feeds, const std::vector<TensorArg>& fetches, const Plan& plan) {
  constexpr size_t kChunkBytes = size_t(32) << 20;    // per-chunk budget of the widest column
  if (plan.steps.size() != 1) return false;
  const Step& st = plan.steps[0];
  const bool is_mm = st.kind == Step::MATMUL;
  if (st.kind != Step::ELT && !is_mm) return false;
  int64_t n = 0;
  if (is_mm) {
    // x[n, K] (a fed placeholder, not transposed) times a constant matrix (+ constant bias): rows are independent
    const Node& mm = g_.nodes[size_t(st.nodes[0])];
    if (mm.transpose_a || g_.nodes[size_t(mm.in[0])].cls != OpClass::Placeholder ||
        g_.nodes[size_t(mm.in[1])].cls != OpClass::Const)
      return false;
    if (st.bias_node >= 0 && g_.nodes[size_t(st.bias_node)].cls != OpClass::Const) return false;
    n = plan.shape[size_t(st.nodes[0])].d[0];
  } else {
    if (st.shape.rank < 1) return false;
    n = st.shape.d[0];
  }
  for (const TensorArg& f : feeds)
    if (!is_pinned_host_pointer(f.data)) return false;
  for (const TensorArg& f : fetches)
    if (!is_pinned_host_pointer(f.data)) return false;
  // every fed placeholder and every fetch must be a [n, ...] block of the step's row count
  size_t max_row_bytes = 0, total_bytes = 0;
  auto row_bytes = [&](int idx) {
    const Shape& sh = plan.shape[size_t(idx)];
    return size_t(sh.numel() / std::max<int64_t>(n, 1)) * size_t(dtype_size(g_.nodes[size_t(idx)].dtype));
  };
  const int out_node = is_mm ? (st.out_node >= 0 ? st.out_node : st.nodes[0]) : -1;
  for (int idx : plan.needed) {
    const Node& nd = g_.nodes[size_t(idx)];
    if (nd.cls == OpClass::Placeholder) {
      const Shape& sh = plan.shape[size_t(idx)];
      if (sh.rank < 1 || sh.d[0] != n) return false;
      if (!is_mm && sh.numel() != st.shape.numel()) return false;     // only same-shaped columns (no per-row broadcast)
      max_row_bytes = std::max(max_row_bytes, row_bytes(idx));
      total_bytes += size_t(sh.numel()) * size_t(dtype_size(nd.dtype));
    }
  }
  for (int f : plan.fetch_nodes) {
    if (plan.step_of[size_t(f)] != 0) return false;
    if (is_mm ? f != out_node : plan.shape[size_t(f)] != st.shape) return false;
    max_row_bytes = std::max(max_row_bytes, row_bytes(f));
  }
  if (total_bytes < 4 * kChunkBytes || max_row_bytes == 0) return false;   // small blocks: one shot is as good
  // every other operand of the group must be independent of the row count: a constant (or a value computed from
  // constants) as large as the whole block lines up with the block's rows and cannot be cut along with them
  if (!is_mm) {
    std::set<int> members(st.nodes.begin(), st.nodes.end());
    for (int m : st.nodes)
      for (int o : elt_operands(g_.nodes[size_t(m)])) {
        if (members.count(o) || g_.nodes[size_t(o)].cls == OpClass::Placeholder) continue;
        const int64_t on = plan.shape[size_t(o)].numel();
        if (on > 1 && on == st.shape.numel()) return false;
      }
  }

  int64_t chunk_rows = int64_t(kChunkBytes / max_row_bytes);
  chunk_rows = std::max<int64_t>(16, chunk_rows / 16 * 16);    // keeps every chunk start 16-byte aligned
  const int64_t n_chunks = (n + chunk_rows - 1) / chunk_rows;

  // chunk plans: one for a full chunk, one for the last (shorter) chunk.  Planned before anything is allocated or
  // enqueued: a graph that does not plan for a chunk of rows (an operand tied to the row count in a way the test
  // above does not see) goes through the one-shot path instead of failing here.
  std::vector<std::string> names;
  for (auto& f : fetches) names.push_back(f.name);
  auto chunk_feeds = [&](int64_t rows_in_chunk) {
    std::vector<TensorArg> cf = feeds;
    for (auto& f : cf) f.shape.d[0] = rows_in_chunk;
    return cf;
  };
  Plan full_plan, tail_plan;
  const int64_t tail_rows = n - (n_chunks - 1) * chunk_rows;
  const bool has_tail = tail_rows != std::min(chunk_rows, n);
  try {
    build_plan(chunk_feeds(std::min(chunk_rows, n)), names, full_plan, false);
    if (has_tail) build_plan(chunk_feeds(tail_rows), names, tail_plan, false);
  } catch (const Error&) {
    return false;
  }
  if (full_plan.steps.size() != 1 || (has_tail && tail_plan.steps.size() != 1)) return false;

  if (!pipe_stream_[0]) {
    for (int k = 0; k < 2; ++k) TFS_CUDA(cudaStreamCreateWithFlags(&pipe_stream_[k], cudaStreamNonBlocking));
    TFS_CUDA(cudaEventCreateWithFlags(&pipe_event_, cudaEventDisableTiming));
  }

  // workspace: two slots, each holding one chunk of every fed column and every fetched column
  const size_t N = g_.nodes.size();
  std::vector<size_t> off(N, SIZE_MAX);
  size_t slot_bytes = 0;
  for (int idx : plan.needed) {
    const Node& nd = g_.nodes[size_t(idx)];
    const bool is_out = std::find(plan.fetch_nodes.begin(), plan.fetch_nodes.end(), idx) != plan.fetch_nodes.end();
    if (nd.cls == OpClass::Placeholder || is_out || (plan.materialize[size_t(idx)] && nd.cls != OpClass::Const)) {
      off[size_t(idx)] = slot_bytes;
      slot_bytes += align_up(row_bytes(idx) * size_t(chunk_rows));
    }
  }
  reserve_workspace(2 * slot_bytes);
  // constants that are read as tensors
  std::vector<void*> const_ptr(N, nullptr);
  for (int idx : plan.needed)
    if (g_.nodes[size_t(idx)].cls == OpClass::Const && plan.materialize[size_t(idx)]) const_ptr[size_t(idx)] = const_device(idx);
  if (is_mm) {
    // the tf32 hi/lo split of the constant matrix is cached per session: make sure it exists (on the session
    // stream) before two pipeline streams would both try to create it
    // -- under the same conditions as exec_step takes the tensor-core path (workspace chunks are 256-byte aligned)
    const Node& mm = g_.nodes[size_t(st.nodes[0])];
    const Shape& os = plan.shape[size_t(st.nodes[0])];
    const Shape& as = plan.shape[size_t(mm.in[0])];
    if (mm.dtype == TFS_FLOAT && !matmul_force_simt() &&
        matmul_tc_supported(ws_, ws_, std::min(chunk_rows, n), os.d[1], as.d[1], mm.transpose_a))
      ensure_tc_weights(mm, static_cast<const float*>(const_ptr[size_t(mm.in[1])]), os.d[1], as.d[1], stream_);
  }
  // everything enqueued so far on the session stream (constant uploads) precedes the pipeline
  TFS_CUDA(cudaEventRecord(pipe_event_, stream_));
  for (int k = 0; k < 2; ++k) TFS_CUDA(cudaStreamWaitEvent(pipe_stream_[k], pipe_event_, 0));

  for (int64_t c = 0; c < n_chunks; ++c) {
    const int slot = int(c & 1);
    cudaStream_t sx = pipe_stream_[slot];
    const int64_t r0 = c * chunk_rows;
    const int64_t rows_here = std::min(chunk_rows, n - r0);
    const Plan& cp = (has_tail && c == n_chunks - 1) ? tail_plan : full_plan;
    std::vector<void*> dev_ptr(N, nullptr);
    for (int idx : plan.needed) {
      if (off[size_t(idx)] != SIZE_MAX) dev_ptr[size_t(idx)] = ws_ + size_t(slot) * slot_bytes + off[size_t(idx)];
      else if (const_ptr[size_t(idx)]) dev_ptr[size_t(idx)] = const_ptr[size_t(idx)];
    }
    for (int idx : plan.needed) {
      if (g_.nodes[size_t(idx)].cls != OpClass::Placeholder) continue;
      const TensorArg& f = feeds[size_t(plan.feed_of[size_t(idx)])];
      const size_t rb = row_bytes(idx);
      TFS_CUDA(cudaMemcpyAsync(dev_ptr[size_t(idx)], static_cast<const char*>(f.data) + size_t(r0) * rb,
                               size_t(rows_here) * rb, cudaMemcpyHostToDevice, sx));
    }
    exec_step(cp, 0, dev_ptr, sx);
    for (size_t i = 0; i < fetches.size(); ++i) {
      const int idx = plan.fetch_nodes[i];
      const size_t rb = row_bytes(idx);
      TFS_CUDA(cudaMemcpyAsync(static_cast<char*>(fetches[i].data) + size_t(r0) * rb, dev_ptr[size_t(idx)],
                               size_t(rows_here) * rb, cudaMemcpyDeviceToHost, sx));
    }
  }
  for (int k = 0; k < 2; ++k) TFS_CUDA(cudaStreamSynchronize(pipe_stream_[k]));
  if (plan.may_raise) check_device_error();
  return true;
}

// ---------------------------------------------------------------------------
// Execution
// ---------------------------------------------------------------------------
void Session::run(const std::vector<TensorArg>& feeds, const std::vector<TensorArg>& fetches, bool rows) {
  Plan plan;
  std::vector<std::string> names;
  for (auto& f : fetches) names.push_back(f.name);
  build_plan(feeds, names, plan, rows);
  // row mode: a fetch that does not depend on any placeholder is replicated once per row
  std::vector<char> bcast(fetches.size(), 0);
  auto out_numel = [&](size_t i) {
    const int idx = plan.fetch_nodes[i];
    return plan.shape[size_t(idx)].numel() * (bcast[i] ? plan.n_batch : 1);
  };
  for (size_t i = 0; i < fetches.size(); ++i) bcast[i] = rows && !plan.batched[size_t(plan.fetch_nodes[i])];
  for (size_t i = 0; i < fetches.size(); ++i) {
    int idx = plan.fetch_nodes[i];
    const Node& n = g_.nodes[size_t(idx)];
    TFS_CHECK(fetches[i].data != nullptr || out_numel(i) == 0, TFS_ERR_INVALID,
              "fetch '" + n.name + "' has no output buffer");
    TFS_CHECK(fetches[i].dtype == 0 || fetches[i].dtype == n.dtype, TFS_ERR_INVALID,
              "fetch '" + n.name + "' produces " + dtype_name(n.dtype) + " but the output buffer is declared " +
                  dtype_name(fetches[i].dtype));
    if (fetches[i].shape.known_rank())
      TFS_CHECK(fetches[i].shape.fully_known() && fetches[i].shape.numel() == out_numel(i),
                TFS_ERR_INVALID,
                "fetch '" + n.name + "' has shape " + plan.shape[size_t(idx)].str() + " but the output buffer is declared " +
                    fetches[i].shape.str());
  }
  ensure_device();
  if (!rows && run_pipelined(feeds, fetches, plan)) { mark_work(); return; }
  const size_t N = g_.nodes.size();
  std::vector<void*> dev_ptr(N, nullptr);
  auto nbytes = [&](int idx) { return size_t(plan.shape[size_t(idx)].numel()) * size_t(dtype_size(g_.nodes[size_t(idx)].dtype)); };

  // classify caller pointers
  bool any_host = false;
  std::vector<char> feed_dev(feeds.size(), 0), fetch_dev(fetches.size(), 0);
  for (size_t i = 0; i < feeds.size(); ++i) {
    TFS_CHECK(feeds[i].data != nullptr || feeds[i].shape.numel() == 0, TFS_ERR_INVALID,
              "feed '" + feeds[i].name + "' has a null data pointer");
    feed_dev[i] = on_device_or_refuse(feeds[i].data, dev_.device, "feed '" + feeds[i].name + "'");
    if (!feed_dev[i]) any_host = true;
  }
  for (size_t i = 0; i < fetches.size(); ++i) {
    fetch_dev[i] = on_device_or_refuse(fetches[i].data, dev_.device, "fetch buffer '" + fetches[i].name + "'");
    if (!fetch_dev[i]) any_host = true;
  }

  // workspace layout: staged host feeds, then every materialised node not bound to a caller buffer
  size_t ws_bytes = 0;
  std::vector<size_t> ws_off(N, SIZE_MAX);
  std::vector<char> staged_twice(N, 0);      // a host feed that shares the staged copy of an earlier one
  bool needs_reduce_scratch = false;
  size_t red_bytes = 0;
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    if (plan.feed_of[size_t(idx)] >= 0) {
      int f = plan.feed_of[size_t(idx)];
      if (feed_dev[size_t(f)]) dev_ptr[size_t(idx)] = feeds[size_t(f)].data;
      else {
        // one host buffer fed to several placeholders (reduce_sum and reduce_min of the same column: the X / X_input
        // contract gives every fetch its own placeholder) crosses PCIe once: placeholders are only ever read
        for (int prev : plan.needed) {
          if (prev == idx) break;
          const int pf = plan.feed_of[size_t(prev)];
          if (pf >= 0 && !feed_dev[size_t(pf)] && ws_off[size_t(prev)] != SIZE_MAX && feeds[size_t(pf)].data == feeds[size_t(f)].data &&
              nbytes(prev) == nbytes(idx)) {
            ws_off[size_t(idx)] = ws_off[size_t(prev)];
            staged_twice[size_t(idx)] = 1;
            break;
          }
        }
        if (ws_off[size_t(idx)] == SIZE_MAX) { ws_off[size_t(idx)] = ws_bytes; ws_bytes += align_up(nbytes(idx)); }
      }
    } else if (n.cls == OpClass::Const) {
      // uploaded lazily below
    } else if (plan.materialize[size_t(idx)]) {
      for (size_t i = 0; i < fetches.size() && !dev_ptr[size_t(idx)]; ++i)
        if (plan.fetch_nodes[i] == idx && fetch_dev[i] && !bcast[i]) dev_ptr[size_t(idx)] = fetches[i].data;
      if (!dev_ptr[size_t(idx)]) { ws_off[size_t(idx)] = ws_bytes; ws_bytes += align_up(std::max<size_t>(nbytes(idx), 16)); }
    }
  }
  std::vector<size_t> bc_off(fetches.size(), SIZE_MAX);
  for (size_t i = 0; i < fetches.size(); ++i)
    if (bcast[i] && !fetch_dev[i]) { bc_off[i] = ws_bytes; ws_bytes += align_up(std::max<size_t>(nbytes(plan.fetch_nodes[i]) * size_t(plan.n_batch), 16)); }
  for (const Step& s : plan.steps) {
    if (s.kind == Step::REDUCE_AXIS0) {
      needs_reduce_scratch = true;
      red_bytes = std::max(red_bytes, reduce_scratch_bytes(g_.nodes[size_t(s.nodes[0])].dtype, s.outer, s.inner, dev_));
    }
  }
  reserve_workspace(ws_bytes);
  for (int idx : plan.needed)
    if (ws_off[size_t(idx)] != SIZE_MAX) dev_ptr[size_t(idx)] = ws_ + ws_off[size_t(idx)];
  if (needs_reduce_scratch && red_bytes > red_scratch_cap_) {
    TFS_CUDA(cudaStreamSynchronize(stream_));
    if (red_scratch_) TFS_CUDA(cudaFree(red_scratch_));
    red_scratch_ = nullptr;
    red_scratch_cap_ = 0;
    TFS_CUDA(cudaMalloc(&red_scratch_, red_bytes));
    TFS_CUDA(cudaMemsetAsync(red_scratch_, 0, 256, stream_));
    red_scratch_cap_ = red_bytes;
  }
  // constants that are read as tensors (anything but scalar immediates inside a group)
  for (int idx : plan.needed) {
    const Node& n = g_.nodes[size_t(idx)];
    if (n.cls == OpClass::Const && plan.materialize[size_t(idx)]) dev_ptr[size_t(idx)] = const_device(idx);
  }

  // host -> device
  for (int idx : plan.needed) {
    if (plan.feed_of[size_t(idx)] < 0) continue;
    int f = plan.feed_of[size_t(idx)];
    if (!feed_dev[size_t(f)] && nbytes(idx) > 0 && !staged_twice[size_t(idx)])
      copy_in(dev_ptr[size_t(idx)], feeds[size_t(f)].data, nbytes(idx));
  }

  // steps
  for (size_t k = 0; k < plan.steps.size(); ++k) exec_step(plan, k, dev_ptr, stream_);

  // results
  for (size_t i = 0; i < fetches.size(); ++i) {
    int idx = plan.fetch_nodes[i];
    size_t bytes = nbytes(idx);
    if (bytes == 0) continue;
    void* src = dev_ptr[size_t(idx)];
    TFS_CHECK(src != nullptr, TFS_ERR_INVALID, "internal: fetch '" + g_.nodes[size_t(idx)].name + "' was not produced");
    if (bcast[i]) {
      if (plan.n_batch == 0) continue;
      // replicate the row-independent value once per row: out[flat] = value[flat % numel]
      void* dst = fetch_dev[i] ? fetches[i].data : static_cast<void*>(ws_ + bc_off[i]);
      EltProgram prog{};
      prog.err = err_flag_;
      const int di = dtype_index(g_.nodes[size_t(idx)].dtype);
      prog.n_bufs = 2;
      prog.bufs[0] = BufRef{src, plan.shape[size_t(idx)].numel(), di, 0, 0};
      prog.bufs[1] = BufRef{dst, 0, di, 0, 0};
      prog.n_instr = 2;
      prog.ins[0] = Instr{};
      prog.ins[0].code = C_LOAD_PER;
      prog.ins[0].arg = 0;
      prog.ins[1] = Instr{};
      prog.ins[1].code = C_STORE;
      prog.ins[1].arg = 1;
      launches_ += launch_eltwise(prog, out_numel(i), dtype_size(g_.nodes[size_t(idx)].dtype) == 8, dev_, stream_);
      if (!fetch_dev[i])
        TFS_CUDA(cudaMemcpyAsync(fetches[i].data, dst, bytes * size_t(plan.n_batch), cudaMemcpyDeviceToHost, stream_));
      continue;
    }
    if (fetch_dev[i]) {
      if (src != fetches[i].data)
        TFS_CUDA(cudaMemcpyAsync(fetches[i].data, src, bytes, cudaMemcpyDeviceToDevice, stream_));
    } else {
      copy_out(fetches[i].data, src, bytes);
    }
  }
  // results on the host: the caller may read them as soon as this returns, so a raised flag must fail the call now
  // (device-resident results are checked by tfs_session_sync); graphs that cannot raise skip the round trip
  mark_work();
  if (any_host) {
    if (plan.may_raise) check_device_error();
    else TFS_CUDA(cudaStreamSynchronize(stream_));     // pinned destinations are written asynchronously
  }
}

// ---------------------------------------------------------------------------
// aggregate: the TensorFlowUDAF of the reference for one block (impl/DebugRowOps.scala:601-695)
// ---------------------------------------------------------------------------
// Blocks whose payload spreads over more than kAggChunkBytes are cut into row ranges that are reduced one after the
// other; their partial groups are then merged by one more keyed reduction (the graph is re-applied to partial rows,
// as the reference's UDAF does when it compacts its buffers, impl/DebugRowOps.scala:644-676).  The gather pass reads
// every payload row once, at an address that is random over the block: measured on B200, it runs at 6.6 TB/s while
// the block spans 64 GB and drops to 5.4 TB/s at 90 GB and 3.1 TB/s at 128 GB (address translation stops covering the
// block), so two 64 GB passes and a merge of 2 x 1e6 partial rows are twice as fast as one 128 GB pass.
constexpr size_t kAggChunkBytes = size_t(64) << 30;

void Session::aggregate(int64_t n_rows, const int64_t* keys, const std::vector<TensorArg>& feeds,
                        const std::vector<TensorArg>& fetches, int64_t max_groups, int64_t* out_keys,
                        int64_t* n_groups_out, const std::vector<int>* feed_of_fetch) {
  TFS_CHECK(n_rows >= 0 && max_groups >= 0 && n_groups_out != nullptr, TFS_ERR_INVALID, "negative row or group count");
  size_t widest = 0;
  for (const TensorArg& f : feeds)
    if (f.shape.fully_known() && f.shape.rank >= 1 && f.shape.d[0] == n_rows && n_rows > 0)
      widest = std::max(widest, size_t(f.shape.numel() / n_rows) * size_t(std::max(dtype_size(f.dtype), 1)));
  size_t chunk_bytes = kAggChunkBytes;
  if (const char* e = std::getenv("TFS_AGG_CHUNK_MB")) {       // (tests cut small blocks with it)
    const long v = std::atol(e);
    if (v > 0) chunk_bytes = size_t(v) << 20;
  }
  if (widest == 0 || size_t(n_rows) * widest <= chunk_bytes) {
    aggregate_block(n_rows, keys, feeds, fetches, max_groups, out_keys, n_groups_out, feed_of_fetch);
    return;
  }
  int64_t chunk_rows = std::max<int64_t>(1024, int64_t(chunk_bytes / widest) / 1024 * 1024);
  const int64_t n_chunks = (n_rows + chunk_rows - 1) / chunk_rows;
  chunk_rows = ((n_rows + n_chunks - 1) / n_chunks + 1023) / 1024 * 1024;       // even chunks, starts stay 16-byte aligned
  ensure_device();
  // per-chunk partial groups: [keys][one column per fetch], each with room for n_chunks * max_groups rows
  std::vector<TensorArg> resolved = fetches;
  {
    std::vector<TensorArg> one = feeds;
    for (TensorArg& f : one) f.shape.d[0] = 1;
    resolve(one, resolved, false);          // cell shape and dtype of every fetch
  }
  const size_t cap_rows = size_t(n_chunks) * size_t(max_groups);
  size_t off = align_up(cap_rows * 8);
  std::vector<size_t> off_col(fetches.size()), row_bytes(fetches.size());
  for (size_t i = 0; i < fetches.size(); ++i) {
    row_bytes[i] = size_t(resolved[i].shape.numel()) * size_t(dtype_size(resolved[i].dtype));
    off_col[i] = off;
    off += align_up(cap_rows * row_bytes[i]);
  }
  char* parts = reserve_exchange(2, off);
  int64_t total = 0;
  for (int64_t c = 0; c < n_chunks; ++c) {
    const int64_t r0 = c * chunk_rows, rows_here = std::min(chunk_rows, n_rows - r0);
    std::vector<TensorArg> cf = feeds, co = fetches;
    for (TensorArg& f : cf) {
      const size_t rb = size_t(f.shape.numel() / n_rows) * size_t(dtype_size(f.dtype));
      f.data = static_cast<char*>(f.data) + size_t(r0) * rb;
      f.shape.d[0] = rows_here;
    }
    for (size_t i = 0; i < co.size(); ++i) {
      co[i].data = parts + off_col[i] + size_t(total) * row_bytes[i];
      co[i].shape = with_lead(max_groups, resolved[i].shape);
      co[i].dtype = resolved[i].dtype;
    }
    int64_t g = 0;
    {
      // the partial groups of a range need no order: the merge below sorts its own output, and the partial rows of a key
      // are folded in range order whatever their position inside a range (the row partition is stable)
      struct Flag { bool& f; explicit Flag(bool& x) : f(x) { f = true; } ~Flag() { f = false; } } table_order(agg_table_order_);
      aggregate_block(rows_here, keys + r0, cf, co, max_groups, reinterpret_cast<int64_t*>(parts) + total, &g, feed_of_fetch);
    }
    total += g;
  }
  // merge: the partial rows of a key, in chunk order, through the same graph
  std::vector<TensorArg> mf(fetches.size());
  for (size_t i = 0; i < fetches.size(); ++i) {
    const Node& n = g_.nodes[size_t(g_.find(fetches[i].name))];
    mf[i].name = g_.find(n.name + "_input") >= 0 ? n.name + "_input" : g_.nodes[size_t(n.in[0])].name;
    mf[i].dtype = resolved[i].dtype;
    mf[i].shape = with_lead(total, resolved[i].shape);
    mf[i].data = parts + off_col[i];
  }
  std::vector<int> identity(fetches.size());
  for (size_t i = 0; i < identity.size(); ++i) identity[i] = int(i);
  aggregate_block(total, reinterpret_cast<const int64_t*>(parts), mf, fetches, max_groups, out_keys, n_groups_out, &identity);
}

// What the keyed passes work on: one column of n_rows cells to be reduced per key into `out` (max_groups cells).
struct Session::AggColumn {
  std::string name;      // for messages
  int dtype = 0, red_op = 0;
  int64_t cell = 0;
  const void* payload = nullptr;   // [n_rows, cell], host or device
  void* out = nullptr;             // [max_groups, cell], host or device
};

// aggregate for one block.  The reference hands each key's buffered rows to the user's reducer graph
// (TensorFlowUDAF.compact -> performReduceBlock, impl/DebugRowOps.scala:644-656), so any graph that keeps the X / X_input
// contract works there.  Here a fetch is taken apart into
//     fetch = post( Reduce_axis0( pre(placeholders) ) )
// with `pre` and `post` element-wise (or empty): `pre` runs once over the whole block as an ordinary map, the keyed
// passes reduce its result per key, and `post` runs once over the [n_groups, cell] table of reduced rows, every row being
// one cell (row mode, with the Reduce node fed).  A bare Sum / Min / Max of a placeholder -- the usual case, and
// BASELINE config 4 -- has neither and touches the payload exactly once.
void Session::aggregate_block(int64_t n_rows, const int64_t* keys, const std::vector<TensorArg>& feeds,
                              const std::vector<TensorArg>& fetches, int64_t max_groups, int64_t* out_keys,
                              int64_t* n_groups_out, const std::vector<int>* feed_of_fetch) {
  TFS_CHECK(n_rows >= 0 && max_groups >= 0, TFS_ERR_INVALID, "negative row or group count");
  TFS_CHECK(n_rows < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "aggregate: blocks of 2^32 rows or more are not supported");
  TFS_CHECK(n_rows == 0 || keys != nullptr, TFS_ERR_INVALID, "null key column");
  TFS_CHECK(max_groups == 0 || out_keys != nullptr, TFS_ERR_INVALID, "null out_keys");
  TFS_CHECK(!fetches.empty(), TFS_ERR_INVALID, "Output cannot be empty: no fetch requested");
  const size_t N = g_.nodes.size();
  // which nodes depend on a placeholder
  std::vector<signed char> dep(N, -1);
  std::function<bool(int)> depends = [&](int u) -> bool {
    if (dep[size_t(u)] >= 0) return dep[size_t(u)] != 0;
    const Node& n = g_.nodes[size_t(u)];
    bool d = n.cls == OpClass::Placeholder;
    for (int k : n.in) d = depends(k) || d;
    dep[size_t(u)] = d ? 1 : 0;
    return d;
  };
  struct Item { int fetch_node, red_node, src_node, feed; bool pre, post; };
  std::vector<Item> items;
  std::vector<AggColumn> cols;
  std::vector<Shape> red_in_shape(fetches.size());     // [n_rows, cell...] shape of every reduction's input
  bool any_pre = false, any_post = false;
  for (size_t i = 0; i < fetches.size(); ++i) {
    const int f_idx = g_.find(fetches[i].name);
    TFS_CHECK(f_idx >= 0, TFS_ERR_INVALID, "fetch '" + fetches[i].name + "' is not a node of the graph");
    // walk down from the fetch through element-wise nodes to the one reduction
    int cur = f_idx;
    while (g_.nodes[size_t(cur)].cls != OpClass::Reduce) {
      const Node& c = g_.nodes[size_t(cur)];
      TFS_CHECK(c.cls == OpClass::Unary || c.cls == OpClass::Binary, TFS_ERR_UNSUPPORTED,
                "aggregate: fetch '" + g_.nodes[size_t(f_idx)].name + "' must be a Sum, Min or Max over axis 0, optionally inside "
                "element-wise operations; found '" + c.name + "' (" + c.op + ") on the way to the reduction");
      int data = -1;
      for (int k : c.in)
        if (depends(k)) {
          TFS_CHECK(data < 0 || data == k, TFS_ERR_UNSUPPORTED,
                    "aggregate: '" + c.name + "' (" + c.op + ") combines two reduced values; one reduction per fetch is supported");
          data = k;
        }
      TFS_CHECK(data >= 0, TFS_ERR_UNSUPPORTED,
                "aggregate: fetch '" + g_.nodes[size_t(f_idx)].name + "' does not depend on any placeholder");
      cur = data;
    }
    const Node& r = g_.nodes[size_t(cur)];
    const Node& src = g_.nodes[size_t(r.in[0])];
    Item it{f_idx, cur, r.in[0], -1, src.cls != OpClass::Placeholder, cur != f_idx};
    AggColumn col;
    col.name = g_.nodes[size_t(f_idx)].name;
    col.dtype = r.dtype;
    col.red_op = r.red_op;
    if (!it.pre) {
      if (feed_of_fetch != nullptr) it.feed = (*feed_of_fetch)[i];      // merge of partial rows: column i belongs to fetch i
      else
        for (size_t f = 0; f < feeds.size(); ++f)
          if (g_.find(feeds[f].name) == r.in[0]) it.feed = int(f);
      TFS_CHECK(it.feed >= 0, TFS_ERR_INVALID, "placeholder '" + src.name + "' is needed by the fetches but was not fed");
      const TensorArg& fa = feeds[size_t(it.feed)];
      TFS_CHECK(fa.dtype == src.dtype, TFS_ERR_INVALID,
                "The type of node '" + src.name + "' (" + dtype_name(src.dtype) +
                    ") is not compatible with the data type of the feed (" + dtype_name(fa.dtype) + ")");
      TFS_CHECK(fa.shape.fully_known() && fa.shape.rank >= 1 && fa.shape.d[0] == n_rows, TFS_ERR_INVALID,
                "feed '" + src.name + "' must have " + std::to_string(n_rows) + " rows, got shape " + fa.shape.str());
      std::vector<int> axes = normalize_axes(r, fa.shape.rank);
      TFS_CHECK(axes.size() == 1 && axes[0] == 0, TFS_ERR_UNSUPPORTED, "aggregate: fetch '" + col.name + "' must reduce axis 0 only");
      col.cell = 1;
      for (int d = 1; d < fa.shape.rank; ++d) col.cell *= fa.shape.d[d];
      col.payload = fa.data;
      red_in_shape[i] = fa.shape;
    }
    if (!it.post) {
      col.out = fetches[i].data;
      TFS_CHECK(col.out != nullptr || it.pre || max_groups * col.cell == 0, TFS_ERR_INVALID, "fetch '" + col.name + "' has no output buffer");
    }
    any_pre = any_pre || it.pre;
    any_post = any_post || it.post;
    items.push_back(it);
    cols.push_back(col);
  }
  ensure_device();
  *n_groups_out = 0;
  if (n_rows == 0) return;
  if (!any_pre && !any_post) {
    aggregate_columns(n_rows, keys, cols, max_groups, out_keys, n_groups_out);
    return;
  }
  // ---- pre: one map run over the block for every reduction whose input is computed
  std::vector<TensorArg> pre_fetch;
  std::vector<size_t> pre_of;
  for (size_t i = 0; i < items.size(); ++i)
    if (items[i].pre) {
      TensorArg a;
      a.name = g_.nodes[size_t(items[i].src_node)].name;
      pre_fetch.push_back(a);
      pre_of.push_back(i);
    }
  if (!pre_fetch.empty()) resolve(feeds, pre_fetch, false);
  size_t tmp = 0;
  std::vector<size_t> off_pre(items.size(), SIZE_MAX), off_red(items.size(), SIZE_MAX);
  for (size_t k = 0; k < pre_fetch.size(); ++k) {
    const size_t i = pre_of[k];
    const Shape& sh = pre_fetch[k].shape;
    const Node& r = g_.nodes[size_t(items[i].red_node)];
    TFS_CHECK(sh.rank >= 1 && sh.d[0] == n_rows, TFS_ERR_UNSUPPORTED,
              "aggregate: the input of '" + r.name + "' has shape " + sh.str() + ", not one cell per row of the block");
    std::vector<int> axes = normalize_axes(r, sh.rank);
    TFS_CHECK(axes.size() == 1 && axes[0] == 0, TFS_ERR_UNSUPPORTED, "aggregate: fetch '" + cols[i].name + "' must reduce axis 0 only");
    cols[i].cell = sh.numel() / n_rows;
    red_in_shape[i] = sh;
    off_pre[i] = tmp;
    tmp += align_up(size_t(sh.numel()) * size_t(dtype_size(pre_fetch[k].dtype)));
  }
  for (size_t i = 0; i < items.size(); ++i)
    if (items[i].post) { off_red[i] = tmp; tmp += align_up(size_t(max_groups) * size_t(cols[i].cell) * size_t(dtype_size(cols[i].dtype))); }
  char* tbuf = reserve_exchange(3, std::max<size_t>(tmp, 256));
  if (!pre_fetch.empty()) {
    for (size_t k = 0; k < pre_fetch.size(); ++k) pre_fetch[k].data = tbuf + off_pre[pre_of[k]];
    run(feeds, pre_fetch, false);
    for (size_t k = 0; k < pre_fetch.size(); ++k) cols[pre_of[k]].payload = pre_fetch[k].data;
  }
  for (size_t i = 0; i < items.size(); ++i)
    if (items[i].post) cols[i].out = tbuf + off_red[i];
  int64_t n_groups = 0;
  aggregate_columns(n_rows, keys, cols, max_groups, out_keys, &n_groups);
  // ---- post: the element-wise tail of every fetch over the table of reduced rows, the Reduce node fed
  for (size_t i = 0; i < items.size(); ++i) {
    if (!items[i].post || n_groups == 0) continue;
    const Node& r = g_.nodes[size_t(items[i].red_node)];
    TensorArg feed;
    feed.name = r.name;
    feed.dtype = r.dtype;
    const Shape cell_shape = drop_lead(red_in_shape[i]);
    feed.shape = with_lead(n_groups, cell_shape);
    feed.data = cols[i].out;
    TensorArg out = fetches[i];
    out.shape = Shape();                       // the caller's buffer holds max_groups rows; n_groups of them are written
    out.dtype = 0;
    struct Allow { bool& f; explicit Allow(bool& x) : f(x) { f = true; } ~Allow() { f = false; } } allow(allow_internal_feeds_);
    run({feed}, {out}, true);
  }
  TFS_CUDA(cudaStreamSynchronize(stream_));
  *n_groups_out = n_groups;
}

// The keyed passes (K3) over prepared columns.  Synchronous.
void Session::aggregate_columns(int64_t n_rows, const int64_t* keys, const std::vector<AggColumn>& items, int64_t max_groups,
                                int64_t* out_keys, int64_t* n_groups_out) {
  (void)on_device_or_refuse(out_keys, dev_.device, "out_keys");     // refused before any work, not at the final copy
  TFS_CHECK(max_groups > 0, TFS_ERR_INVALID, "aggregate: max_groups must be positive for a non-empty block");

  AggLayout L = agg_layout(n_rows, max_groups, dev_);
  auto up = [](size_t x) { return align_up(x); };
  // workspace: [agg scratch][out_keys][staged keys][staged feeds][staged outputs]
  size_t off = up(L.bytes);
  const size_t off_out_keys = off; off += up(size_t(max_groups) * 8);
  int64_t widest = 1;
  for (const AggColumn& it : items) widest = std::max(widest, it.cell);
  const size_t split_bytes = agg_split_scratch_bytes(widest, dev_);
  const size_t off_split = off; off += up(split_bytes);
  // scalar payloads (at most two of them) ride through their own sort instead of being gathered
  std::vector<char> carried(items.size(), 0);
  size_t carried_bytes = 0;
  {
    size_t n_scalar = 0;
    auto row_bytes = [&](const AggColumn& it) { return it.cell * int64_t(dtype_size(it.dtype)); };
    for (const AggColumn& it : items) n_scalar += row_bytes(it) <= 8 ? 1 : 0;
    static const bool off_switch = [] { const char* e = std::getenv("TFS_AGG_CARRY"); return e && std::string(e) == "off"; }();
    if (n_scalar > 0 && n_scalar <= 2 && n_rows >= (int64_t(1) << 16) && !off_switch)
      for (size_t k = 0; k < items.size(); ++k)
        if (row_bytes(items[k]) <= 8) {          // one f64 / i64, or up to two f32 / i32 per row
          carried[k] = 1;
          carried_bytes = std::max(carried_bytes, agg_carried_bytes(n_rows, int(row_bytes(items[k])), L));
        }
  }
  const size_t off_carried = off; off += up(carried_bytes);
  // gather-only blocks whose table needs two partition passes or more: the rows of a slot are not counted at insert
  // (one scattered atomic per row), the last partition pass records where every slot's segment begins and ends
  {
    const char* e = std::getenv("TFS_AGG_COUNT");        // (A/B switch, read per call: "insert" counts the rows at insert)
    const bool count_at_insert = e && std::string(e) == "insert";
    if (L.bounds_eligible && carried_bytes == 0 && !count_at_insert) L.rows_counted = false;
  }
  const bool keys_dev = on_device_or_refuse(keys, dev_.device, "the key column");
  size_t off_keys = 0;
  if (!keys_dev) { off_keys = off; off += up(size_t(n_rows) * 8); }
  std::vector<size_t> off_feed(items.size(), SIZE_MAX), off_out(items.size(), SIZE_MAX);
  for (size_t k = 0; k < items.size(); ++k) {
    const size_t es = size_t(dtype_size(items[k].dtype));
    TFS_CHECK(items[k].out != nullptr || max_groups * items[k].cell == 0, TFS_ERR_INVALID, "fetch '" + items[k].name + "' has no output buffer");
    if (!on_device_or_refuse(items[k].payload, dev_.device, "the feed of '" + items[k].name + "'")) { off_feed[k] = off; off += up(size_t(n_rows) * items[k].cell * es); }
    if (!on_device_or_refuse(items[k].out, dev_.device, "fetch buffer '" + items[k].name + "'")) { off_out[k] = off; off += up(size_t(max_groups) * items[k].cell * es); }
  }
  reserve_workspace(off);
  char* scratch = ws_;
  int64_t* d_out_keys = reinterpret_cast<int64_t*>(ws_ + off_out_keys);
  const int64_t* d_keys = keys;
  if (!keys_dev) {
    copy_in(ws_ + off_keys, keys, size_t(n_rows) * 8);
    d_keys = reinterpret_cast<const int64_t*>(ws_ + off_keys);
  }
  launches_ += launch_agg_build(d_keys, n_rows, max_groups, scratch, L, dev_, stream_);
  // the number of groups sizes every later launch: one small readback
  unsigned int header[4] = {0, 0, 0, 0};
  TFS_CUDA(cudaMemcpyAsync(header, scratch + L.off_header, sizeof(header), cudaMemcpyDeviceToHost, stream_));
  TFS_CUDA(cudaStreamSynchronize(stream_));
  TFS_CHECK(header[1] == 0 && int64_t(header[2]) <= max_groups, TFS_ERR_INVALID,
            "aggregate: the block holds more than max_groups = " + std::to_string(max_groups) + " distinct keys");
  const int64_t n_groups = header[2];
  bool any_gather = false;
  for (size_t k = 0; k < items.size(); ++k) any_gather = any_gather || !carried[k];
  launches_ += launch_agg_segments(d_keys, n_rows, n_groups, scratch, L, d_out_keys, agg_table_order_ ? kGroupsTableOrder : kGroupsByKey, false,
                                   dev_, stream_);   // groups by key (partial results of a row range: table order)
  // carried fetches first: they read the unsorted slot column, which the row sort below recycles
  for (int pass = 0; pass < 2; ++pass) {
   if (pass == 1 && any_gather)
     launches_ += launch_agg_segments(d_keys, n_rows, n_groups, scratch, L, d_out_keys, kGroupsKeep, true, dev_, stream_);  // rows by slot
   for (size_t k = 0; k < items.size(); ++k) {
    if ((pass == 0) != (carried[k] != 0)) continue;
    const AggColumn& n = items[k];
    const size_t es = size_t(dtype_size(n.dtype));
    const void* payload = items[k].payload;
    if (off_feed[k] != SIZE_MAX) {
      copy_in(ws_ + off_feed[k], payload, size_t(n_rows) * items[k].cell * es);
      payload = ws_ + off_feed[k];
    }
    void* out = off_out[k] != SIZE_MAX ? static_cast<void*>(ws_ + off_out[k]) : items[k].out;
    if (carried[k])
      launches_ += launch_agg_reduce_carried(n.dtype, n.red_op, payload, items[k].cell, n_groups, n_rows, scratch, L, out, ws_ + off_carried,
                                             carried_bytes, ws_ + off_split, split_bytes, dev_, stream_);
    else
      launches_ += launch_agg_reduce(n.dtype, n.red_op, payload, items[k].cell, n_groups, n_rows, scratch, L, out, ws_ + off_split,
                                     split_bytes, dev_, stream_);
    if (off_out[k] != SIZE_MAX && n_groups * items[k].cell > 0)
      TFS_CUDA(cudaMemcpyAsync(items[k].out, out, size_t(n_groups) * items[k].cell * es,
                               cudaMemcpyDeviceToHost, stream_));
   }
  }
  if (n_groups > 0) {
    const bool ok_dev = on_device_or_refuse(out_keys, dev_.device, "out_keys");
    TFS_CUDA(cudaMemcpyAsync(out_keys, d_out_keys, size_t(n_groups) * 8,
                             ok_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, stream_));
  }
  TFS_CUDA(cudaStreamSynchronize(stream_));
  *n_groups_out = n_groups;
}

// ---------------------------------------------------------------------------
// Exchange steps (one rank per partition)
// ---------------------------------------------------------------------------
char* Session::reserve_exchange(int which, size_t bytes) {
  if (bytes > xchg_cap_[which]) {
    TFS_CUDA(cudaStreamSynchronize(stream_));
    if (xchg_[which]) TFS_CUDA(cudaFree(xchg_[which]));
    xchg_[which] = nullptr;
    xchg_cap_[which] = 0;
    const size_t cap = align_up(bytes + bytes / 4);
    TFS_CUDA(cudaMalloc(reinterpret_cast<void**>(&xchg_[which]), cap));
    xchg_cap_[which] = cap;
  }
  return xchg_[which];
}

void Session::combine_wait(Comm& comm) {
  ensure_device();
  comm.wait_side(stream_);
  mark_work();
}

void Session::reduce_combine(Comm& comm, const std::vector<TensorArg>& partials, const std::vector<TensorArg>& merged, bool overlap) {
  TFS_CHECK(!partials.empty() && partials.size() == merged.size(), TFS_ERR_INVALID,
            "reduce_combine: one merged buffer per partial row is needed");
  ensure_device();
  TFS_CHECK(comm.device() == dev_.device, TFS_ERR_INVALID,
            "reduce_combine: the communicator belongs to device " + std::to_string(comm.device()) + ", the session to device " +
                std::to_string(dev_.device));
  struct Item { int node; size_t count, bytes; bool bare; };
  std::vector<Item> items;
  bool all_bare = true;
  size_t ring_bytes = 0;
  for (size_t i = 0; i < partials.size(); ++i) {
    const int idx = g_.find(partials[i].name);
    TFS_CHECK(idx >= 0, TFS_ERR_INVALID, "fetch '" + partials[i].name + "' is not a node of the graph");
    TFS_CHECK(merged[i].name == partials[i].name, TFS_ERR_INVALID, "reduce_combine: partials and merged must name the same fetches in the same order");
    const Node& n = g_.nodes[size_t(idx)];
    TFS_CHECK(partials[i].shape.fully_known() && merged[i].shape.fully_known() && partials[i].shape.numel() == merged[i].shape.numel(),
              TFS_ERR_INVALID, "reduce_combine: fetch '" + n.name + "': partial " + partials[i].shape.str() + " and merged " +
                                   merged[i].shape.str() + " rows differ in size");
    TFS_CHECK((partials[i].dtype == 0 || partials[i].dtype == n.dtype) && (merged[i].dtype == 0 || merged[i].dtype == n.dtype),
              TFS_ERR_INVALID, "reduce_combine: fetch '" + n.name + "' is " + dtype_name(n.dtype));
    const size_t count = size_t(partials[i].shape.numel());
    if (count > 0) {
      TFS_CHECK(on_device_or_refuse(partials[i].data, dev_.device, "partial row '" + n.name + "'") &&
                    on_device_or_refuse(merged[i].data, dev_.device, "merged row '" + n.name + "'"),
                TFS_ERR_INVALID, "reduce_combine: partial and merged rows must be device memory");
    }
    // Sum / Min / Max taken directly from the placeholder: merging partial rows is the same operation element by element
    const bool bare = n.cls == OpClass::Reduce && g_.nodes[size_t(n.in[0])].cls == OpClass::Placeholder;
    all_bare = all_bare && bare;
    items.push_back({idx, count, count * size_t(dtype_size(n.dtype)), bare});
    ring_bytes += align_up(items.back().bytes);
  }
  if (comm.world() == 1) {       // a single partition: RDD.reduce returns its row untouched
    for (size_t i = 0; i < items.size(); ++i)
      if (items[i].bytes > 0 && partials[i].data != merged[i].data)
        TFS_CUDA(cudaMemcpyAsync(merged[i].data, partials[i].data, items[i].bytes, cudaMemcpyDeviceToDevice, stream_));
    mark_work();
    return;
  }
  if (all_bare) {
    if (overlap && ring_bytes <= Comm::kSlotBytes) {
      // The exchange runs on the communicator's side stream and overlaps whatever the caller enqueues next.  The rows
      // are first copied to a ring slot on the launch stream, so the caller may overwrite `partials` right away.
      int slot = 0;
      char* buf = comm.ring_acquire(stream_, &slot);
      size_t off = 0;
      for (size_t i = 0; i < items.size(); ++i) {
        if (items[i].bytes > 0)
          TFS_CUDA(cudaMemcpyAsync(buf + off, partials[i].data, items[i].bytes, cudaMemcpyDeviceToDevice, stream_));
        off += align_up(items[i].bytes);
      }
      comm.ring_produced(slot, stream_);
      comm.group_start();
      off = 0;
      for (size_t i = 0; i < items.size(); ++i) {
        const Node& n = g_.nodes[size_t(items[i].node)];
        if (items[i].count > 0) comm.all_reduce(buf + off, merged[i].data, items[i].count, n.dtype, n.red_op, comm.side_stream());
        off += align_up(items[i].bytes);
      }
      comm.group_end();
      comm.ring_consumed(slot);
    } else {
      comm.group_start();
      for (size_t i = 0; i < items.size(); ++i) {
        const Node& n = g_.nodes[size_t(items[i].node)];
        if (items[i].count > 0) comm.all_reduce(partials[i].data, merged[i].data, items[i].count, n.dtype, n.red_op, stream_);
      }
      comm.group_end();
    }
    mark_work();
    return;
  }
  // Any other reducer graph: gather every rank's partial rows and run the graph once more on the stacked
  // [world, cell...] block -- the reference's pairwise reducePairBlock merge, done in one run.  In stream order.
  const size_t world = size_t(comm.world());
  size_t total = 0;
  std::vector<size_t> off(items.size());
  for (size_t i = 0; i < items.size(); ++i) { off[i] = total; total += align_up(items[i].bytes * world); }
  char* buf = reserve_exchange(0, total);
  comm.group_start();
  for (size_t i = 0; i < items.size(); ++i)
    if (items[i].bytes > 0) comm.all_gather_bytes(partials[i].data, buf + off[i], items[i].bytes, stream_);
  comm.group_end();
  std::vector<TensorArg> feeds(items.size());
  for (size_t i = 0; i < items.size(); ++i) {
    const Node& n = g_.nodes[size_t(items[i].node)];
    feeds[i].name = n.name + "_input";       // the reduce contract: fetch X <- placeholder X_input (impl/DebugRowOps.scala:89-113)
    TFS_CHECK(g_.find(feeds[i].name) >= 0, TFS_ERR_INVALID,
              "reduce_combine: the graph has no placeholder '" + feeds[i].name + "' for fetch '" + n.name + "'");
    feeds[i].dtype = n.dtype;
    feeds[i].shape = with_lead(int64_t(world), partials[i].shape);
    feeds[i].data = buf + off[i];
  }
  run(feeds, merged, false);
}

// aggregate over several GPUs: after the per-GPU keyed reduction every rank holds partial groups; each group goes to the
// rank that owns its key (owner = (mix64(key) >> 32) mod world) and is merged there by one more keyed reduction.  The reference
// does this with Spark's hash shuffle between the UDAF's partial and final aggregation (impl/DebugRowOps.scala:573-576).
// Partition (stable, by owner), one grouped send / receive per peer, merge: all enqueued from here, no host work per
// group; one small readback (the per-peer counts) sizes the receives.
void Session::aggregate_combine(Comm& comm, int64_t n_groups, const int64_t* keys, const std::vector<TensorArg>& partials,
                                const std::vector<TensorArg>& merged, int64_t max_groups, int64_t* out_keys, int64_t* n_owned) {
  TFS_CHECK(n_groups >= 0 && max_groups >= 0 && n_owned != nullptr, TFS_ERR_INVALID, "aggregate_combine: bad arguments");
  TFS_CHECK(!partials.empty() && partials.size() == merged.size(), TFS_ERR_INVALID,
            "aggregate_combine: one merged buffer per partial column is needed");
  TFS_CHECK(n_groups < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "aggregate_combine: 2^32 or more groups are not supported");
  ensure_device();
  TFS_CHECK(comm.device() == dev_.device, TFS_ERR_INVALID,
            "aggregate_combine: the communicator belongs to device " + std::to_string(comm.device()) + ", the session to device " +
                std::to_string(dev_.device));
  const int world = comm.world(), rank = comm.rank();
  struct Col { int node; std::string feed; size_t row_bytes; Shape cell; };
  std::vector<Col> cols;
  for (size_t i = 0; i < partials.size(); ++i) {
    const int idx = g_.find(partials[i].name);
    TFS_CHECK(idx >= 0, TFS_ERR_INVALID, "fetch '" + partials[i].name + "' is not a node of the graph");
    TFS_CHECK(merged[i].name == partials[i].name, TFS_ERR_INVALID, "aggregate_combine: partials and merged must name the same fetches in the same order");
    const Node& n = g_.nodes[size_t(idx)];
    const Shape& ps = partials[i].shape;
    TFS_CHECK(ps.fully_known() && ps.rank >= 1 && ps.d[0] == n_groups, TFS_ERR_INVALID,
              "aggregate_combine: partial column '" + n.name + "' must have " + std::to_string(n_groups) + " rows, got " + ps.str());
    TFS_CHECK(partials[i].dtype == 0 || partials[i].dtype == n.dtype, TFS_ERR_INVALID, "aggregate_combine: fetch '" + n.name + "' is " + dtype_name(n.dtype));
    if (n_groups > 0)
      TFS_CHECK(on_device_or_refuse(partials[i].data, dev_.device, "partial column '" + n.name + "'"), TFS_ERR_INVALID,
                "aggregate_combine: partial columns must be device memory");
    Col c;
    c.node = idx;
    c.cell = drop_lead(ps);
    c.row_bytes = size_t(c.cell.numel()) * size_t(dtype_size(n.dtype));
    // the reduce contract: fetch X <- placeholder X_input (impl/DebugRowOps.scala:89-113)
    c.feed = n.name + "_input";
    if (g_.find(c.feed) < 0) {
      TFS_CHECK(n.cls == OpClass::Reduce && g_.nodes[size_t(n.in[0])].cls == OpClass::Placeholder, TFS_ERR_INVALID,
                "aggregate_combine: the graph has no placeholder '" + c.feed + "' for fetch '" + n.name + "'");
      c.feed = g_.nodes[size_t(n.in[0])].name;
    }
    cols.push_back(c);
  }
  if (n_groups > 0)
    TFS_CHECK(on_device_or_refuse(keys, dev_.device, "the partial key column"), TFS_ERR_INVALID, "aggregate_combine: keys must be device memory");

  // ---- send side: owner of every group, stable partition by owner, keys and rows gathered in that order
  const size_t ng = size_t(n_groups);
  size_t off = 0;
  const size_t off_counts = off; off += align_up(size_t(world) * 8);
  const size_t off_owner = off; off += align_up(ng * 4);
  const size_t off_perm = off; off += align_up(ng * 4);
  const size_t off_radix = off; off += align_up(radix_scratch_bytes(dev_));
  const size_t off_keys = off; off += align_up(ng * 8);
  std::vector<size_t> off_rows(cols.size());
  for (size_t i = 0; i < cols.size(); ++i) { off_rows[i] = off; off += align_up(ng * cols[i].row_bytes); }
  char* sb = reserve_exchange(0, off);
  unsigned long long* d_counts = reinterpret_cast<unsigned long long*>(sb + off_counts);
  unsigned int* d_owner = reinterpret_cast<unsigned int*>(sb + off_owner);
  unsigned int* d_perm = reinterpret_cast<unsigned int*>(sb + off_perm);
  launches_ += launch_agg_owner(keys, n_groups, world, d_owner, d_counts, dev_, stream_);
  if (n_groups > 0) {
    int bits = 1;
    while ((1 << bits) < world) ++bits;
    launches_ += launch_radix_pass(4, 4, d_owner, nullptr, nullptr, d_perm, n_groups, 0, 0, bits, sb + off_radix, dev_, stream_);
    launches_ += launch_gather_rows(keys, d_perm, n_groups, 8, sb + off_keys, dev_, stream_);
    for (size_t i = 0; i < cols.size(); ++i)
      launches_ += launch_gather_rows(partials[i].data, d_perm, n_groups, int64_t(cols[i].row_bytes), sb + off_rows[i], dev_, stream_);
  }
  // ---- every rank learns how much it receives from whom: all-gather of the count vectors, one small readback
  comm.all_gather_bytes(d_counts, comm.counts_dev(), size_t(world) * 8, stream_);
  TFS_CUDA(cudaMemcpyAsync(comm.counts_host(), comm.counts_dev(), size_t(world) * size_t(world) * 8, cudaMemcpyDeviceToHost, stream_));
  TFS_CUDA(cudaStreamSynchronize(stream_));
  const int64_t* all = comm.counts_host();          // all[src * world + dst]
  std::vector<int64_t> send_off(size_t(world) + 1, 0), recv_off(size_t(world) + 1, 0);
  for (int p = 0; p < world; ++p) {
    send_off[size_t(p) + 1] = send_off[size_t(p)] + all[size_t(rank) * world + p];
    recv_off[size_t(p) + 1] = recv_off[size_t(p)] + all[size_t(p) * world + rank];
  }
  TFS_CHECK(send_off[size_t(world)] == n_groups, TFS_ERR_RUNTIME, "aggregate_combine: owner counts do not add up");
  const int64_t n_recv = recv_off[size_t(world)];
  TFS_CHECK(n_recv < (int64_t(1) << 32), TFS_ERR_UNSUPPORTED, "aggregate_combine: 2^32 or more received groups");
  size_t roff = 0;
  const size_t roff_keys = roff; roff += align_up(size_t(n_recv) * 8);
  std::vector<size_t> roff_rows(cols.size());
  for (size_t i = 0; i < cols.size(); ++i) { roff_rows[i] = roff; roff += align_up(size_t(n_recv) * cols[i].row_bytes); }
  char* rb = reserve_exchange(1, std::max<size_t>(roff, 256));
  comm.group_start();
  for (int p = 0; p < world; ++p) {
    const size_t ns = size_t(send_off[size_t(p) + 1] - send_off[size_t(p)]), nr = size_t(recv_off[size_t(p) + 1] - recv_off[size_t(p)]);
    if (ns > 0) comm.send_bytes(sb + off_keys + size_t(send_off[size_t(p)]) * 8, ns * 8, p, stream_);
    if (nr > 0) comm.recv_bytes(rb + roff_keys + size_t(recv_off[size_t(p)]) * 8, nr * 8, p, stream_);
    for (size_t i = 0; i < cols.size(); ++i) {
      if (cols[i].row_bytes == 0) continue;
      if (ns > 0) comm.send_bytes(sb + off_rows[i] + size_t(send_off[size_t(p)]) * cols[i].row_bytes, ns * cols[i].row_bytes, p, stream_);
      if (nr > 0) comm.recv_bytes(rb + roff_rows[i] + size_t(recv_off[size_t(p)]) * cols[i].row_bytes, nr * cols[i].row_bytes, p, stream_);
    }
  }
  comm.group_end();
  // ---- merge: the keyed reduction once more, on what this rank owns (rows arrive ordered by source rank, each
  // source's rows in ascending key order, so the merge order is reproducible)
  std::vector<TensorArg> feeds(cols.size());
  for (size_t i = 0; i < cols.size(); ++i) {
    feeds[i].name = cols[i].feed;
    feeds[i].dtype = g_.nodes[size_t(cols[i].node)].dtype;
    feeds[i].shape = with_lead(n_recv, cols[i].cell);
    feeds[i].data = rb + roff_rows[i];
  }
  *n_owned = 0;
  if (n_recv == 0) { mark_work(); return; }
  std::vector<int> identity(cols.size());
  for (size_t i = 0; i < identity.size(); ++i) identity[i] = int(i);
  aggregate(n_recv, reinterpret_cast<const int64_t*>(rb + roff_keys), feeds, merged, max_groups, out_keys, n_owned, &identity);
}

}  // namespace tfs
