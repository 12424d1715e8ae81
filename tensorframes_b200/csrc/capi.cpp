// extern "C" surface of libtfsb200.so (declared in include/tfs_b200.h).
// Nothing but plain C types crosses this boundary; C++ exceptions are turned
// into status codes + a thread-local message here.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "comm.h"
#include "graph.h"
#include "session.h"

struct tfs_graph { tfs::Graph g; };
struct tfs_session { tfs::Session* s; };
struct tfs_comm { tfs::Comm* c; };

namespace {

thread_local std::string g_last_error;

template <class F>
int guarded(F&& f) {
  try {
    f();
    g_last_error.clear();
    return TFS_OK;
  } catch (const tfs::Error& e) {
    g_last_error = e.what();
    return e.code;
  } catch (const std::bad_alloc&) {
    g_last_error = "out of host memory";
    return TFS_ERR_RUNTIME;
  } catch (const std::exception& e) {
    g_last_error = std::string("unexpected: ") + e.what();
    return TFS_ERR_RUNTIME;
  } catch (...) {                     // nothing may unwind into a C / JNI caller
    g_last_error = "unexpected: unknown exception";
    return TFS_ERR_RUNTIME;
  }
}

tfs::Shape shape_from(int rank, const int64_t* dims) {
  tfs::Shape s;
  TFS_CHECK(rank >= -1 && rank <= TFS_MAX_RANK, TFS_ERR_INVALID, "tensor rank out of range: " + std::to_string(rank));
  s.rank = rank;
  for (int i = 0; i < rank; ++i) s.d[i] = dims[i];
  return s;
}

std::vector<tfs::TensorArg> args_from(int n, const tfs_tensor* t, const char* what) {
  TFS_CHECK(n >= 0 && (n == 0 || t != nullptr), TFS_ERR_INVALID, std::string("null ") + what + " array");
  std::vector<tfs::TensorArg> out(static_cast<size_t>(n));
  for (int i = 0; i < n; ++i) {
    TFS_CHECK(t[i].name != nullptr, TFS_ERR_INVALID, std::string(what) + " without a name");
    out[size_t(i)].name = t[i].name;
    out[size_t(i)].dtype = t[i].dtype;
    out[size_t(i)].shape = shape_from(t[i].rank, t[i].dims);
    out[size_t(i)].data = t[i].data;
  }
  return out;
}

}  // namespace

namespace {
// rows <-> block copies, split over a few host threads when the block is large
template <class F>
void for_row_ranges(int64_t n_rows, size_t row_bytes, F&& body) {
  const size_t total = size_t(n_rows) * row_bytes;
  unsigned hw = std::thread::hardware_concurrency();
  int threads = total < (size_t(8) << 20) ? 1 : int(std::min<unsigned>(hw ? hw : 1, 8));
  if (threads <= 1) { body(int64_t(0), n_rows); return; }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] { body(n_rows * t / threads, n_rows * (t + 1) / threads); });
  for (auto& th : pool) th.join();
}
}  // namespace

namespace tfs {
// lets an entry point defined in another translation unit report through tfs_last_error()
void set_last_error(const std::string& msg) { g_last_error = msg; }
}  // namespace tfs

extern "C" {

const char* tfs_last_error(void) { return g_last_error.c_str(); }

int tfs_version(void) { return 100; /* 0.1.0 */ }

int tfs_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int tfs_graph_import(const void* bytes, size_t len, tfs_graph** out) {
  return guarded([&] {
    TFS_CHECK(out != nullptr, TFS_ERR_INVALID, "null output handle");
    *out = nullptr;
    tfs::Graph g = tfs::import_graphdef(static_cast<const uint8_t*>(bytes), len);
    *out = new tfs_graph{std::move(g)};
  });
}

void tfs_graph_free(tfs_graph* g) { delete g; }

int tfs_graph_num_nodes(const tfs_graph* g) { return g ? int(g->g.nodes.size()) : 0; }

const char* tfs_graph_node_name(const tfs_graph* g, int i) {
  if (!g || i < 0 || size_t(i) >= g->g.nodes.size()) return nullptr;
  return g->g.nodes[size_t(i)].name.c_str();
}

const char* tfs_graph_node_op(const tfs_graph* g, int i) {
  if (!g || i < 0 || size_t(i) >= g->g.nodes.size()) return nullptr;
  return g->g.nodes[size_t(i)].op.c_str();
}

int tfs_graph_node_info(const tfs_graph* g, const char* name, tfs_node_info* out) {
  return guarded([&] {
    TFS_CHECK(g && name && out, TFS_ERR_INVALID, "null argument");
    int idx = g->g.find(name);
    TFS_CHECK(idx >= 0, TFS_ERR_INVALID, std::string("node '") + name + "' is not in the graph");
    const tfs::Node& n = g->g.nodes[size_t(idx)];
    out->dtype = n.dtype;
    out->rank = n.sshape.rank;
    for (int i = 0; i < TFS_MAX_RANK; ++i) out->dims[i] = (i < n.sshape.rank) ? n.sshape.d[i] : 0;
    out->is_placeholder = n.cls == tfs::OpClass::Placeholder;
    out->n_inputs = int(n.in.size());
  });
}

int tfs_graph_analyze(const tfs_graph* g, int n_hints, const tfs_shape_hint* hints, int n_fetches,
                      const char* const* requested_fetches, int capacity, tfs_node_summary* out, int* n_out) {
  return guarded([&] {
    TFS_CHECK(g && n_out && n_hints >= 0 && n_fetches >= 0 && capacity >= 0, TFS_ERR_INVALID, "bad argument");
    TFS_CHECK((n_hints == 0 || hints) && (n_fetches == 0 || requested_fetches) && (capacity == 0 || out), TFS_ERR_INVALID, "null array");
    auto strip0 = [](std::string n) { return n.size() > 2 && n.compare(n.size() - 2, 2, ":0") == 0 ? n.substr(0, n.size() - 2) : n; };
    std::unordered_set<std::string> outputs;
    for (int i = 0; i < n_fetches; ++i) {
      TFS_CHECK(requested_fetches[i] != nullptr, TFS_ERR_INVALID, "null fetch name");
      outputs.insert(strip0(requested_fetches[i]));
    }
    std::unordered_map<std::string, const tfs_shape_hint*> hinted;
    for (int i = 0; i < n_hints; ++i) {
      TFS_CHECK(hints[i].name != nullptr && hints[i].rank >= 0 && hints[i].rank <= TFS_MAX_RANK, TFS_ERR_INVALID, "bad shape hint");
      hinted.emplace(hints[i].name, &hints[i]);
    }
    int count = 0;
    std::unordered_set<std::string> found;
    for (const tfs::Node& n : g->g.nodes) {
      const bool is_ph = n.cls == tfs::OpClass::Placeholder;
      const bool is_input = is_ph && n.in.empty();
      const bool is_output = outputs.count(n.name) != 0;
      if (!is_input && !is_output) continue;
      if (is_output) found.insert(n.name);
      if (count < capacity) {
        tfs_node_summary& o = out[count];
        o.name = n.name.c_str();
        o.dtype = n.dtype;
        o.is_placeholder = is_ph;
        o.is_input = is_input;
        o.is_output = is_output;
        o.pad_ = 0;
        auto h = hinted.find(n.name);
        if (h == hinted.end()) h = hinted.find(n.name + ":0");
        for (int d = 0; d < TFS_MAX_RANK; ++d) o.dims[d] = 0;
        if (h != hinted.end()) {                     // the hinted shape wins (impl/TensorFlowOps.scala:124-133)
          o.rank = h->second->rank;
          for (int d = 0; d < o.rank; ++d) o.dims[d] = h->second->dims[d];
        } else if (n.sshape.known_rank()) {
          o.rank = n.sshape.rank;
          for (int d = 0; d < o.rank; ++d) o.dims[d] = n.sshape.d[d];
        } else {
          o.rank = 0;                                // unknown rank reads as the empty shape (Shape.scala:107-109)
        }
      }
      ++count;
    }
    if (found.size() != outputs.size()) {
      std::vector<std::string> missing;
      for (const std::string& o : outputs)
        if (!found.count(o)) missing.push_back(o);
      std::sort(missing.begin(), missing.end());
      std::string msg = "Requested fetches are not in the graph: ";
      for (size_t i = 0; i < missing.size(); ++i) msg += (i ? ", " : "") + missing[i];
      tfs::fail(TFS_ERR_INVALID, msg);
    }
    *n_out = count;
  });
}

int tfs_session_create(const tfs_graph* g, int device, tfs_session** out) {
  return guarded([&] {
    TFS_CHECK(g && out, TFS_ERR_INVALID, "null argument");
    *out = nullptr;
    TFS_CHECK(device >= 0, TFS_ERR_INVALID, "negative device ordinal");
    *out = new tfs_session{new tfs::Session(g->g, device)};
  });
}

void tfs_session_free(tfs_session* s) {
  if (!s) return;
  delete s->s;
  delete s;
}

int tfs_session_set_stream(tfs_session* s, void* cuda_stream) {
  return guarded([&] {
    TFS_CHECK(s, TFS_ERR_INVALID, "null session");
    s->s->set_stream(static_cast<cudaStream_t>(cuda_stream));
  });
}

int tfs_session_sync(tfs_session* s) {
  return guarded([&] {
    TFS_CHECK(s, TFS_ERR_INVALID, "null session");
    s->s->sync();
  });
}

int64_t tfs_session_launch_count(const tfs_session* s) { return s ? s->s->launches() : 0; }

static int resolve_impl(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, tfs_tensor* fetches, bool rows) {
  return guarded([&] {
    TFS_CHECK(s, TFS_ERR_INVALID, "null session");
    auto f = args_from(n_feeds, feeds, "feed");
    TFS_CHECK(n_fetches >= 0 && (n_fetches == 0 || fetches), TFS_ERR_INVALID, "null fetch array");
    std::vector<tfs::TensorArg> out(static_cast<size_t>(n_fetches));
    for (int i = 0; i < n_fetches; ++i) {
      TFS_CHECK(fetches[i].name != nullptr, TFS_ERR_INVALID, "fetch without a name");
      out[size_t(i)].name = fetches[i].name;
    }
    s->s->resolve(f, out, rows);
    for (int i = 0; i < n_fetches; ++i) {
      fetches[i].dtype = out[size_t(i)].dtype;
      fetches[i].rank = out[size_t(i)].shape.rank;
      for (int d = 0; d < TFS_MAX_RANK; ++d) fetches[i].dims[d] = d < out[size_t(i)].shape.rank ? out[size_t(i)].shape.d[d] : 0;
    }
  });
}

int tfs_session_resolve(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, tfs_tensor* fetches) {
  return resolve_impl(s, n_feeds, feeds, n_fetches, fetches, false);
}

int tfs_session_resolve_rows(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, tfs_tensor* fetches) {
  return resolve_impl(s, n_feeds, feeds, n_fetches, fetches, true);
}

static int run_impl(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, const tfs_tensor* fetches, bool rows) {
  return guarded([&] {
    TFS_CHECK(s, TFS_ERR_INVALID, "null session");
    auto f = args_from(n_feeds, feeds, "feed");
    auto o = args_from(n_fetches, fetches, "fetch");
    s->s->run(f, o, rows);
  });
}

int tfs_session_run(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, const tfs_tensor* fetches) {
  return run_impl(s, n_feeds, feeds, n_fetches, fetches, false);
}

int tfs_session_run_rows(tfs_session* s, int n_feeds, const tfs_tensor* feeds, int n_fetches, const tfs_tensor* fetches) {
  return run_impl(s, n_feeds, feeds, n_fetches, fetches, true);
}

int tfs_session_aggregate(tfs_session* s, int64_t n_rows, const int64_t* keys, int n_feeds, const tfs_tensor* feeds,
                          int n_fetches, const tfs_tensor* fetches, int64_t max_groups, int64_t* out_keys,
                          int64_t* n_groups) {
  return guarded([&] {
    TFS_CHECK(s && n_groups, TFS_ERR_INVALID, "null argument");
    auto f = args_from(n_feeds, feeds, "feed");
    auto o = args_from(n_fetches, fetches, "fetch");
    s->s->aggregate(n_rows, keys, f, o, max_groups, out_keys, n_groups);
  });
}

int tfs_comm_unique_id(void* id_out) {
  return guarded([&] { tfs::Comm::unique_id(id_out); });
}

int tfs_comm_create(int device, int world, int rank, const void* id, tfs_comm** out) {
  return guarded([&] {
    TFS_CHECK(out != nullptr, TFS_ERR_INVALID, "null output handle");
    *out = nullptr;
    *out = new tfs_comm{tfs::Comm::create(device, world, rank, id)};
  });
}

int tfs_comm_create_all(int n, const int* devices, tfs_comm** out) {
  return guarded([&] {
    TFS_CHECK(out != nullptr && n >= 1 && devices != nullptr, TFS_ERR_INVALID, "bad argument");
    std::vector<tfs::Comm*> cs(static_cast<size_t>(n), nullptr);
    tfs::Comm::create_all(n, devices, cs.data());
    for (int i = 0; i < n; ++i) out[i] = new tfs_comm{cs[size_t(i)]};
  });
}

void tfs_comm_free(tfs_comm* c) {
  if (!c) return;
  delete c->c;
  delete c;
}

int tfs_comm_world(const tfs_comm* c) { return c ? c->c->world() : 0; }
int tfs_comm_rank(const tfs_comm* c) { return c ? c->c->rank() : -1; }

int tfs_session_reduce_combine(tfs_session* s, tfs_comm* c, int n_fetches, const tfs_tensor* partials, const tfs_tensor* merged,
                               int overlap) {
  return guarded([&] {
    TFS_CHECK(s && c, TFS_ERR_INVALID, "null argument");
    auto p = args_from(n_fetches, partials, "partial row");
    auto m = args_from(n_fetches, merged, "merged row");
    s->s->reduce_combine(*c->c, p, m, overlap != 0);
  });
}

int tfs_session_combine_wait(tfs_session* s, tfs_comm* c) {
  return guarded([&] {
    TFS_CHECK(s && c, TFS_ERR_INVALID, "null argument");
    s->s->combine_wait(*c->c);
  });
}

int tfs_session_aggregate_combine(tfs_session* s, tfs_comm* c, int64_t n_groups, const int64_t* keys, int n_fetches,
                                  const tfs_tensor* partials, const tfs_tensor* merged, int64_t max_groups, int64_t* out_keys,
                                  int64_t* n_owned) {
  return guarded([&] {
    TFS_CHECK(s && c && n_owned, TFS_ERR_INVALID, "null argument");
    auto p = args_from(n_fetches, partials, "partial column");
    auto m = args_from(n_fetches, merged, "merged column");
    s->s->aggregate_combine(*c->c, n_groups, keys, p, m, max_groups, out_keys, n_owned);
  });
}

int tfs_pack_rows(int32_t dtype, int64_t n_rows, int64_t cell_elems, const void* const* rows, void* block) {
  return guarded([&] {
    TFS_CHECK(tfs::dtype_supported(dtype), TFS_ERR_UNSUPPORTED, "dtype must be float, double, int32 or int64");
    TFS_CHECK(n_rows >= 0 && cell_elems >= 0, TFS_ERR_INVALID, "negative size");
    const size_t rb = size_t(cell_elems) * size_t(tfs::dtype_size(dtype));
    if (n_rows == 0 || rb == 0) return;
    TFS_CHECK(rows != nullptr && block != nullptr, TFS_ERR_INVALID, "null argument");
    for (int64_t i = 0; i < n_rows; ++i) TFS_CHECK(rows[i] != nullptr, TFS_ERR_INVALID, "row " + std::to_string(i) + " is null");
    for_row_ranges(n_rows, rb, [&](int64_t lo, int64_t hi) {
      for (int64_t i = lo; i < hi; ++i) memcpy(static_cast<char*>(block) + size_t(i) * rb, rows[i], rb);
    });
  });
}

int tfs_unpack_rows(int32_t dtype, int64_t n_rows, int64_t cell_elems, const void* block, void* const* rows) {
  return guarded([&] {
    TFS_CHECK(tfs::dtype_supported(dtype), TFS_ERR_UNSUPPORTED, "dtype must be float, double, int32 or int64");
    TFS_CHECK(n_rows >= 0 && cell_elems >= 0, TFS_ERR_INVALID, "negative size");
    const size_t rb = size_t(cell_elems) * size_t(tfs::dtype_size(dtype));
    if (n_rows == 0 || rb == 0) return;
    TFS_CHECK(rows != nullptr && block != nullptr, TFS_ERR_INVALID, "null argument");
    for (int64_t i = 0; i < n_rows; ++i) TFS_CHECK(rows[i] != nullptr, TFS_ERR_INVALID, "row " + std::to_string(i) + " is null");
    for_row_ranges(n_rows, rb, [&](int64_t lo, int64_t hi) {
      for (int64_t i = lo; i < hi; ++i) memcpy(rows[i], static_cast<const char*>(block) + size_t(i) * rb, rb);
    });
  });
}

int tfs_host_alloc(size_t bytes, void** out) {
  return guarded([&] {
    TFS_CHECK(out, TFS_ERR_INVALID, "null argument");
    *out = nullptr;
    TFS_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
  });
}

int tfs_host_free(void* p) {
  return guarded([&] {
    if (p) TFS_CUDA(cudaFreeHost(p));
  });
}

}  // extern "C"
