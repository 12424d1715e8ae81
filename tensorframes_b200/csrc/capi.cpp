// extern "C" surface of libtfsb200.so (declared in include/tfs_b200.h).
// Nothing but plain C types crosses this boundary; C++ exceptions are turned
// into status codes + a thread-local message here.
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "graph.h"
#include "session.h"

struct tfs_graph { tfs::Graph g; };
struct tfs_session { tfs::Session* s; };

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
