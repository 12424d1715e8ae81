// Session: one imported graph bound to one CUDA device.  It plans a run for the
// concrete feed shapes (fused element-wise groups, reductions, matmul), owns the
// stream / workspace / uploaded constants, and executes the steps.
//
// Stand-in for TF-Java's Session + Runner as the reference uses them
// (impl/TensorFlowOps.scala:76-85, impl/DebugRowOps.scala:783-802, 900-917).
#pragma once
#include <memory>
#include <unordered_map>
#include <vector>

#include "comm.h"
#include "graph.h"
#include "kernels.h"

namespace tfs {

struct TensorArg {
  std::string name;
  int dtype = 0;
  Shape shape;
  void* data = nullptr;
};

class Session {
 public:
  Session(const Graph& g, int device);
  ~Session();
  Session(const Session&) = delete;
  Session& operator=(const Session&) = delete;

  void set_stream(cudaStream_t s);
  void sync();
  int64_t launches() const { return launches_; }

  // Fills fetches[i].dtype / .shape for the given feed shapes. No CUDA work.
  // rows = true: the graph describes ONE row (placeholders are cell-shaped) and is mapped over the
  // leading dimension of the feeds; every fetch comes back with one cell per row (mapRows).
  void resolve(const std::vector<TensorArg>& feeds, std::vector<TensorArg>& fetches, bool rows = false);
  void run(const std::vector<TensorArg>& feeds, const std::vector<TensorArg>& fetches, bool rows = false);
  void aggregate(int64_t n_rows, const int64_t* keys, const std::vector<TensorArg>& feeds,
                 const std::vector<TensorArg>& fetches, int64_t max_groups, int64_t* out_keys,
                 int64_t* n_groups, const std::vector<int>* feed_of_fetch = nullptr);

  // one block of at most ~64 GB of payload through the keyed-reduction passes (aggregate cuts larger blocks)
  void aggregate_block(int64_t n_rows, const int64_t* keys, const std::vector<TensorArg>& feeds,
                       const std::vector<TensorArg>& fetches, int64_t max_groups, int64_t* out_keys,
                       int64_t* n_groups, const std::vector<int>* feed_of_fetch = nullptr);

  struct AggColumn;
  void aggregate_columns(int64_t n_rows, const int64_t* keys, const std::vector<AggColumn>& cols, int64_t max_groups,
                         int64_t* out_keys, int64_t* n_groups);

  // ---- the exchange steps over a communicator (one rank per partition) ----------------------------------------
  // reduceBlocks: merge this rank's partial result rows with every other rank's; every rank receives the merged
  // row (RDD.reduce + reducePairBlock, impl/DebugRowOps.scala:524-525, 741-750).  partials / merged: device memory.
  void reduce_combine(Comm& comm, const std::vector<TensorArg>& partials, const std::vector<TensorArg>& merged, bool overlap);
  // the session's stream waits for the exchanges that were left running on the communicator's side stream
  void combine_wait(Comm& comm);
  // aggregate: every partial group travels to the rank that owns its key and is merged there
  // (the UDAF's shuffle between partial and final aggregation, impl/DebugRowOps.scala:573-576)
  void aggregate_combine(Comm& comm, int64_t n_groups, const int64_t* keys, const std::vector<TensorArg>& partials,
                         const std::vector<TensorArg>& merged, int64_t max_groups, int64_t* out_keys, int64_t* n_owned);

 private:
  struct Plan;
  void ensure_device();
  void build_plan(const std::vector<TensorArg>& feeds, const std::vector<std::string>& fetch_names, Plan& plan,
                  bool rows = false);
  std::unordered_map<int, std::pair<float*, float*>>::iterator ensure_tc_weights(const Node& n, const float* b, int64_t Nc, int64_t K,
                                                                                 cudaStream_t stream);
  void exec_step(const Plan& plan, size_t step_index, std::vector<void*>& dev_ptr, cudaStream_t stream);
  bool run_pipelined(const std::vector<TensorArg>& feeds, const std::vector<TensorArg>& fetches, const Plan& plan);
  void* const_device(int node);
  void copy_in(void* dst_dev, const void* src_host, size_t bytes);     // host -> device on stream_
  void copy_out(void* dst_host, const void* src_dev, size_t bytes);    // device -> host on stream_
  void reserve_workspace(size_t bytes);
  void check_device_error();
  void mark_work();
  char* reserve_exchange(int which, size_t bytes);

  Graph g_;
  DeviceInfo dev_;
  bool device_ready_ = false;
  cudaStream_t own_stream_ = nullptr;
  cudaStream_t stream_ = nullptr;
  cudaStream_t pipe_stream_[2] = {nullptr, nullptr};   // host ingest pipeline (chunked H2D / kernel / D2H)
  cudaEvent_t pipe_event_ = nullptr;
  cudaEvent_t work_event_ = nullptr;   // recorded after everything a call enqueues: orders a stream switch and the destructor
  char* ws_ = nullptr;
  size_t ws_cap_ = 0;
  void* red_scratch_ = nullptr;
  size_t red_scratch_cap_ = 0;
  // staging grown on demand: [0] gathered rows / send side of an exchange, [1] receive side, [2] per-chunk partial groups
  char* xchg_[4] = {nullptr, nullptr, nullptr, nullptr};      // ([3]: computed reduction inputs / reduced rows of general reducer graphs)
  size_t xchg_cap_[4] = {0, 0, 0, 0};
  bool agg_table_order_ = false;        // aggregate over row ranges: the partial groups of a range stay in table order
  bool allow_internal_feeds_ = false;   // the post-map of aggregate feeds the Reduce node itself
  char* seg_scratch_ = nullptr;        // UnsortedSegmentSum: keyed-reduction scratch, grown on demand
  size_t seg_scratch_cap_ = 0;
  int32_t* err_flag_ = nullptr;       // device
  int32_t* err_host_ = nullptr;       // pinned
  std::unordered_map<int, void*> const_dev_;
  std::unordered_map<int, std::pair<float*, float*>> tc_weights_;   // MatMul constants split into tf32 hi / lo parts
  int64_t launches_ = 0;
};

}  // namespace tfs
