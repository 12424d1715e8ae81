/*
 * tfs_b200.h -- C ABI of libtfsb200.so, the B200-native replacement for the
 * TensorFlow-Java layer that TensorFrames' block executors sit on.
 *
 * Every entry point below stands in for one TF-Java call made by the reference
 * (paths are under /root/reference/src/main/scala/org/tensorframes/):
 *
 *   reference call                                        replaced by
 *   ----------------------------------------------------  -------------------------
 *   new Graph(); g.importGraphDef(bytes)                  tfs_graph_import
 *     (impl/TensorFlowOps.scala:87-95)
 *   g.operation(name).output(0).dataType()/.shape()       tfs_graph_node_info
 *     (impl/TensorFlowOps.scala:115-150)
 *   g.close()                                             tfs_graph_free
 *   new Session(g) / s.close()                            tfs_session_create/_free
 *     (impl/TensorFlowOps.scala:76-85)
 *   Tensor.create(shape, buffer)  (feed)                  tfs_tensor{data,dims} in
 *     (impl/datatypes.scala:344-347,407-410,468-471,526-529)  tfs_session_run feeds
 *   runner.fetch(n)...feed(n,t)...run()                   tfs_session_run
 *     (impl/DebugRowOps.scala:787-794, 900-917)
 *   ... the same, once per row (mapRows)                  tfs_session_run_rows
 *     (impl/DebugRowOps.scala:832-856)
 *   t.numElements(); t.writeTo(buffer)   (fetch)          tfs_session_resolve +
 *     (impl/datatypes.scala:365-370,438-443,488-493,546-551) fetch buffers of _run
 *   TensorFlowUDAF update/merge/evaluate                  tfs_session_aggregate
 *     (impl/DebugRowOps.scala:601-695)
 *
 * Conventions
 *  - plain C, no callbacks, no C++/torch types; all sizes are int64.
 *  - every function returns TFS_OK (0) or an error code; the message is
 *    retrieved with tfs_last_error() (thread local, valid until the next call
 *    on that thread).  The reference throws JVM exceptions instead
 *    (impl/DebugRowOps.scala:54-60); a binding turns a non-zero status into one.
 *  - tensors are row-major, little-endian, densely packed; a block of a
 *    DataFrame column is [n_rows, cell dims...] exactly as TFDataOps.convert
 *    produces (impl/TFDataOps.scala:27-59).
 *  - tensor memory is caller-owned.  `data` may be a host pointer (pageable or
 *    pinned) or a CUDA device pointer on the session's device; the library
 *    inspects it with cudaPointerGetAttributes.  Feed data is never aliased
 *    after the call returns (same contract as Tensor.create, which copies).
 *  - distinct handles may be used concurrently from distinct threads; one
 *    session must not run two calls at the same time (the reference uses one
 *    Session per task thread, impl/DebugRowOps.scala:783).
 *  - there is no CPU execution path: every run needs a CUDA device.
 *  - several GPUs: one session per device (any mix of processes and threads).  The library moves no data between
 *    devices; the exchange steps of the reference stay where they are -- RDD.reduce / the UDAF's shuffle on the JVM
 *    side (impl/DebugRowOps.scala:524-525, 573-576), torch.distributed in this repository's harness -- and their
 *    merge halves are the same two entry points on the gathered data: tfs_session_run on the stacked
 *    [n_partitions, cell...] partial rows (reducePairBlock, :741-750) and tfs_session_aggregate on the
 *    concatenated (key, partial row) pairs (TensorFlowUDAF.merge, :659-672).
 */
#ifndef TFS_B200_H
#define TFS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TFS_OK              0
#define TFS_ERR_INVALID     1  /* bad argument, malformed graph, failed validation */
#define TFS_ERR_UNSUPPORTED 2  /* op / dtype / shape outside the supported subset   */
#define TFS_ERR_CUDA        3  /* a CUDA runtime call failed                        */
#define TFS_ERR_RUNTIME     4  /* data dependent failure (integer division by zero) */

/* dtype codes are TensorFlow DataType enum values
 * (src/main/protobuf/tensorflow/core/framework/types.proto:16-24), the four the
 * reference supports for blocks (impl/datatypes.scala:265-267). */
#define TFS_FLOAT   1
#define TFS_DOUBLE  2
#define TFS_INT32   3
#define TFS_INT64   9

#define TFS_MAX_RANK 8

typedef struct tfs_graph tfs_graph;
typedef struct tfs_session tfs_session;

/* Static description of a graph node's first output, as analyzeGraphTF reads it
 * (impl/TensorFlowOps.scala:101-150).  rank == -1 means unknown rank; a dim of
 * -1 means unknown (org.tensorframes.Shape.Unknown, Shape.scala:96). */
typedef struct tfs_node_info {
  int32_t dtype;
  int32_t rank;
  int64_t dims[TFS_MAX_RANK];
  int32_t is_placeholder; /* op == "Placeholder"                                  */
  int32_t n_inputs;       /* number of data inputs (0 for Placeholder and Const)   */
} tfs_node_info;

/* A dense tensor argument: a feed (name = placeholder) or a fetch (name = node). */
typedef struct tfs_tensor {
  const char* name;
  int32_t dtype;
  int32_t rank;
  int64_t dims[TFS_MAX_RANK];
  void* data;             /* host or device pointer; may be NULL for _resolve      */
} tfs_tensor;

/* ---- errors / library info ------------------------------------------------ */
const char* tfs_last_error(void);
/* library version, major*10000 + minor*100 + patch */
int tfs_version(void);
/* number of visible CUDA devices (0 when there is none); never fails */
int tfs_device_count(void);

/* ---- graph: GraphDef bytes -> analysed graph -------------------------------
 * Replaces Graph.importGraphDef (impl/TensorFlowOps.scala:88-89).  `bytes` is a
 * binary GraphDef (graph.proto:14-112).  Import fails with TFS_ERR_INVALID on
 * malformed protobuf / dangling inputs, and with TFS_ERR_UNSUPPORTED naming the
 * first node whose op is outside the subset.  No CUDA call is made. */
int tfs_graph_import(const void* bytes, size_t len, tfs_graph** out);
void tfs_graph_free(tfs_graph* g);
int tfs_graph_num_nodes(const tfs_graph* g);
/* name / op of the i-th node in GraphDef order; NULL when out of range */
const char* tfs_graph_node_name(const tfs_graph* g, int i);
const char* tfs_graph_node_op(const tfs_graph* g, int i);
/* dtype + statically inferred shape of `name` (a trailing ":0" is accepted, any
 * other output index is TFS_ERR_INVALID: only output 0 is addressable,
 * impl/TensorFlowOps.scala:127-129). */
int tfs_graph_node_info(const tfs_graph* g, const char* name, tfs_node_info* out);

/* ---- session: one graph bound to one device --------------------------------
 * Replaces new Session(graph) (impl/TensorFlowOps.scala:78).  The session owns
 * a stream, a device workspace that grows on demand and the uploaded Const
 * tensors.  `device` is a CUDA ordinal. */
int tfs_session_create(const tfs_graph* g, int device, tfs_session** out);
void tfs_session_free(tfs_session* s);
/* Run on a caller-provided cudaStream_t instead of the session's own stream.
 * NULL restores the own stream -- so the legacy default stream (handle 0) cannot be
 * selected; pass a created stream, cudaStreamLegacy or cudaStreamPerThread.  Lets a
 * caller time with events on the stream the kernels run on.
 * The session's own stream is non-blocking: it does not wait for the legacy default stream.  Device-resident feeds
 * must therefore be complete before a call, or be produced on the stream handed over here.
 * Switching streams is ordered: the new stream first waits for the work earlier calls left pending on the old one
 * (runs share the session's workspace).  A caller-provided stream may be destroyed before the session is freed;
 * tfs_session_free waits for the session's own pending work through an event, not through that stream handle. */
int tfs_session_set_stream(tfs_session* s, void* cuda_stream);
/* Wait for everything enqueued by this session; reports deferred device-side
 * errors (TFS_ERR_RUNTIME) of asynchronous runs. */
int tfs_session_sync(tfs_session* s);
/* kernels launched by this session since creation (own kernels only, memcpys
 * are not counted) */
int64_t tfs_session_launch_count(const tfs_session* s);

/* Concrete output shapes for concrete feed shapes: fills dtype/rank/dims of each
 * fetch (matched by name).  Feed `data` is ignored.  This is what lets a caller
 * size the buffers it hands to tfs_session_run, the way the reference sizes its
 * arrays from Tensor.numElements (impl/datatypes.scala:365-367) and
 * DataOps.inferPhysicalShape (impl/DataOps.scala:105-144). */
int tfs_session_resolve(tfs_session* s, int n_feeds, const tfs_tensor* feeds,
                        int n_fetches, tfs_tensor* fetches);

/* One Session.run: feed every placeholder the fetches depend on, compute the
 * fetches, write each into fetches[i].data (capacity = product(dims) elements of
 * the resolved shape; dims given by the caller are checked against it).
 *   - validation (TFS_ERR_INVALID): unknown names, a needed placeholder not fed,
 *     dtype different from the placeholder's (no implicit cast,
 *     impl/DebugRowOps.scala:328-333), shape not matching the placeholder's
 *     declared shape (Shape.checkMorePreciseThan, Shape.scala:54-59).
 *   - if every feed and fetch pointer is a device pointer the call only
 *     enqueues work on the session stream and returns; otherwise it returns
 *     after the host buffers have been written. */
int tfs_session_run(tfs_session* s, int n_feeds, const tfs_tensor* feeds,
                    int n_fetches, const tfs_tensor* fetches);

/* Row-wise execution = DebugRowOpsImpl.performMapRows (impl/DebugRowOps.scala:819-857), which runs one
 * Session.run per row.  Here the graph still describes ONE row -- placeholders have the cell's shape
 * (impl/DebugRowOps.scala:424-428) -- but a whole block of rows is evaluated in one call: every feed is
 * [n_rows, cell...], every fetch comes back as [n_rows, fetch cell...] (a fetch that does not depend on
 * any placeholder is replicated per row).  All rows of one call share a cell shape; a caller with
 * variable-length cells (TFDataOps.convert(Row,...), impl/TFDataOps.scala:86-105) calls once per shape.
 * Supported per-row graphs: element-wise ops between same-shaped cells, scalars and constants, and
 * reductions over all axes of a cell; anything else is TFS_ERR_UNSUPPORTED (use the block form). */
int tfs_session_resolve_rows(tfs_session* s, int n_feeds, const tfs_tensor* feeds,
                             int n_fetches, tfs_tensor* fetches);
int tfs_session_run_rows(tfs_session* s, int n_feeds, const tfs_tensor* feeds,
                         int n_fetches, const tfs_tensor* fetches);

/* Keyed block reduction = the whole TensorFlowUDAF (impl/DebugRowOps.scala:601-695)
 * for one block: rows with equal keys[i] are reduced together by the session's
 * graph, which must follow the reduce contract (fetch X <- placeholder X_input,
 * impl/DebugRowOps.scala:80-170) and reduce along axis 0 with Sum, Min or Max.
 *   keys      : int64[n_rows], host or device
 *   feeds     : one per placeholder, dims[0] == n_rows
 *   fetches   : dims[0] of each fetch buffer is the group capacity `max_groups`;
 *               row g of every fetch belongs to out_keys[g]
 *   out_keys  : int64[max_groups]
 *   n_groups  : number of distinct keys found (host int64)
 * Keys are compared bit-exactly; group order is unspecified (as in the
 * reference, whose tests sort rows before comparing). Synchronous. */
int tfs_session_aggregate(tfs_session* s, int64_t n_rows, const int64_t* keys,
                          int n_feeds, const tfs_tensor* feeds,
                          int n_fetches, const tfs_tensor* fetches,
                          int64_t max_groups, int64_t* out_keys, int64_t* n_groups);

/* ---- block reduction without any allocation ------------------------------------
 * (No counterpart in the reference: TF-Java gives its callers no say over allocation.)
 *
 * For callers that may not allocate device memory when they call in -- an executor whose GPU
 * memory is wholly owned by a pool allocator, a task that must not cause the device-wide
 * synchronisation a cudaMalloc/cudaFree implies, a process with a fixed memory budget.
 * Reduces in[outer, inner] (row-major, dtype = TFS_FLOAT/DOUBLE/INT32/INT64) along axis 0 into
 * out[inner] with Sum, Min or Max -- what tfs_session_run does for `reduce_sum(x_input, [0])` --
 * but `in` and `out` must already be device pointers on `device`, and the kernel's scratch is a
 * fixed array inside the library image instead of a grown workspace.
 *   - limit: inner <= 256 elements (TFS_ERR_UNSUPPORTED beyond; outer is unlimited); the array
 *     (about 0.5 MiB) stays resident on every GPU that loads the library until process exit;
 *   - no cudaMalloc / cudaFree / allocating call is made; the work is enqueued on `cuda_stream`
 *     (NULL = the legacy default stream) and the call returns without synchronising;
 *   - the first call for a device looks up the array's device address: make it at start-up,
 *     before any stream capture;
 *   - every call on a device shares that one array: calls whose kernels could overlap on the
 *     GPU (other streams, other threads) must be serialised by the caller. */
#define TFS_REDUCE_SUM 0
#define TFS_REDUCE_MIN 1
#define TFS_REDUCE_MAX 2
int tfs_reduce_axis0_noalloc(int device, void* cuda_stream, int32_t dtype, int32_t red_op,
                             const void* in, int64_t outer, int64_t inner, void* out);

/* ---- host memory helpers ----------------------------------------------------
 * Page-locked host buffers so that packed column blocks DMA at full PCIe rate;
 * the counterpart of TensorConverter.reserve (impl/datatypes.scala:334-338). */
int tfs_host_alloc(size_t bytes, void** out);
int tfs_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* TFS_B200_H */
