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
 *   analyzeGraphTF: inputs, fetches, hinted shapes        tfs_graph_analyze
 *     (impl/TensorFlowOps.scala:101-141)
 *   transformRdd.reduce(reducePairBlock(...))             tfs_session_reduce_combine
 *     (impl/DebugRowOps.scala:524-525, 741-750)
 *   data.agg(udaf(...)): partial -> shuffle -> final      tfs_session_aggregate_combine
 *     (impl/DebugRowOps.scala:573-576)
 *   TensorConverter.append / convertBuffer1-3 loops       tfs_pack_rows / tfs_unpack_rows
 *     (impl/DataOps.scala:63-81, impl/datatypes.scala:114-127, 204-256)
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
 *  - several GPUs: one session per device (any mix of processes and threads) and, for the operators that have an
 *    exchange step, one communicator per session (tfs_comm_*, below).  mapBlocks needs none: partitions are
 *    independent (impl/DebugRowOps.scala:377-391).  The exchange steps the reference hands to Spark -- RDD.reduce with
 *    reducePairBlock (impl/DebugRowOps.scala:524-525, 741-750) and the UDAF's shuffle (:573-576) -- are
 *    tfs_session_reduce_combine and tfs_session_aggregate_combine: NCCL over NVLink, driven from inside the library.
 *    Feeds, fetch buffers and keys must be host memory or memory of the session's own device; memory of another
 *    device is TFS_ERR_INVALID.
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

/* analyzeGraphTF (impl/TensorFlowOps.scala:101-141) in one call: the summary of every input -- a Placeholder without
 * inputs (:106-108) -- and of every requested fetch (names with or without a trailing ":0", :110), in GraphDef order.
 * A shape hint for a node (looked up under `name`, then `name:0`) overrides the statically inferred shape (:124-133).
 * A node whose rank is unknown and that has no hint is reported as a scalar (rank 0), which is what Shape.from does
 * with TF's "unknown shape" (Shape.scala:107-109).  A requested fetch that is not in the graph is TFS_ERR_INVALID.
 * `out` has room for `capacity` entries; *n_out is the number of summaries (also when it exceeds capacity, in which
 * case only `capacity` are written).  `name` points into the graph and lives as long as it. */
typedef struct tfs_shape_hint {
  const char* name;
  int32_t rank;
  int64_t dims[TFS_MAX_RANK];
} tfs_shape_hint;
typedef struct tfs_node_summary {
  const char* name;
  int32_t dtype;
  int32_t rank;
  int64_t dims[TFS_MAX_RANK];
  int32_t is_placeholder;
  int32_t is_input;
  int32_t is_output;
  int32_t pad_;
} tfs_node_summary;
int tfs_graph_analyze(const tfs_graph* g, int n_hints, const tfs_shape_hint* hints, int n_fetches,
                      const char* const* requested_fetches, int capacity, tfs_node_summary* out, int* n_out);

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
 *     after the host buffers have been written.
 *   - feeds are only read.  Several placeholders may be fed from one buffer
 *     (reduce_sum and reduce_min of the same column: the X / X_input contract,
 *     impl/DebugRowOps.scala:89-113, gives every fetch its own placeholder); a
 *     host buffer fed more than once is copied to the device once. */
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

/* ---- several GPUs: communicators and the exchange steps ---------------------------
 * One partition per GPU.  A communicator is this rank's end of an NCCL clique (NCCL is loaded at run time --
 * libnccl.so.2, or the name in TFS_NCCL_LIB; a process that never creates a communicator never loads it).
 *   - one rank per process or thread: rank 0 calls tfs_comm_unique_id, the caller carries the 128 bytes to the other
 *     ranks (Spark broadcast, a file, MPI...), then every rank calls tfs_comm_create -- collectively, like
 *     ncclCommInitRank;
 *   - one thread creating the communicators of n devices of one process: tfs_comm_create_all.  The exchange calls of
 *     different ranks must still be made concurrently, one thread per device. */
typedef struct tfs_comm tfs_comm;
#define TFS_COMM_ID_BYTES 128
int tfs_comm_unique_id(void* id_out);
int tfs_comm_create(int device, int world, int rank, const void* id, tfs_comm** out);
int tfs_comm_create_all(int n, const int* devices, tfs_comm** out);
void tfs_comm_free(tfs_comm* c);
int tfs_comm_world(const tfs_comm* c);
int tfs_comm_rank(const tfs_comm* c);

/* The exchange step of reduceBlocks = transformRdd.reduce(reducePairBlock(...)) (impl/DebugRowOps.scala:524-525,
 * 741-750): `partials` are the fetches of this rank's tfs_session_run on its partition (device memory, named by fetch),
 * `merged` receives the row reduced over every rank's partition -- on every rank (same shapes; may alias partials).
 *   - fetches that are a Sum / Min / Max taken directly from their placeholder are combined by one grouped
 *     ncclAllReduce; any other reducer graph by an all-gather of the rows and one more run of the graph on the stacked
 *     [world, cell...] block (fetch X <- placeholder X_input, the reduce contract).
 *   - overlap = 0: enqueued on the session's stream, in order.  overlap = 1 (all-reduce path only, rows of at most
 *     64 KiB in total): the rows are copied to a ring slot and exchanged on the communicator's own stream, so the
 *     exchange overlaps whatever the session runs next -- the next partition's kernels -- and `partials` may be
 *     overwritten at once; `merged` is complete after tfs_session_combine_wait + tfs_session_sync.
 * Collective: every rank calls it with the same fetches.  Returns without synchronising. */
int tfs_session_reduce_combine(tfs_session* s, tfs_comm* c, int n_fetches, const tfs_tensor* partials,
                               const tfs_tensor* merged, int overlap);
/* the session's stream waits for the overlapped exchanges enqueued so far */
int tfs_session_combine_wait(tfs_session* s, tfs_comm* c);

/* The exchange step of aggregate = the shuffle between the UDAF's partial and final aggregation
 * (impl/DebugRowOps.scala:573-576): `keys[n_groups]` and `partials` (dims[0] == n_groups, device memory) are this
 * rank's tfs_session_aggregate result; every group is sent to the rank that owns its key (a fixed 64-bit mix of the
 * key, modulo the number of ranks) and merged there by one more keyed reduction.  `merged` (dims[0] == max_groups,
 * device or host), out_keys and *n_owned receive the groups this rank owns, in ascending key order; the result stays
 * sharded by owner (the reference leaves row order unspecified).  Collective and synchronous. */
int tfs_session_aggregate_combine(tfs_session* s, tfs_comm* c, int64_t n_groups, const int64_t* keys,
                                  int n_fetches, const tfs_tensor* partials, const tfs_tensor* merged,
                                  int64_t max_groups, int64_t* out_keys, int64_t* n_owned);

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

/* Row <-> block packing = the loops of DataOps.convertFast0 / TensorConverter.append (impl/DataOps.scala:63-81,
 * impl/datatypes.scala:85-127) and of convertBuffer1-3 (impl/datatypes.scala:204-256), for a caller that holds one
 * array per row (a JVM binding pins each row's primitive array): rows[i] points at the `cell_elems` scalars of row i,
 * `block` is the dense row-major [n_rows, cell_elems] buffer (ideally from tfs_host_alloc).  Plain copies on a few
 * host threads; rows of a block share one cell shape, as the reference requires (impl/datatypes.scala:80-83). */
int tfs_pack_rows(int32_t dtype, int64_t n_rows, int64_t cell_elems, const void* const* rows, void* block);
int tfs_unpack_rows(int32_t dtype, int64_t n_rows, int64_t cell_elems, const void* block, void* const* rows);

#ifdef __cplusplus
}
#endif
#endif /* TFS_B200_H */
