/*
 * Native.java -- Panama (java.lang.foreign, JDK 22+) binding of libtfsb200.so for the TensorFrames JVM side.
 *
 * NOT COMPILED IN THIS REPOSITORY: the build environment has no JDK.  It is shipped as the source a maintainer of
 * yupbank/tensorframes drops next to org/tensorframes/impl/DebugRowOps.scala; every downcall below has a ctypes twin
 * in tensorframes_b200/_native.py that the test-suite runs, and the struct layouts are pinned by
 * tests/test_c_caller.py::test_struct_layouts_a_binding_is_written_against (tfs_tensor = 88 bytes: name 0, dtype 8,
 * rank 12, dims 16, data 80; tfs_shape_hint = 80 bytes; tfs_node_summary = 96 bytes).
 *
 * Each method replaces one TensorFlow-Java call of the reference (paths under src/main/scala/org/tensorframes/):
 *   graphImport          new Graph(); g.importGraphDef(bytes)            impl/TensorFlowOps.scala:87-95
 *   graphAnalyze         op.output(0).dataType()/.shape() + shape hints   impl/TensorFlowOps.scala:101-150
 *   sessionCreate/Free   new Session(g) / close()                         impl/TensorFlowOps.scala:76-85
 *   sessionResolve/Run   runner.fetch(..).feed(..).run(); Tensor.writeTo  impl/DebugRowOps.scala:787-794, 900-917
 *   sessionAggregate     TensorFlowUDAF update / merge / evaluate         impl/DebugRowOps.scala:601-695
 *   commCreate, reduceCombine, aggregateCombine   RDD.reduce / the UDAF shuffle    impl/DebugRowOps.scala:524-525, 573-576
 */
package org.tensorframes.impl.b200;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemoryLayout;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;
import java.lang.invoke.VarHandle;

public final class Native {
  private Native() {}

  public static final int TFS_MAX_RANK = 8;
  public static final int DT_FLOAT = 1, DT_DOUBLE = 2, DT_INT32 = 3, DT_INT64 = 9;   // types.proto:16-24
  public static final int COMM_ID_BYTES = 128;

  private static final Linker LINKER = Linker.nativeLinker();
  private static final SymbolLookup LIB = SymbolLookup.libraryLookup(
      System.getProperty("tensorframes.b200.lib", "libtfsb200.so"), Arena.global());

  private static MethodHandle h(String name, FunctionDescriptor d) {
    return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
  }

  // struct tfs_tensor { const char* name; int32 dtype; int32 rank; int64 dims[8]; void* data; }
  public static final MemoryLayout TENSOR = MemoryLayout.structLayout(
      ADDRESS.withName("name"), JAVA_INT.withName("dtype"), JAVA_INT.withName("rank"),
      MemoryLayout.sequenceLayout(TFS_MAX_RANK, JAVA_LONG).withName("dims"), ADDRESS.withName("data"));
  // struct tfs_shape_hint { const char* name; int32 rank; (4 bytes padding) int64 dims[8]; }
  public static final MemoryLayout SHAPE_HINT = MemoryLayout.structLayout(
      ADDRESS.withName("name"), JAVA_INT.withName("rank"), MemoryLayout.paddingLayout(4),
      MemoryLayout.sequenceLayout(TFS_MAX_RANK, JAVA_LONG).withName("dims"));
  // struct tfs_node_summary { const char* name; int32 dtype, rank; int64 dims[8]; int32 is_placeholder, is_input, is_output, pad_; }
  public static final MemoryLayout NODE_SUMMARY = MemoryLayout.structLayout(
      ADDRESS.withName("name"), JAVA_INT.withName("dtype"), JAVA_INT.withName("rank"),
      MemoryLayout.sequenceLayout(TFS_MAX_RANK, JAVA_LONG).withName("dims"), JAVA_INT.withName("is_placeholder"),
      JAVA_INT.withName("is_input"), JAVA_INT.withName("is_output"), JAVA_INT.withName("pad_"));

  static final VarHandle T_NAME = TENSOR.varHandle(MemoryLayout.PathElement.groupElement("name"));
  static final VarHandle T_DTYPE = TENSOR.varHandle(MemoryLayout.PathElement.groupElement("dtype"));
  static final VarHandle T_RANK = TENSOR.varHandle(MemoryLayout.PathElement.groupElement("rank"));
  static final VarHandle T_DATA = TENSOR.varHandle(MemoryLayout.PathElement.groupElement("data"));
  static final long T_DIMS_OFFSET = TENSOR.byteOffset(MemoryLayout.PathElement.groupElement("dims"));

  private static final MethodHandle LAST_ERROR = h("tfs_last_error", FunctionDescriptor.of(ADDRESS));
  private static final MethodHandle DEVICE_COUNT = h("tfs_device_count", FunctionDescriptor.of(JAVA_INT));
  private static final MethodHandle GRAPH_IMPORT = h("tfs_graph_import", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS));
  private static final MethodHandle GRAPH_FREE = h("tfs_graph_free", FunctionDescriptor.ofVoid(ADDRESS));
  private static final MethodHandle GRAPH_NUM_NODES = h("tfs_graph_num_nodes", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  private static final MethodHandle GRAPH_ANALYZE = h("tfs_graph_analyze",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS));
  private static final MethodHandle SESSION_CREATE = h("tfs_session_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
  private static final MethodHandle SESSION_FREE = h("tfs_session_free", FunctionDescriptor.ofVoid(ADDRESS));
  private static final MethodHandle SESSION_SYNC = h("tfs_session_sync", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  private static final MethodHandle SESSION_RESOLVE = h("tfs_session_resolve",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
  private static final MethodHandle SESSION_RUN = h("tfs_session_run",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
  private static final MethodHandle SESSION_RESOLVE_ROWS = h("tfs_session_resolve_rows",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
  private static final MethodHandle SESSION_RUN_ROWS = h("tfs_session_run_rows",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
  // int tfs_session_aggregate(session, n_rows, keys, n_feeds, feeds, n_fetches, fetches, max_groups, out_keys, &n_groups)
  private static final MethodHandle SESSION_AGGREGATE = h("tfs_session_aggregate",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
  private static final MethodHandle COMM_UNIQUE_ID = h("tfs_comm_unique_id", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  private static final MethodHandle COMM_CREATE = h("tfs_comm_create",
      FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS));
  private static final MethodHandle COMM_FREE = h("tfs_comm_free", FunctionDescriptor.ofVoid(ADDRESS));
  private static final MethodHandle REDUCE_COMBINE = h("tfs_session_reduce_combine",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_INT));
  private static final MethodHandle COMBINE_WAIT = h("tfs_session_combine_wait", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
  private static final MethodHandle AGGREGATE_COMBINE = h("tfs_session_aggregate_combine",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
  private static final MethodHandle HOST_ALLOC = h("tfs_host_alloc", FunctionDescriptor.of(JAVA_INT, JAVA_LONG, ADDRESS));
  private static final MethodHandle HOST_FREE = h("tfs_host_free", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  private static final MethodHandle PACK_ROWS = h("tfs_pack_rows",
      FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS));
  private static final MethodHandle UNPACK_ROWS = h("tfs_unpack_rows",
      FunctionDescriptor.of(JAVA_INT, JAVA_INT, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS));

  /** The reference throws plain Exceptions for every failure (impl/DebugRowOps.scala:54-60); so does the binding. */
  static void check(int status) {
    if (status == 0) return;
    try {
      MemorySegment msg = (MemorySegment) LAST_ERROR.invokeExact();
      throw new RuntimeException(msg.reinterpret(Long.MAX_VALUE).getString(0));
    } catch (RuntimeException e) {
      throw e;
    } catch (Throwable t) {
      throw new RuntimeException("libtfsb200 call failed with status " + status, t);
    }
  }

  private static int call(MethodHandle mh, Object... args) {
    try {
      return (int) mh.invokeWithArguments(args);
    } catch (Throwable t) {
      throw new RuntimeException(t);
    }
  }

  public static int deviceCount() { return call(DEVICE_COUNT); }

  public static MemorySegment graphImport(Arena arena, byte[] graphDef) {
    MemorySegment bytes = arena.allocate(Math.max(1, graphDef.length));
    MemorySegment.copy(graphDef, 0, bytes, java.lang.foreign.ValueLayout.JAVA_BYTE, 0, graphDef.length);
    MemorySegment out = arena.allocate(ADDRESS);
    check(call(GRAPH_IMPORT, bytes, (long) graphDef.length, out));
    return out.get(ADDRESS, 0);
  }

  public static void graphFree(MemorySegment graph) { try { GRAPH_FREE.invokeExact(graph); } catch (Throwable t) { throw new RuntimeException(t); } }

  public static int graphNumNodes(MemorySegment graph) { return call(GRAPH_NUM_NODES, graph); }

  /** tfs_graph_analyze: `hints` / `out` are arrays of SHAPE_HINT / NODE_SUMMARY; returns the number of summaries. */
  public static int graphAnalyze(Arena arena, MemorySegment graph, int nHints, MemorySegment hints, String[] fetches,
                                 int capacity, MemorySegment out) {
    MemorySegment names = arena.allocate(ADDRESS, Math.max(1, fetches.length));
    for (int i = 0; i < fetches.length; ++i) names.setAtIndex(ADDRESS, i, arena.allocateFrom(fetches[i]));
    MemorySegment n = arena.allocate(JAVA_INT);
    check(call(GRAPH_ANALYZE, graph, nHints, hints, fetches.length, names, capacity, out, n));
    return n.get(JAVA_INT, 0);
  }

  public static MemorySegment sessionCreate(Arena arena, MemorySegment graph, int device) {
    MemorySegment out = arena.allocate(ADDRESS);
    check(call(SESSION_CREATE, graph, device, out));
    return out.get(ADDRESS, 0);
  }

  public static void sessionFree(MemorySegment session) { try { SESSION_FREE.invokeExact(session); } catch (Throwable t) { throw new RuntimeException(t); } }

  public static void sessionSync(MemorySegment session) { check(call(SESSION_SYNC, session)); }

  /** Fills entry `i` of a tfs_tensor array. `data` is a direct buffer's segment (or NULL for resolve). */
  public static void setTensor(Arena arena, MemorySegment array, int i, String name, int dtype, long[] dims, MemorySegment data) {
    MemorySegment t = array.asSlice((long) i * TENSOR.byteSize(), TENSOR.byteSize());
    T_NAME.set(t, 0L, arena.allocateFrom(name));
    T_DTYPE.set(t, 0L, dtype);
    T_RANK.set(t, 0L, dims.length);
    for (int d = 0; d < dims.length; ++d) t.set(JAVA_LONG, T_DIMS_OFFSET + 8L * d, dims[d]);
    T_DATA.set(t, 0L, data == null ? MemorySegment.NULL : data);
  }

  public static long[] tensorDims(MemorySegment array, int i) {
    MemorySegment t = array.asSlice((long) i * TENSOR.byteSize(), TENSOR.byteSize());
    int rank = (int) T_RANK.get(t, 0L);
    long[] dims = new long[rank];
    for (int d = 0; d < rank; ++d) dims[d] = t.get(JAVA_LONG, T_DIMS_OFFSET + 8L * d);
    return dims;
  }

  public static int tensorDtype(MemorySegment array, int i) {
    return (int) T_DTYPE.get(array.asSlice((long) i * TENSOR.byteSize(), TENSOR.byteSize()), 0L);
  }

  public static void sessionResolve(MemorySegment s, int nFeeds, MemorySegment feeds, int nFetches, MemorySegment fetches, boolean rows) {
    check(call(rows ? SESSION_RESOLVE_ROWS : SESSION_RESOLVE, s, nFeeds, feeds, nFetches, fetches));
  }

  public static void sessionRun(MemorySegment s, int nFeeds, MemorySegment feeds, int nFetches, MemorySegment fetches, boolean rows) {
    check(call(rows ? SESSION_RUN_ROWS : SESSION_RUN, s, nFeeds, feeds, nFetches, fetches));
  }

  /** Returns the number of groups; out_keys[0..n) ascending, row g of every fetch belongs to out_keys[g]. */
  public static long sessionAggregate(Arena arena, MemorySegment s, long nRows, MemorySegment keys, int nFeeds, MemorySegment feeds,
                                      int nFetches, MemorySegment fetches, long maxGroups, MemorySegment outKeys) {
    MemorySegment n = arena.allocate(JAVA_LONG);
    check(call(SESSION_AGGREGATE, s, nRows, keys, nFeeds, feeds, nFetches, fetches, maxGroups, outKeys, n));
    return n.get(JAVA_LONG, 0);
  }

  // ---- several GPUs: one communicator per session; the id travels by Spark broadcast ----------------------------------
  public static byte[] commUniqueId(Arena arena) {
    MemorySegment id = arena.allocate(COMM_ID_BYTES);
    check(call(COMM_UNIQUE_ID, id));
    return id.toArray(java.lang.foreign.ValueLayout.JAVA_BYTE);
  }

  public static MemorySegment commCreate(Arena arena, int device, int world, int rank, byte[] id) {
    MemorySegment seg = arena.allocate(COMM_ID_BYTES);
    MemorySegment.copy(id, 0, seg, java.lang.foreign.ValueLayout.JAVA_BYTE, 0, COMM_ID_BYTES);
    MemorySegment out = arena.allocate(ADDRESS);
    check(call(COMM_CREATE, device, world, rank, seg, out));
    return out.get(ADDRESS, 0);
  }

  public static void commFree(MemorySegment comm) { try { COMM_FREE.invokeExact(comm); } catch (Throwable t) { throw new RuntimeException(t); } }

  public static void reduceCombine(MemorySegment s, MemorySegment comm, int n, MemorySegment partials, MemorySegment merged, boolean overlap) {
    check(call(REDUCE_COMBINE, s, comm, n, partials, merged, overlap ? 1 : 0));
  }

  public static void combineWait(MemorySegment s, MemorySegment comm) { check(call(COMBINE_WAIT, s, comm)); }

  public static long aggregateCombine(Arena arena, MemorySegment s, MemorySegment comm, long nGroups, MemorySegment keys, int n,
                                      MemorySegment partials, MemorySegment merged, long maxGroups, MemorySegment outKeys) {
    MemorySegment owned = arena.allocate(JAVA_LONG);
    check(call(AGGREGATE_COMBINE, s, comm, nGroups, keys, n, partials, merged, maxGroups, outKeys, owned));
    return owned.get(JAVA_LONG, 0);
  }

  // ---- host memory: pinned blocks DMA at PCIe rate; pageable direct buffers are staged by the library ------------------
  public static MemorySegment hostAlloc(Arena arena, long bytes) {
    MemorySegment out = arena.allocate(ADDRESS);
    check(call(HOST_ALLOC, bytes, out));
    return out.get(ADDRESS, 0).reinterpret(bytes);
  }

  public static void hostFree(MemorySegment p) { check(call(HOST_FREE, p)); }

  /** rows[i] -> block[i, :]; `rows` is an array of n pointers (pinned primitive arrays or direct buffers). */
  public static void packRows(int dtype, long nRows, long cellElems, MemorySegment rows, MemorySegment block) {
    check(call(PACK_ROWS, dtype, nRows, cellElems, rows, block));
  }

  public static void unpackRows(int dtype, long nRows, long cellElems, MemorySegment block, MemorySegment rows) {
    check(call(UNPACK_ROWS, dtype, nRows, cellElems, block, rows));
  }
}
