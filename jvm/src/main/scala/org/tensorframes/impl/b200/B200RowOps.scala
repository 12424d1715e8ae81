/*
 * B200RowOps.scala -- OperationsInterface over libtfsb200 (the B200 block path) in place of DebugRowOps.
 *
 * NOT COMPILED IN THIS REPOSITORY: there is no JDK, Scala or Spark in the build environment.  This is the source a
 * maintainer of yupbank/tensorframes adds under src/main/scala/org/tensorframes/impl/b200/, next to Native.java, and
 * wires in at the two places the implementation object is chosen:
 *     dsl/Ops.scala:13              private[this] val ops = new org.tensorframes.impl.b200.B200RowOps
 *     python tensorframes/core.py:23   jvm class name "org.tensorframes.impl.b200.B200RowOps"
 * The C entry points it calls are exercised through the same signatures by tensorframes_b200/_native.py (ctypes) and
 * tests/c_caller.c (C99); tensorframes_b200/ops.py is the tested Python mirror of the control flow below.
 *
 * What stays: everything DebugRowOps does on the driver -- analysing the graph against the frame's schema, the error
 * messages, the output schema (fetches sorted by name first, impl/DebugRowOps.scala:351, 368-372), reduceBlocksSchema
 * (:80-170), explain, the DSL and Python interfaces it mixes in.  B200RowOps extends DebugRowOps and replaces only the
 * per-partition workers, i.e. the code that touched org.tensorflow.{Graph, Session, Tensor}:
 *     DebugRowOpsImpl.performMap          (:766-803)  ->  B200Workers.performMap
 *     DebugRowOpsImpl.performReduceBlock  (:872-895)  ->  B200Workers.performReduceBlock
 *     DebugRowOpsImpl.reducePairBlock     (:741-750)  ->  B200Workers.reducePairBlock   (or the NCCL combine, below)
 *     TensorFlowUDAF                      (:601-695)  ->  B200Workers.aggregatePartition + Spark's own shuffle
 *     TensorFlowOps.analyzeGraphTF        (impl/TensorFlowOps.scala:101-141)  ->  B200Workers.analyzeGraph
 */
package org.tensorframes.impl.b200

import java.lang.foreign.{Arena, MemorySegment, ValueLayout}
import java.nio.ByteOrder

import scala.collection.mutable

import org.apache.spark.TaskContext
import org.apache.spark.sql.{DataFrame, RelationalGroupedDataset, Row}
import org.apache.spark.sql.catalyst.expressions.GenericRowWithSchema
import org.apache.spark.sql.functions.col
import org.apache.spark.sql.types._
import org.tensorflow.framework.GraphDef

import org.tensorframes._
import org.tensorframes.impl._

class B200RowOps extends DebugRowOps {
  import SchemaTransforms._

  override def toString: String = "B200RowOps"

  /** mapBlocks / mapBlocksTrimmed: the driver part is DebugRowOps.mapBlocks (:305-372) verbatim except for its two
    * TF touch points -- analyzeGraphTF and performMap -- so it is restated here around the new workers. */
  override def mapBlocks(df: DataFrame, graph: GraphDef, hints: ShapeDescription): DataFrame =
    mapBlocksB200(df, graph, hints, appendInput = true)

  override def mapBlocksTrimmed(df: DataFrame, graph: GraphDef, hints: ShapeDescription): DataFrame =
    mapBlocksB200(df, graph, hints, appendInput = false)

  private def mapBlocksB200(df: DataFrame, graph: GraphDef, hints: ShapeDescription, appendInput: Boolean): DataFrame = {
    val sc = df.sqlContext.sparkContext
    val summary = B200Workers.analyzeGraph(graph.toByteArray, hints).map(x => x.name -> x).toMap
    val inputs = summary.filter(_._2.isInput)
    val outputs = summary.filter(_._2.isOutput)
    val fieldsByName = df.schema.fields.map(f => f.name -> f).toMap
    val cols = df.schema.fieldNames.mkString(", ")
    inputs.values.foreach { in =>                                  // DebugRowOps.scala:320-344, same messages
      val f = fieldsByName.getOrElse(in.name, throw new Exception(
        s"Graph input ${in.name} found, but no column to match it. Dataframe columns: $cols"))
      val stf = ColumnInformation(f).stf.getOrElse(throw new Exception(
        s"Data column ${f.name} has not been analyzed yet, cannot run TF on this dataframe"))
      if (stf.dataType != in.scalarType) throw new Exception(
        s"The type of node '${in.name}' (${stf.dataType}) is not compatible with the data type of the column (${in.scalarType})")
      if (!stf.shape.checkMorePreciseThan(in.shape)) throw new Exception(
        s"The data column '${f.name}' has shape ${stf.shape} (not compatible) with shape ${in.shape} requested by the TF graph")
      if (!in.isPlaceholder) throw new Exception(s"Invalid type for input node ${in.name}. It has to be a placeholder")
    }
    val outputTFSchema = StructType(outputs.values.toSeq.sortBy(_.name).map { out =>
      if (fieldsByName.contains(out.name)) throw new Exception(
        s"TF graph has an output node called '${out.name}', but this column already exists. Input columns: $cols")
      ColumnInformation.structField(out.name, out.scalarType, out.shape)
    }.toArray)
    val colIdxs = df.schema.fieldNames.zipWithIndex.toMap
    val requestedTFInput = inputs.keys.toArray.map(colIdxs)
    val outputSchema = if (appendInput) StructType(outputTFSchema ++ df.schema.fields) else StructType(outputTFSchema)
    val inputSchema = df.schema
    val gBytes = sc.broadcast(graph.toByteArray)
    val rdd = df.rdd.mapPartitions { it =>
      if (it.hasNext) B200Workers.performMap(it.toArray, inputSchema, requestedTFInput, gBytes.value, outputTFSchema, appendInput)
      else Iterator.empty                                           // empty partitions never reach the device (:380-390)
    }
    df.sqlContext.createDataFrame(rdd, outputSchema)
  }

  /** reduceBlocks (:503-526): per-partition block reduction, then the pairwise merge with the same graph. */
  override def reduceBlocks(df: DataFrame, graph: GraphDef, hints: ShapeDescription): Row = {
    val sc = df.sqlContext.sparkContext
    val all = reduceBlocksSchema(df.schema, graph, hints)
    val gBytes = sc.broadcast(graph.toByteArray)
    val partials = df.rdd.mapPartitions { it =>
      val arr = it.toArray
      // (the reference passes a single row through untouched, :515-516, keeping ALL its columns, which the later merge
      //  then indexes wrongly when the frame has extra columns; a 1-row block is simply reduced here)
      if (arr.isEmpty) Iterator.empty
      else Iterator.single(B200Workers.performReduceBlock(arr, all.mapInput, all.mapTFCols, all.output, gBytes.value))
    }
    partials.reduce(B200Workers.reducePairBlock(all.reduceInput, all.output, gBytes.value))
  }

  /** aggregate (:547-592).  Spark keeps the grouping and the shuffle; what changes is that a partition's rows are
    * reduced per key by ONE tfs_session_aggregate call instead of one TF session per 10 buffered rows (:639-656). */
  override def aggregate(data: RelationalGroupedDataset, graph: GraphDef, hints: ShapeDescription): DataFrame = {
    val df = DebugRowOpsImpl.backingDF(data).get
    val all = reduceBlocksSchema(df.schema, graph, hints)
    val keyCols = B200Workers.groupingColumns(data)
    val gBytes = df.sqlContext.sparkContext.broadcast(graph.toByteArray)
    val outFields = all.output.fields
    val keyFields = keyCols.map(df.schema(_))
    val partialSchema = StructType(keyFields ++ outFields)
    // map side: partial groups of every partition; shuffle on the key columns; reduce side: merge the partial groups
    val partial = df.sqlContext.createDataFrame(
      df.rdd.mapPartitions(it => B200Workers.aggregatePartition(it.toArray, df.schema, keyCols, all, gBytes.value)), partialSchema)
    val shuffled = partial.repartition(keyCols.map(col): _*)
    val merged = shuffled.rdd.mapPartitions(it => B200Workers.aggregatePartition(it.toArray, partialSchema, keyCols, all, gBytes.value, merge = true))
    val withMeta = StructType(keyFields ++ outFields.map(f => ColumnInformation(f).merged))     // metadata re-attached (:581-586)
    df.sqlContext.createDataFrame(merged, withMeta)
  }
}

/** The per-partition workers: the only code that talks to the native library. */
object B200Workers extends Logging {
  private def dtypeCode(t: NumericType): Int = t match {            // impl/datatypes.scala:265-324
    case FloatType => Native.DT_FLOAT
    case DoubleType => Native.DT_DOUBLE
    case IntegerType => Native.DT_INT32
    case LongType => Native.DT_INT64
    case other => throw new Exception(s"Type $other is not supported")
  }
  private def elemSize(code: Int): Int = if (code == Native.DT_DOUBLE || code == Native.DT_INT64) 8 else 4

  /** One GPU per executor: the ordinal comes from the executor's environment (or the partition id modulo the count). */
  private def device(): Int =
    sys.env.get("TFS_B200_DEVICE").map(_.toInt).getOrElse(TaskContext.getPartitionId() % math.max(1, Native.deviceCount()))

  /** analyzeGraphTF (impl/TensorFlowOps.scala:101-141) in one native call: tfs_graph_analyze applies the hints. */
  def analyzeGraph(graphDef: Array[Byte], hints: ShapeDescription): Seq[GraphNodeSummary] = {
    val arena = Arena.ofConfined()
    try {
      val g = Native.graphImport(arena, graphDef)
      try {
        val hs = hints.out.toSeq
        val hseg = arena.allocate(Native.SHAPE_HINT, math.max(1, hs.size))
        hs.zipWithIndex.foreach { case ((name, shape), i) =>
          val h = hseg.asSlice(i * Native.SHAPE_HINT.byteSize(), Native.SHAPE_HINT.byteSize())
          h.set(ValueLayout.ADDRESS, 0, arena.allocateFrom(name))
          h.set(ValueLayout.JAVA_INT, 8, shape.numDims)
          shape.dims.zipWithIndex.foreach { case (d, k) => h.set(ValueLayout.JAVA_LONG, 16 + 8L * k, d) }
        }
        val cap = Native.graphNumNodes(g)
        val out = arena.allocate(Native.NODE_SUMMARY, math.max(1, cap))
        val n = Native.graphAnalyze(arena, g, hs.size, hseg, hints.requestedFetches.toArray, cap, out)
        (0 until n).map { i =>
          val s = out.asSlice(i * Native.NODE_SUMMARY.byteSize(), Native.NODE_SUMMARY.byteSize())
          val name = s.get(ValueLayout.ADDRESS, 0).reinterpret(Long.MaxValue).getString(0)
          val rank = s.get(ValueLayout.JAVA_INT, 12)
          val dims = (0 until rank).map(k => s.get(ValueLayout.JAVA_LONG, 16 + 8L * k))
          val st = SupportedOperations.opsFor(org.tensorflow.framework.DataType.forNumber(s.get(ValueLayout.JAVA_INT, 8))).scalarType
          GraphNodeSummary(s.get(ValueLayout.JAVA_INT, 80) != 0, s.get(ValueLayout.JAVA_INT, 84) != 0,
            s.get(ValueLayout.JAVA_INT, 88) != 0, st, Shape(dims: _*), name)
        }
      } finally Native.graphFree(g)
    } finally arena.close()
  }

  /** A block of one column packed row-major into a pinned host buffer: TFDataOps.convert + convertFast0
    * (impl/TFDataOps.scala:27-59, impl/DataOps.scala:63-81) writing into memory the DMA engine can read directly. */
  private def packColumn(arena: Arena, rows: Array[Row], idx: Int, field: StructField): (Int, Array[Long], MemorySegment) = {
    val info = ColumnInformation(field).stf.getOrElse(throw new Exception(s"Could not column information for column $field"))
    val cell = info.shape.tail
    require(!cell.hasUnknown, s"Shape $cell has unknown values.")
    val code = dtypeCode(info.dataType)
    val cellElems = cell.numElements.get
    val seg = Native.hostAlloc(arena, math.max(1L, rows.length.toLong * cellElems * elemSize(code)))
    val buf = seg.asByteBuffer().order(ByteOrder.LITTLE_ENDIAN)
    val conv = SupportedOperations.opsFor(info.dataType)
    rows.foreach(r => conv.appendCell(buf, r, idx, cell))          // the reference's TensorConverter.append, on `buf`
    (code, (rows.length.toLong +: cell.dims).toArray, seg)
  }

  def performMap(input: Array[Row], inputSchema: StructType, inputTFCols: Array[Int], graphDef: Array[Byte],
                 tfOutputSchema: StructType, appendInput: Boolean): Iterator[Row] = {
    if (input.isEmpty) return Iterator.empty
    val arena = Arena.ofConfined()
    try {
      val g = Native.graphImport(arena, graphDef)
      val s = Native.sessionCreate(arena, g, device())
      try {
        val feeds = arena.allocate(Native.TENSOR, math.max(1, inputTFCols.length))
        inputTFCols.zipWithIndex.foreach { case (c, i) =>
          val (code, dims, seg) = packColumn(arena, input, c, inputSchema(c))
          Native.setTensor(arena, feeds, i, inputSchema(c).name, code, dims, seg)
        }
        val names = tfOutputSchema.fieldNames
        val fetches = arena.allocate(Native.TENSOR, names.length)
        names.zipWithIndex.foreach { case (n, i) => Native.setTensor(arena, fetches, i, n, 0, Array.empty[Long], null) }
        Native.sessionResolve(s, inputTFCols.length, feeds, names.length, fetches, false)       // sizes the outputs
        val outs = names.indices.map { i =>
          val dims = Native.tensorDims(fetches, i)
          val code = Native.tensorDtype(fetches, i)
          val seg = Native.hostAlloc(arena, math.max(1L, dims.product * elemSize(code)))
          Native.setTensor(arena, fetches, i, names(i), code, dims, seg)
          (code, dims, seg)
        }
        Native.sessionRun(s, inputTFCols.length, feeds, names.length, fetches, false)          // H2D, kernels, D2H
        // convertBack (impl/TFDataOps.scala:128-169): same row-count checks, rows = [tf columns..., input columns...]
        B200Convert.convertBack(outs, tfOutputSchema, input, inputSchema, appendInput)
      } finally { Native.sessionFree(s); Native.graphFree(g) }
    } finally arena.close()
  }

  def performReduceBlock(input: Array[Row], inputSchema: StructType, inputTFCols: Seq[Int], schema: StructType,
                         graphDef: Array[Byte]): Row = {
    val it = performMap(input, inputSchema, inputTFCols.toArray, graphDef, schema, appendInput = false)
    assert(it.hasNext)
    val r = it.next()
    assert(!it.hasNext)                                            // exactly one output row (:889-891)
    r
  }

  def reducePairBlock(reduceInput: StructType, output: StructType, graphDef: Array[Byte])(a: Row, b: Row): Row =
    performReduceBlock(Array(a, b), reduceInput, reduceInput.indices, output, graphDef)       // :741-750

  def groupingColumns(data: RelationalGroupedDataset): Seq[String] = {
    // the grouping expressions are a private field of RelationalGroupedDataset, read the same way backingDF reads `df` (:707-730)
    val f = data.getClass.getDeclaredField("groupingExprs")
    f.setAccessible(true)
    f.get(data).asInstanceOf[Seq[org.apache.spark.sql.catalyst.expressions.NamedExpression]].map(_.name)
  }

  /** Partial (or, with merge = true, final) groups of one partition: keys coded to int64 -- an integer key column as it
    * is, anything else by its dense rank within the partition -- payload columns packed once, one tfs_session_aggregate. */
  def aggregatePartition(rows: Array[Row], schema: StructType, keyCols: Seq[String], all: ReduceBlockSchema,
                         graphDef: Array[Byte], merge: Boolean = false): Iterator[Row] = {
    if (rows.isEmpty) return Iterator.empty
    val arena = Arena.ofConfined()
    try {
      val keyIdx = keyCols.map(schema.fieldIndex)
      val table = mutable.LinkedHashMap.empty[Seq[Any], Long]
      val keys = Native.hostAlloc(arena, rows.length * 8L)
      rows.zipWithIndex.foreach { case (r, i) =>
        val k = keyIdx.map(r.get)
        keys.setAtIndex(ValueLayout.JAVA_LONG, i, table.getOrElseUpdate(k, table.size.toLong))
      }
      val decode = table.keys.toArray
      val g = Native.graphImport(arena, graphDef)
      val s = Native.sessionCreate(arena, g, device())
      try {
        val outNames = all.output.fieldNames
        val feeds = arena.allocate(Native.TENSOR, outNames.length)
        outNames.zipWithIndex.foreach { case (n, i) =>
          val c = schema.fieldIndex(n)                              // (merge: the partial column X is fed as X_input)
          val (code, dims, seg) = packColumn(arena, rows, c, if (merge) all.output(n) else schema(c))
          Native.setTensor(arena, feeds, i, n + "_input", code, dims, seg)
        }
        val cap = decode.length.toLong
        val fetches = arena.allocate(Native.TENSOR, outNames.length)
        val outs = outNames.zipWithIndex.map { case (n, i) =>
          val info = ColumnInformation(all.output(n)).stf.get
          val code = dtypeCode(info.dataType)
          val dims = (cap +: info.shape.dims).toArray
          val seg = Native.hostAlloc(arena, math.max(1L, dims.product * elemSize(code)))
          Native.setTensor(arena, fetches, i, n, code, dims, seg)
          (code, dims, seg)
        }
        val outKeys = Native.hostAlloc(arena, cap * 8L)
        val n = Native.sessionAggregate(arena, s, rows.length, keys, outNames.length, feeds, outNames.length, fetches, cap, outKeys)
        val cells = B200Convert.columns(outs, all.output, n.toInt)
        (0 until n.toInt).iterator.map { gi =>
          val k = decode(outKeys.getAtIndex(ValueLayout.JAVA_LONG, gi).toInt)
          Row.fromSeq(k ++ cells.map(_(gi)))
        }.toArray.iterator                                          // materialised before the arena closes
      } finally { Native.sessionFree(s); Native.graphFree(g) }
    } finally arena.close()
  }
}

/** Tensor bytes -> Rows: convertBack / getColumnFast0 / convertBuffer1-3 of the reference
  * (impl/TFDataOps.scala:128-204, impl/DataOps.scala:20-166, impl/datatypes.scala:204-256) reading from the pinned
  * result buffers instead of a Tensor. */
object B200Convert {
  def columns(outs: Seq[(Int, Array[Long], MemorySegment)], schema: StructType, nRows: Int): Seq[Int => Any] =
    outs.zip(schema.fields).map { case ((code, dims, seg), f) =>
      val cell = ColumnInformation(f).stf.get.shape.tail
      val buf = seg.asByteBuffer().order(ByteOrder.LITTLE_ENDIAN)
      SupportedOperations.opsFor(ColumnInformation(f).stf.get.dataType).cellReader(buf, cell, nRows)   // row index -> scalar / Seq / Seq[Seq]
    }

  def convertBack(outs: Seq[(Int, Array[Long], MemorySegment)], tfSchema: StructType, input: Array[Row], inputSchema: StructType,
                  appendInput: Boolean): Iterator[Row] = {
    val rowCounts = outs.map(_._2.headOption.getOrElse(1L)).distinct
    require(rowCounts.size == 1, s"Multiple number of rows detected. tf_struct=$tfSchema, tfSizes = $rowCounts")      // :150-157
    val n = rowCounts.head.toInt
    if (appendInput) require(n == input.length, s"Incompatible sizes detected: tf num rows = $n, input num rows = ${input.length}")
    val cols = columns(outs, tfSchema, n)
    val outSchema = if (appendInput) StructType(tfSchema ++ inputSchema.fields) else tfSchema
    (0 until n).iterator.map { i =>
      val tf = cols.map(_(i))
      val values = if (appendInput) tf ++ input(i).toSeq else tf
      new GenericRowWithSchema(values.toArray, outSchema): Row
    }.toArray.iterator
  }
}
