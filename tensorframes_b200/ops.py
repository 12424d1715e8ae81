"""Block executors over the C ABI: the counterpart of the reference's DebugRowOps.

Reference: src/main/scala/org/tensorframes/Operations.scala:20-135 (the operator
interface), src/main/scala/org/tensorframes/impl/DebugRowOps.scala (schema
validation :80-170 and :305-372, mapBlocks :290-393, reduceBlocks :503-526,
aggregate :547-592, performMap :766-803, performReduceBlock :872-895),
impl/TensorFlowOps.scala:101-141 (analyzeGraphTF), impl/TFDataOps.scala:27-59
and :128-169 (block packing / unpacking rules), impl/DataOps.scala:105-144
(inferPhysicalShape).

What stays on the host is what the reference does on the Spark driver: name and
type validation with the reference's error messages, output schema, per-partition
bookkeeping.  Every tensor operation goes through tfs_session_run /
tfs_session_aggregate on the device; nothing in this module computes on data.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import _native, dsl
from .frame import (ColumnInformation, GroupedData, Row, SparkTFColInfo, StructField, StructType, TensorFrame,
                    analyze as _analyze, explain as _explain, is_device_column, _to_python)
from .shape import ScalarType, Shape, ShapeDescription, Unknown


@dataclass(frozen=True)
class GraphNodeSummary:
    """impl/TensorFlowOps.scala:163-169"""
    isPlaceholder: bool
    isInput: bool
    isOutput: bool
    scalarType: ScalarType
    shape: Shape
    name: str


def _strip0(name: str) -> str:
    return name[:-2] if name.endswith(":0") else name


def analyze_graph(graph: _native.Graph, hints: ShapeDescription) -> List[GraphNodeSummary]:
    """Inputs = zero-input Placeholders; outputs = requested fetches; dtype and shape from the
    library's static analysis, the hinted shape winning (impl/TensorFlowOps.scala:101-141)."""
    try:
        summaries = graph.analyze({name: list(shape.dims) for name, shape in hints.out.items()}, list(hints.requestedFetches))
    except _native.NativeError as exc:
        if "Requested fetches are not in the graph" in str(exc):
            raise Exception(str(exc)) from None
        raise
    return [GraphNodeSummary(is_input, is_input, is_output, dtype, shape, name)
            for name, dtype, shape, _is_ph, is_input, is_output in summaries]


def infer_physical_shape(num_scalars: int, cell: Shape, expected_rows: Optional[int]) -> Tuple[int, Shape]:
    """Row count and cell shape of an output block from its element count (impl/DataOps.scala:105-144)."""
    if sum(1 for d in cell.dims if d == Unknown) > 1:
        raise ValueError(f"Shape has too many unkown values to perform inference: {cell}")
    h = []
    for d in cell.dims:
        if d < 0:
            break
        h.append(d)
    t = list(cell.dims[len(h) + 1:]) if len(h) < len(cell.dims) else []
    p = int(np.prod(h, dtype=np.int64)) * int(np.prod(t, dtype=np.int64))
    if expected_rows is None and cell.hasUnknown:
        raise ValueError(f"Cannot infer the final cell shape: {cell}")
    if expected_rows is not None and not cell.hasUnknown:
        if num_scalars != p * expected_rows:
            raise ValueError(f"Expected {p * expected_rows} elements in the final buffer,"
                             f"but got {num_scalars} instead. shape={cell}, numRows={expected_rows}")
        return expected_rows, cell
    if expected_rows is not None:
        inferred = num_scalars // (p * expected_rows) if p * expected_rows else 0
        shape = Shape(h + [inferred] + t)
        assert shape.numElements * expected_rows == num_scalars, \
            f"Incompatible values: shape={cell} numRows={expected_rows} numScalars={num_scalars}"
        return expected_rows, shape
    rows = num_scalars // p if p else 0
    assert (cell.numElements or 0) * rows == num_scalars, \
        f"Incompatible values: shape={cell} numRows={rows} numScalars={num_scalars}"
    return rows, cell


def _as_graph(fetches) -> Tuple[bytes, ShapeDescription]:
    nodes: List[dsl.Node] = []
    for f in fetches:
        nodes.extend(f if isinstance(f, (list, tuple)) else [f])
    g = dsl.build_graph(*nodes)
    return g, dsl.hints(nodes)


@dataclass(frozen=True)
class ReduceBlockSchema:
    """impl/DebugRowOps.scala:36-40"""
    mapInput: StructType
    mapTFCols: Tuple[int, ...]
    output: StructType
    reduceInput: StructType


def _widen_lead(f: StructField) -> StructField:
    ci = ColumnInformation.of(f)
    if ci.stf is not None and ci.stf.shape.numDims >= 1:
        s = ci.stf.shape.tail.prepend(Unknown)
        return ColumnInformation(f, SparkTFColInfo(s, ci.stf.dataType)).merged
    return f


def reduce_blocks_schema(schema: StructType, graph: _native.Graph, hints: ShapeDescription) -> ReduceBlockSchema:
    """The X / X_input naming contract and its checks (impl/DebugRowOps.scala:80-170)."""
    from dataclasses import replace
    summary = {s.name: s for s in analyze_graph(graph, hints)}
    fields_by_name = {f.name: f for f in schema.fields}
    field_names = ", ".join(sorted(fields_by_name))
    outputs = {k: v for k, v in summary.items() if v.isOutput}
    output_names = ", ".join(sorted(outputs))
    suffix = "_input"

    missing_cols = sorted(set(outputs) - set(fields_by_name))
    if missing_cols:
        raise Exception(f"Based on the TF graph, some inputs are missing: {', '.join(missing_cols)}. "
                        f"Dataframe columns: {field_names}; Outputs: {output_names}")
    inputs = {k: v for k, v in summary.items() if v.isInput}
    expected = {o + suffix for o in outputs}
    extra = sorted(set(inputs) - expected)
    if extra:
        raise Exception(f"Extra graph inputs have been found: {', '.join(extra)}. Dataframe columns: {field_names}")
    missing = sorted(expected - set(inputs))
    if missing:
        raise Exception(f"Some inputs are missing in the graph: {', '.join(missing)}. Dataframe columns: {field_names}")

    out_fields = []
    for f in schema.fields:  # keep the DataFrame's column order
        if f.name not in outputs:
            continue
        stf = ColumnInformation.of(f).stf
        if stf is None:
            raise Exception(f"Data column '{f.name}' has not been analyzed yet, cannot run TF on this dataframe")
        out = summary[f.name]
        if stf.dataType != out.scalarType:
            raise Exception(f"Output '{f.name}' has type {out.scalarType} but the column type is {stf.dataType}")
        cell = stf.shape.tail
        if not out.shape.checkMorePreciseThan(cell):
            raise Exception(f"Output '{f.name}' has shape {out.shape}, not compatible with the shape "
                            f"of field elements {cell}")
        in_shape = cell.prepend(Unknown)
        in_name = f.name + suffix
        inp = summary.get(in_name)
        if inp is None:
            raise Exception(f"The graph needs to have a placeholder input called {in_name}.")
        assert inp.isPlaceholder, f"Node {in_name} should be a placeholder"
        if not in_shape.checkMorePreciseThan(inp.shape):
            raise Exception(f"The data column '{f.name}' has shape {in_shape}, not compatible with shape"
                            f" {inp.shape} requested by the TF graph")
        if stf.dataType != inp.scalarType:
            raise Exception(f"The type of node '{inp.name}' ({stf.dataType}) is not compatible with the data"
                            f" type of the column ({inp.scalarType})")
        out_fields.append(ColumnInformation(f, SparkTFColInfo(in_shape, stf.dataType)).merged)
    output_schema = StructType(tuple(out_fields))
    map_input = StructType(tuple(_widen_lead(replace(f, name=f.name + suffix)) if f.name in outputs else f
                                 for f in schema.fields))
    reduce_input = StructType(tuple(_widen_lead(replace(f, name=f.name + suffix))
                                    for f in schema.fields if f.name in outputs))
    idx = tuple(i for i, f in enumerate(schema.fields) if f.name in outputs)
    return ReduceBlockSchema(map_input, idx, output_schema, reduce_input)


def _has_device_columns(df: TensorFrame) -> bool:
    return any(is_device_column(c) for p in df.partitions for c in p.values())


def _numel(arr) -> int:
    return int(arr.numel()) if is_device_column(arr) else int(arr.size)


def _as_feed(col):
    """A device-resident block as the C ABI wants it (pointer + dtype + shape); host blocks pass through."""
    if is_device_column(col):
        import torch
        torch.cuda.current_stream(col.device).synchronize()      # produced on torch's stream, read on the session's
        return _native.DeviceArray.from_torch(col.contiguous())
    return col


class B200Ops:
    """OperationsInterface (Operations.scala:20-135) executed on B200s through libtfsb200.

    `devices=[...]`: partition i of a frame runs on devices[i % len(devices)], every device driven by its own thread
    (the reference runs one task thread per partition, impl/DebugRowOps.scala:377-391).  mapBlocks needs no exchange;
    reduceBlocks and aggregate combine the per-device results through the library's communicators
    (tfs_session_reduce_combine / tfs_session_aggregate_combine: NCCL over NVLink), which stand in for RDD.reduce and
    the UDAF's shuffle (impl/DebugRowOps.scala:524-525, 573-576)."""

    def __init__(self, device: int = 0, devices: Optional[List[int]] = None):
        self.devices = list(devices) if devices else [device]
        self.device = self.devices[0]
        self._comms = None

    def __repr__(self) -> str:
        return f"B200Ops(devices={self.devices})"

    def _communicators(self):
        if self._comms is None:
            self._comms = _native.Comm.create_all(self.devices)
        return self._comms

    def _per_device(self, work):
        """Runs work(slot) for every device slot on its own thread; returns the results in slot order."""
        n = len(self.devices)
        if n == 1:
            return [work(0)]
        import threading
        results, errors = [None] * n, []

        def runner(slot):
            try:
                results[slot] = work(slot)
            except BaseException as exc:      # surfaced in the caller's thread
                errors.append(exc)

        threads = [threading.Thread(target=runner, args=(slot,)) for slot in range(n)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results

    # ---- helpers -------------------------------------------------------------------
    @staticmethod
    def _graph_args(args) -> Tuple[bytes, ShapeDescription]:
        """(graph_bytes, hints) given either DSL nodes or an explicit (bytes, ShapeDescription)."""
        if len(args) == 2 and isinstance(args[0], (bytes, bytearray)) and isinstance(args[1], ShapeDescription):
            return bytes(args[0]), args[1]
        return _as_graph(args)

    @staticmethod
    def _dense_block(df: TensorFrame, part: Dict[str, object], name: str, n_rows: int) -> np.ndarray:
        """The [n_rows, cell...] block of a column, with the checks of TFDataOps.convert
        (impl/TFDataOps.scala:36-52) and TensorConverter (impl/datatypes.scala:80-83)."""
        f = df.schema[name]
        ci = ColumnInformation.of(f).stf
        if ci is None:
            raise Exception(f"Could not column information for column {f}")
        if ci.shape.numDims < 1:
            raise Exception(f"Column {f} found to be scalar, but its dimensions should be >= 1")
        lead = ci.shape.dims[0]
        if lead != Unknown and lead != n_rows:
            raise Exception(f"Lead dimension for column {f} (found to be {lead}) is not compatible with a block of "
                            f"size {n_rows}. Expected block structure: {df.schema}, meta info = {ci}")
        cell = ci.shape.tail
        if cell.hasUnknown:
            raise AssertionError(f"Shape {cell} has unknown values.")
        col = part[name]
        if not isinstance(col, np.ndarray) and not is_device_column(col):
            raise Exception(f"Column '{name}' holds ragged or non-numeric cells and cannot be packed as a block")
        if tuple(col.shape[1:]) != cell.dims:
            raise Exception(f"Column '{name}' cells have shape {Shape(col.shape[1:])} but the metadata says {cell}")
        if cell.numDims > 2:
            raise Exception(f"Higher order dimension {cell.numDims} from shape {cell} is not supported")
        return col

    # ---- mapBlocks (impl/DebugRowOps.scala:290-393, 766-803) ---------------------------
    def mapBlocks(self, df: TensorFrame, *fetches) -> TensorFrame:
        g, h = self._graph_args(fetches)
        return self._map_blocks(df, g, h, append_input=True)

    def mapBlocksTrimmed(self, df: TensorFrame, *fetches) -> TensorFrame:
        g, h = self._graph_args(fetches)
        return self._map_blocks(df, g, h, append_input=False)

    def _map_blocks(self, df: TensorFrame, graph_bytes: bytes, hints: ShapeDescription, append_input: bool) -> TensorFrame:
        with _native.Graph(graph_bytes) as graph:
            summary = {s.name: s for s in analyze_graph(graph, hints)}
            inputs = {k: v for k, v in summary.items() if v.isInput}
            outputs = {k: v for k, v in summary.items() if v.isOutput}
            fields_by_name = {f.name: f for f in df.schema.fields}
            cols = ", ".join(df.schema.fieldNames)
            for inp in inputs.values():
                f = fields_by_name.get(inp.name)
                if f is None:
                    raise Exception(f"Graph input {inp.name} found, but no column to match it. Dataframe columns: {cols}")
                stf = ColumnInformation.of(f).stf
                if stf is None:
                    raise Exception(f"Data column {f.name} has not been analyzed yet, cannot run TF on this dataframe")
                if stf.dataType != inp.scalarType:
                    raise Exception(f"The type of node '{inp.name}' ({stf.dataType}) is not compatible with the data type "
                                    f"of the column ({inp.scalarType})")
                if not stf.shape.checkMorePreciseThan(inp.shape):
                    raise Exception(f"The data column '{f.name}' has shape {stf.shape} (not compatible) with shape"
                                    f" {inp.shape} requested by the TF graph")
                if not inp.isPlaceholder:
                    raise Exception(f"Invalid type for input node {inp.name}. It has to be a placeholder")
            out_fields = []
            for out in sorted(outputs.values(), key=lambda o: o.name):
                if out.name in fields_by_name:
                    raise Exception(f"TF graph has an output node called '{out.name}', but this column already exists. "
                                    f"Input columns: {cols}")
                out_fields.append(ColumnInformation.structField(out.name, out.scalarType, out.shape))
            tf_schema = StructType(tuple(out_fields))
            out_schema = StructType(tf_schema.fields + (df.schema.fields if append_input else ()))
            fetch_names = [f.name for f in tf_schema.fields]
            input_names = list(inputs.keys())

            def map_partition(session, device, part):
                n = df._part_len(part)
                feeds = {name: self._dense_block(df, part, name, n) for name in input_names}
                # blocks that live on the device stay there, and so do the columns computed from them
                on_device = any(is_device_column(v) for v in feeds.values())
                outs = session.run_torch(feeds, fetch_names, device) if on_device else session.run(feeds, fetch_names)
                new_part: Dict[str, object] = {}
                rows_seen = []
                for f, arr in zip(tf_schema.fields, outs):
                    info = ColumnInformation.of(f).stf
                    n_rows, cell = infer_physical_shape(_numel(arr), info.shape.tail, n if append_input else None)
                    rows_seen.append(n_rows)
                    if cell.numDims > 2:
                        raise ValueError(f"Operations for tensors of order {cell.numDims + 1} are not supported")
                    new_part[f.name] = arr.reshape((n_rows,) + cell.dims)
                if len(set(rows_seen)) != 1:
                    raise Exception(f"Multiple number of rows detected. tf_struct={tf_schema}, tfSizes = {rows_seen}")
                assert (not append_input) or rows_seen[0] == n, \
                    f"Incompatible sizes detected: appendInput={append_input}, tf num rows = {rows_seen[0]}, " \
                    f"input num rows = {n}"
                if append_input:
                    for name in df.schema.fieldNames:
                        new_part[name] = part[name]
                return new_part

            # empty partitions never reach the runtime (:380-390, :777-779)
            live = [(i, part) for i, part in enumerate(df.partitions) if df._part_len(part) > 0]
            done: Dict[int, Dict[str, object]] = {}

            def work(slot):
                mine = [(i, part) for k, (i, part) in enumerate(live) if k % len(self.devices) == slot]
                if not mine:
                    return
                with _native.Session(graph, self.devices[slot]) as session:
                    for i, part in mine:
                        done[i] = map_partition(session, self.devices[slot], part)

            self._per_device(work)
            parts = [done[i] for i, _ in live]
        return TensorFrame(out_schema, parts)

    # ---- reduceBlocks (impl/DebugRowOps.scala:503-526, 872-895) -----------------------
    def reduceBlocks(self, df: TensorFrame, *fetches) -> Row:
        g, h = self._graph_args(fetches)
        with _native.Graph(g) as graph:
            schemas = reduce_blocks_schema(df.schema, graph, h)
            out_names = [f.name for f in schemas.output.fields]
            live = [part for part in df.partitions if df._part_len(part) > 0]
            if not live:
                raise ValueError("empty collection")  # RDD.reduce on an empty RDD (:524)
            n_dev = min(len(self.devices), len(live))

            def local_reduce(session, mine):
                """The partial row of one device: its partitions one by one, then -- the reference merges partition
                results pairwise by running the same graph on 2-row blocks (:741-750) -- one run over the stacked
                [P, cell...] block, which is the same merge for the associative reducers the contract allows."""
                partials = []
                for part in mine:
                    n = df._part_len(part)
                    feeds = {name + "_input": _as_feed(self._dense_block(df, part, name, n)) for name in out_names}
                    partials.append(session.run(feeds, out_names))      # the partial row is small: it comes to the host
                if len(partials) == 1:
                    return partials[0]
                feeds = {name + "_input": np.stack([p[i] for p in partials]) for i, name in enumerate(out_names)}
                return session.run(feeds, out_names)

            if n_dev == 1:
                with _native.Session(graph, self.device) as session:
                    result = local_reduce(session, live)
            else:
                import torch
                comms = self._communicators() if n_dev == len(self.devices) else _native.Comm.create_all(self.devices[:n_dev])
                DA = _native.DeviceArray.from_torch

                def work(slot):
                    if slot >= n_dev:
                        return None
                    dev = self.devices[slot]
                    with _native.Session(graph, dev) as session:
                        rows = local_reduce(session, [p for k, p in enumerate(live) if k % n_dev == slot])
                        part_t = {n_: torch.from_numpy(np.ascontiguousarray(r)).to(f"cuda:{dev}") for n_, r in zip(out_names, rows)}
                        merged_t = {n_: torch.empty_like(t) for n_, t in part_t.items()}
                        torch.cuda.synchronize(dev)
                        session.reduce_combine(comms[slot], {k: DA(v) for k, v in part_t.items()}, {k: DA(v) for k, v in merged_t.items()})
                        session.sync()
                        return [merged_t[n_].cpu().numpy() for n_ in out_names]

                result = self._per_device(work)[0]
                if comms is not self._comms:
                    for c in comms:
                        c.close()
        for f, arr in zip(schemas.output.fields, result):
            cell = ColumnInformation.of(f).stf.shape.tail
            if cell.numElements is not None and int(arr.size) != cell.numElements:
                raise ValueError(f"Expected {cell.numElements} elements in the final buffer,but got {arr.size} instead. "
                                 f"shape={cell}, numRows=1")
        return Row.of(out_names, [_to_python(a) for a in result])

    # ---- aggregate (impl/DebugRowOps.scala:547-592, TensorFlowUDAF :601-695) ------------
    @staticmethod
    def _key_ids(df: TensorFrame, key_names: List[str], parts: List[Dict[str, object]]):
        """The key columns of the given partitions as one int64 id per row, plus the way back from ids to key values.
        A single integer key column is used as it is, bit for bit (no host pass over its values); anything else is
        coded column by column with numpy (`np.unique(..., return_inverse=True)`) and the codes are combined --
        the grouping itself is Spark's job in the reference."""
        cols = {k: [p[k] for p in parts] for k in key_names}
        if len(key_names) == 1 and all(isinstance(c, np.ndarray) and c.dtype.kind == "i" for c in cols[key_names[0]]):
            ids = [c.astype(np.int64, copy=False) for c in cols[key_names[0]]]
            return ids, None, None
        codes, tables = [], []
        for k in key_names:
            flat = np.concatenate([np.asarray(c) if isinstance(c, np.ndarray) else np.asarray(list(c), dtype=object) for c in cols[k]])
            if flat.dtype == object:
                flat = flat.astype(str) if all(isinstance(v, str) for v in flat[:1]) else flat
            values, inv = np.unique(flat, return_inverse=True)
            codes.append(inv.astype(np.int64))
            tables.append(values)
        radix = [len(t) for t in tables]
        span = 1
        for r in radix:
            span *= max(r, 1)
        if span >= 2**62:
            raise Exception("aggregate: the key columns have too many distinct value combinations to be coded in 64 bits")
        ids_flat = np.zeros(len(codes[0]), dtype=np.int64)
        for c, r in zip(codes, radix):
            ids_flat = ids_flat * r + c
        sizes = np.cumsum([0] + [df._part_len(p) for p in parts])
        ids = [ids_flat[sizes[i]:sizes[i + 1]] for i in range(len(parts))]
        return ids, tables, radix

    def aggregate(self, grouped: GroupedData, *fetches) -> TensorFrame:
        g, h = self._graph_args(fetches)
        # (grouping builds key ids on the host: device-resident frames come back for this operator)
        df = grouped.df.to_host() if _has_device_columns(grouped.df) else grouped.df
        with _native.Graph(g) as graph:
            schemas = reduce_blocks_schema(df.schema, graph, h)
            out_names = [f.name for f in schemas.output.fields]
            key_names = list(grouped.keys)
            out_schema = StructType(tuple(df.schema[k] for k in key_names) + schemas.output.fields)
            live = [part for part in df.partitions if df._part_len(part) > 0]
            if not live or not key_names:
                return TensorFrame(out_schema, [])
            ids, tables, radix = self._key_ids(df, key_names, live)
            # a bound on the number of groups: every row its own group, or every combination of coded key values
            total = int(sum(len(i) for i in ids))
            bound = total
            if radix is not None:
                combos = 1
                for r in radix:
                    combos *= max(r, 1)
                bound = min(total, combos)
            n_dev = min(len(self.devices), len(live))

            def local_aggregate(session, slot, capacity):
                mine = [k for k in range(len(live)) if k % n_dev == slot]
                keys = np.concatenate([ids[k] for k in mine])
                feeds = {}
                for name in out_names:
                    blocks = [self._dense_block(df, live[k], name, df._part_len(live[k])) for k in mine]
                    feeds[name + "_input"] = blocks[0] if len(blocks) == 1 else np.concatenate(blocks)
                cap = min(len(keys), capacity)
                while True:           # the capacity is a guess for raw integer keys: grow it when the block has more groups
                    try:
                        return session.aggregate(keys, feeds, out_names, cap)
                    except _native.NativeError as exc:
                        if "more than max_groups" not in str(exc) or cap >= len(keys):
                            raise
                        cap = min(len(keys), cap * 8)

            first_guess = bound if radix is not None else min(bound, 1 << 20)
            if n_dev == 1:
                with _native.Session(graph, self.device) as session:
                    shards = [local_aggregate(session, 0, first_guess)]
            else:
                import torch
                comms = self._communicators() if n_dev == len(self.devices) else _native.Comm.create_all(self.devices[:n_dev])
                DA = _native.DeviceArray.from_torch
                tdt = {"float64": torch.float64, "float32": torch.float32, "int32": torch.int32, "int64": torch.int64}

                def work(slot):
                    if slot >= n_dev:
                        return None
                    dev = self.devices[slot]
                    with _native.Session(graph, dev) as session:
                        keys_out, outs = local_aggregate(session, slot, first_guess)
                        # the partial groups of this device go to the devices that own their keys and are merged there
                        d_keys = torch.from_numpy(keys_out).to(f"cuda:{dev}")
                        d_outs = {n_: torch.from_numpy(np.ascontiguousarray(o)).to(f"cuda:{dev}") for n_, o in zip(out_names, outs)}
                        cap = max(1, bound)
                        m_keys = torch.empty(cap, dtype=torch.int64, device=f"cuda:{dev}")
                        m_outs = {n_: torch.empty((cap,) + tuple(o.shape[1:]), dtype=tdt[o.dtype.name], device=f"cuda:{dev}")
                                  for n_, o in zip(out_names, outs)}
                        torch.cuda.synchronize(dev)
                        owned = session.aggregate_combine(comms[slot], len(keys_out), DA(d_keys), {k: DA(v) for k, v in d_outs.items()},
                                                          {k: DA(v) for k, v in m_outs.items()}, DA(m_keys))
                        return m_keys[:owned].cpu().numpy(), [m_outs[n_][:owned].cpu().numpy() for n_ in out_names]

                shards = [r for r in self._per_device(work) if r is not None]
                if comms is not self._comms:
                    for c in comms:
                        c.close()
        # one output partition per device: the result stays sharded by key owner, as the reference's stays sharded by
        # its shuffle (row order is unspecified there, T/TensorFlossTestSparkContext.scala:37-41)
        parts_out = []
        for keys_out, outs in shards:
            if len(keys_out) == 0:
                continue
            part: Dict[str, object] = {}
            if tables is None:
                kf = df.schema[key_names[0]]
                part[key_names[0]] = keys_out.astype(kf.dataType.np_dtype) if isinstance(kf.dataType, ScalarType) else keys_out
            else:
                rest = keys_out.copy()
                decoded = {}
                for k, t, r in reversed(list(zip(key_names, tables, radix))):
                    decoded[k] = t[rest % r]
                    rest //= r
                for k in key_names:
                    kf = df.schema[k]
                    vals = decoded[k]
                    part[k] = (np.asarray(vals, dtype=kf.dataType.np_dtype) if isinstance(kf.dataType, ScalarType)
                               else [v.item() if hasattr(v, "item") else v for v in vals])
            for f, arr in zip(schemas.output.fields, outs):
                part[f.name] = arr
            parts_out.append(part)
        return TensorFrame(out_schema, parts_out)

    # ---- mapRows (impl/DebugRowOps.scala:396-477, 819-857) ------------------------------
    def mapRows(self, df: TensorFrame, *fetches, feed_dict: Optional[Dict[str, str]] = None) -> TensorFrame:
        """Row-wise map: placeholders have the shape of ONE cell and are bound to columns through
        `feed_dict` (placeholder -> column; default: the column of the same name).  Rows whose cells differ
        in length (first cell dimension unknown) are allowed.  The device evaluates all rows of a partition
        that share a cell shape in one call (tfs_session_run_rows) instead of one Session.run per row."""
        if _has_device_columns(df):          # row operators pack on the host
            df = df.to_host()
        g, h = self._graph_args(fetches)
        feed_dict = dict(feed_dict or {})
        with _native.Graph(g) as graph:
            summary = {s.name: s for s in analyze_graph(graph, h)}
            inputs = {k: v for k, v in summary.items() if v.isInput}
            outputs = {k: v for k, v in summary.items() if v.isOutput}
            fields_by_name = {f.name: f for f in df.schema.fields}
            cols = ", ".join(df.schema.fieldNames)
            hint_inputs = dict(h.inputs)
            hint_inputs.update(feed_dict)
            binding: Dict[str, str] = {}
            for inp in inputs.values():
                fname = hint_inputs.get(inp.name)
                if fname is None:
                    raise Exception(f"The graph placeholder {inp.name} was not given a corresponding dataframe field "
                                    f"name as input:hints: {hint_inputs}")
                f = fields_by_name.get(fname)
                if f is None:
                    raise Exception(f"Graph input {inp.name} found, but no column to match it. Dataframe columns: {cols}")
                stf = ColumnInformation.of(f).stf
                if stf is None:
                    raise Exception(f"Data column {f.name} has not been analyzed yet, cannot run TF on this dataframe")
                if stf.dataType != inp.scalarType:
                    raise Exception(f"The type of node '{inp.name}' ({stf.dataType}) is not compatible with the data type "
                                    f"of the column ({inp.scalarType})")
                cell = stf.shape.tail
                if not cell.checkMorePreciseThan(inp.shape):
                    raise Exception(f"The data column '{f.name}' has shape {stf.shape} (not compatible) with shape"
                                    f" {inp.shape} requested by the TF graph")
                binding[inp.name] = fname
            out_fields = []
            for out in sorted(outputs.values(), key=lambda o: o.name):
                if out.name in fields_by_name:
                    raise Exception(f"TF graph has an output node called '{out.name}', but this column already exists. "
                                    f"Input columns: {cols}")
                out_fields.append(ColumnInformation.structField(out.name, out.scalarType, out.shape.prepend(Unknown)))
            tf_schema = StructType(tuple(out_fields))
            out_schema = StructType(tf_schema.fields + df.schema.fields)
            fetch_names = [f.name for f in tf_schema.fields]
            parts = []
            with _native.Session(graph, self.device) as session:
                for part in df.partitions:
                    n = df._part_len(part)
                    if n == 0:
                        continue
                    # group the rows by the shapes of their fed cells (one group when the columns are dense)
                    groups: Dict[tuple, List[int]] = {}
                    dense = all(isinstance(part[c], np.ndarray) for c in binding.values())
                    if dense or not binding:
                        groups[()] = list(range(n))
                    else:
                        for i in range(n):
                            key = tuple(np.shape(part[c][i]) for c in binding.values())
                            groups.setdefault(key, []).append(i)
                    per_row: Dict[str, List[object]] = {name: [None] * n for name in fetch_names}
                    for idxs in groups.values():
                        feeds = {}
                        for ph, c in binding.items():
                            col = part[c]
                            st = ColumnInformation.of(df.schema[c]).stf.dataType
                            feeds[ph] = (col if isinstance(col, np.ndarray) and len(idxs) == n
                                         else np.asarray([col[i] for i in idxs], dtype=st.np_dtype))
                        if feeds:
                            outs = session.run(feeds, fetch_names, rows=True)
                        else:   # no placeholder at all: every row gets the graph's constant result
                            outs = [np.broadcast_to(o, (len(idxs),) + o.shape) for o in session.run({}, fetch_names)]
                        for name, arr in zip(fetch_names, outs):
                            for j, i in enumerate(idxs):
                                per_row[name][i] = arr[j]
                    new_part: Dict[str, object] = {}
                    for f in tf_schema.fields:
                        cells = per_row[f.name]
                        shapes = {np.shape(c) for c in cells}
                        new_part[f.name] = np.stack(cells) if len(shapes) == 1 else [np.asarray(c).tolist() for c in cells]
                    for name in df.schema.fieldNames:
                        new_part[name] = part[name]
                    parts.append(new_part)
        return TensorFrame(out_schema, parts)

    # ---- reduceRows (impl/DebugRowOps.scala:172-262, 479-501, 930-969) ----------------------
    def reduceRows(self, df: TensorFrame, *fetches) -> Row:
        """Pairwise reduction: for every column X the graph reads placeholders X_1 and X_2 (one cell each)
        and produces X.  The reference folds the rows one by one through a session; here each round
        combines the first half of the remaining rows with the second half in one row-mode call, so n rows
        take about log2(n) device calls (the order of combination is unspecified in the reference too)."""
        if _has_device_columns(df):          # row operators pack on the host
            df = df.to_host()
        g, h = self._graph_args(fetches)
        with _native.Graph(g) as graph:
            summary = {s.name: s for s in analyze_graph(graph, h)}
            fields_by_name = {f.name: f for f in df.schema.fields}
            field_names = ", ".join(sorted(fields_by_name))
            outputs = {k: v for k, v in summary.items() if v.isOutput}
            output_names = ", ".join(sorted(outputs))
            extra = sorted(set(outputs) - set(fields_by_name))
            if extra:
                raise Exception(f"Some extra outputs were found in the reducer: {', '.join(extra)}. "
                                f"Dataframe columns: {field_names}; Outputs: {output_names}")
            missing = sorted(set(fields_by_name) - set(outputs))
            if missing:
                raise Exception(f"Some outputs are missing in the reducer: {', '.join(missing)}. "
                                f"Dataframe columns: {field_names}; Outputs: {output_names}")
            inputs = {k: v for k, v in summary.items() if v.isInput}
            expected = {n + s for n in fields_by_name for s in ("_1", "_2")}
            extra = sorted(set(inputs) - expected)
            if extra:
                raise Exception(f"Extra graph inputs have been found: {', '.join(extra)}. Dataframe columns: {field_names}")
            missing = sorted(expected - set(inputs))
            if missing:
                raise Exception(f"Some inputs are missing in th graph: {', '.join(missing)}. Dataframe columns: {field_names}")
            names = df.schema.fieldNames
            for f in df.schema.fields:
                stf = ColumnInformation.of(f).stf
                if stf is None:
                    raise Exception(f"Data column '{f.name}' has not been analyzed yet, cannot run TF on this dataframe")
                out = summary[f.name]
                if stf.dataType != out.scalarType:
                    raise Exception(f"Output '{f.name}' has type {out.scalarType} but the column type is {stf.dataType}")
                cell = stf.shape.tail
                if not out.shape.checkMorePreciseThan(cell):
                    raise Exception(f"Output '{f.name}' has shape {out.shape}, not compatible with the shapesof field "
                                    f"elements {cell}")
                for suffix in ("_1", "_2"):
                    inp = summary[f.name + suffix]
                    if not cell.checkMorePreciseThan(inp.shape):
                        raise Exception(f"The data column '{f.name}' has shape {stf.shape} (not compatible) with shape"
                                        f" {inp.shape} requested by the TF graph")
                    if stf.dataType != inp.scalarType:
                        raise Exception(f"The type of node '{inp.name}' ({stf.dataType}) is not compatible with the data"
                                        f" type of the column ({inp.scalarType})")
            blocks: Dict[str, List[np.ndarray]] = {n: [] for n in names}
            for part in df.partitions:
                n = df._part_len(part)
                if n == 0:
                    continue
                for name in names:
                    col = part[name]
                    if not isinstance(col, np.ndarray):
                        raise Exception(f"Column '{name}' holds ragged cells and cannot be reduced pairwise")
                    blocks[name].append(col)
            if not blocks[names[0]]:
                raise ValueError("empty collection")
            cur = {name: np.concatenate(blocks[name]) for name in names}
            with _native.Session(graph, self.device) as session:
                n = cur[names[0]].shape[0]
                while n > 1:
                    half = n // 2
                    feeds = {}
                    for name in names:
                        feeds[name + "_1"] = np.ascontiguousarray(cur[name][:half])
                        feeds[name + "_2"] = np.ascontiguousarray(cur[name][half:2 * half])
                    outs = session.run(feeds, names, rows=True)
                    for name, o in zip(names, outs):
                        cur[name] = np.concatenate([o, cur[name][2 * half:]]) if n % 2 else o
                    n = half + (n % 2)
        return Row.of(names, [_to_python(cur[name][0]) for name in names])

    # ---- analysis --------------------------------------------------------------------
    def analyze(self, df: TensorFrame) -> TensorFrame:
        return _analyze(df)

    def explain(self, df: TensorFrame) -> str:
        return _explain(df)


_default: Optional[B200Ops] = None


def default_ops() -> B200Ops:
    global _default
    if _default is None:
        _default = B200Ops(0)
    return _default
