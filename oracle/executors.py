"""Row-by-row restatement of the reference's block executors (test infrastructure only).

Follows /root/reference/src/main/scala/org/tensorframes/ line by line where it matters:

  convert            impl/TFDataOps.scala:27-59 + impl/DataOps.scala:63-81 (convertFast0) +
                     impl/datatypes.scala:85-127 (append by cell rank)
  convert_back       impl/TFDataOps.scala:128-169 + impl/DataOps.scala:20-61, 105-166
  perform_map        impl/DebugRowOps.scala:766-803
  map_blocks         impl/DebugRowOps.scala:305-393 (outputs sorted by name, TF columns first)
  perform_reduce_block  impl/DebugRowOps.scala:872-895
  reduce_blocks      impl/DebugRowOps.scala:503-526 (+ reducePairBlock :741-750)
  TensorFlowUDAF     impl/DebugRowOps.scala:601-695 (buffer of 10 rows, compaction)
  aggregate          impl/DebugRowOps.scala:547-592
  map_rows           impl/DebugRowOps.scala:396-477, 819-857 (one Session.run per row; a cell whose first
                     dimension is unknown takes the row's own length, impl/TFDataOps.scala:86-105)
  reduce_rows        impl/DebugRowOps.scala:479-501, 930-969 (pairwise fold X_1, X_2 -> X)
  analyze            ExperimentalOperations.scala:68-157

Rows are python tuples, partitions are lists of rows, and the arithmetic is
oracle.tf_ops.evaluate.  Pure-python loops: meant for small cases.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .graphdef import DT
from .tf_ops import evaluate

UNKNOWN = -1


@dataclass
class Col:
    name: str
    dtype: Optional[int]              # TF DataType code, None for a non-tensor column (e.g. string keys)
    shape: Optional[Tuple[int, ...]]  # block shape [rows, cell...], -1 = unknown


# ---- analyze ---------------------------------------------------------------------------
def _merge(s1, s2):
    if s1 is None or s2 is None or len(s1) != len(s2):
        return None
    return tuple(a if a == b else UNKNOWN for a, b in zip(s1, s2))


def analyze_data(x):
    if x is None:
        return None
    if isinstance(x, (list, tuple)):
        if not x:
            return (0,)
        m = analyze_data(x[0])
        for e in x[1:]:
            m = _merge(m, analyze_data(e))
        return None if m is None else (len(x),) + m
    if isinstance(x, (int, float)) and not isinstance(x, bool):
        return ()
    return None


def analyze(partitions: Sequence[Sequence[tuple]], n_cols: int) -> List[Optional[Tuple[int, ...]]]:
    """Block shape per column, merged over rows then over non-empty partitions."""
    infos = []
    for rows in partitions:
        if not rows:
            continue
        acc = [analyze_data(v) for v in rows[0]]
        for r in rows[1:]:
            acc = [_merge(a, analyze_data(v)) for a, v in zip(acc, r)]
        infos.append([None if a is None else (len(rows),) + a for a in acc])
    if not infos:
        return [None] * n_cols
    out = infos[0]
    for i in infos[1:]:
        out = [_merge(a, b) for a, b in zip(out, i)]
    return out


# ---- pack / unpack ---------------------------------------------------------------------
def convert(rows: Sequence[tuple], cols: Sequence[Col], requested: Sequence[int]) -> Dict[str, np.ndarray]:
    out = {}
    for ci in requested:
        c = cols[ci]
        if c.shape is None or c.dtype is None:
            raise Exception(f"Could not column information for column {c}")
        if len(c.shape) < 1:
            raise Exception(f"Column {c} found to be scalar, but its dimensions should be >= 1")
        lead = c.shape[0]
        if lead != UNKNOWN and lead != len(rows):
            raise Exception(f"Lead dimension for column {c} (found to be {lead}) is not compatible with a block of "
                            f"size {len(rows)}.")
        cell = c.shape[1:]
        assert UNKNOWN not in cell, f"Shape {cell} has unknown values."
        buf: List[Any] = []
        for r in rows:                       # HOT LOOP 1 of the reference
            v = r[ci]
            if len(cell) == 0:
                buf.append(v)
            elif len(cell) == 1:
                buf.extend(v)
            elif len(cell) == 2:
                for inner in v:
                    buf.extend(inner)
            else:
                raise Exception(f"Higher order dimension {len(cell)} from shape {cell} is not supported")
        n_expected = len(rows) * int(np.prod(cell, dtype=np.int64))
        assert len(buf) == n_expected, (len(buf), n_expected)
        out[c.name] = np.asarray(buf, dtype=DT[c.dtype]).reshape((len(rows),) + tuple(cell))
    return out


def infer_physical_shape(num_scalars: int, cell: Tuple[int, ...], expected_rows: Optional[int]):
    if sum(1 for d in cell if d == UNKNOWN) > 1:
        raise ValueError(f"Shape has too many unkown values to perform inference: {cell}")
    h = []
    for d in cell:
        if d < 0:
            break
        h.append(d)
    t = list(cell[len(h) + 1:]) if len(h) < len(cell) else []
    p = int(np.prod(h, dtype=np.int64) * np.prod(t, dtype=np.int64))
    has_unknown = UNKNOWN in cell
    if expected_rows is None and has_unknown:
        raise ValueError(f"Cannot infer the final cell shape: {cell}")
    if expected_rows is not None and not has_unknown:
        if num_scalars != p * expected_rows:
            raise ValueError(f"Expected {p * expected_rows} elements in the final buffer,but got {num_scalars} instead.")
        return expected_rows, tuple(cell)
    if expected_rows is not None:
        inferred = num_scalars // (p * expected_rows)
        shape = tuple(h) + (inferred,) + tuple(t)
        assert int(np.prod(shape, dtype=np.int64)) * expected_rows == num_scalars
        return expected_rows, shape
    rows = num_scalars // p
    assert int(np.prod(cell, dtype=np.int64)) * rows == num_scalars
    return rows, tuple(cell)


def _cells(t: np.ndarray, cell: Tuple[int, ...], expected_rows: Optional[int]) -> List[Any]:
    n_rows, shape = infer_physical_shape(int(t.size), cell, expected_rows)
    if len(shape) > 2:
        raise ValueError(f"Operations for tensors of order {len(shape) + 1} are not supported")
    arr = t.reshape((n_rows,) + tuple(shape))
    return [arr[i].tolist() if shape else arr[i].item() for i in range(n_rows)]   # HOT LOOP 2


def convert_back(tensors: Sequence[np.ndarray], tf_cols: Sequence[Col], rows: Sequence[tuple],
                 append_input: bool) -> List[tuple]:
    per_col = []
    for t, c in zip(tensors, tf_cols):
        per_col.append(_cells(t, tuple(c.shape[1:]), len(rows) if append_input else None))
    sizes = {len(x) for x in per_col}
    if not sizes:
        raise Exception("Output cannot be empty.")
    if len(sizes) != 1:
        raise Exception(f"Multiple number of rows detected. tfSizes = {[len(x) for x in per_col]}")
    n = sizes.pop()
    assert (not append_input) or n == len(rows)
    out = []
    for i in range(n):                       # HOT LOOP 3 of the reference
        tf_part = tuple(col[i] for col in per_col)
        out.append(tf_part + (tuple(rows[i]) if append_input else ()))
    return out


# ---- map --------------------------------------------------------------------------------
def perform_map(rows, cols, requested, graph: bytes, tf_cols: Sequence[Col], append_input: bool) -> List[tuple]:
    if not rows:
        return []
    feeds = convert(rows, cols, requested)
    outs = evaluate(graph, feeds, [c.name for c in tf_cols])
    return convert_back(outs, tf_cols, rows, append_input)


def map_blocks(partitions, cols: Sequence[Col], graph: bytes, inputs: Sequence[str],
               fetch_info: Dict[str, Tuple[int, Tuple[int, ...]]], append_input: bool = True):
    """fetch_info: fetch name -> (dtype code, block shape).  Returns (column names, partitions of rows)."""
    by_name = {c.name: i for i, c in enumerate(cols)}
    for name in inputs:
        if name not in by_name:
            raise Exception(f"Graph input {name} found, but no column to match it.")
    tf_cols = [Col(n, fetch_info[n][0], tuple(fetch_info[n][1])) for n in sorted(fetch_info)]
    for c in tf_cols:
        if c.name in by_name:
            raise Exception(f"TF graph has an output node called '{c.name}', but this column already exists.")
    requested = [by_name[n] for n in inputs]
    names = [c.name for c in tf_cols] + ([c.name for c in cols] if append_input else [])
    return names, [perform_map(rows, cols, requested, graph, tf_cols, append_input) for rows in partitions]


# ---- reduce ------------------------------------------------------------------------------
def perform_reduce_block(rows, in_cols: Sequence[Col], requested: Sequence[int], out_cols: Sequence[Col],
                         graph: bytes) -> tuple:
    feeds = convert(rows, in_cols, requested)
    outs = evaluate(graph, feeds, [c.name for c in out_cols])
    back = convert_back(outs, out_cols, [()], append_input=False)
    assert len(back) == 1, back
    return back[0]


def _reduce_cols(cols: Sequence[Col], outputs: Sequence[str]):
    idx = [i for i, c in enumerate(cols) if c.name in outputs]
    in_cols = [Col(c.name + "_input", c.dtype, (UNKNOWN,) + tuple(c.shape[1:])) if c.name in outputs else c
               for c in cols]
    out_cols = [Col(cols[i].name, cols[i].dtype, (UNKNOWN,) + tuple(cols[i].shape[1:])) for i in idx]
    red_cols = [in_cols[i] for i in idx]
    return idx, in_cols, out_cols, red_cols


def reduce_blocks(partitions, cols: Sequence[Col], graph: bytes, outputs: Sequence[str]) -> tuple:
    idx, in_cols, out_cols, red_cols = _reduce_cols(cols, outputs)
    partials = []
    for rows in partitions:
        if not rows:
            continue
        if len(rows) == 1:
            # the reference passes a lone row through untouched (:515-516); only the reduced
            # columns are kept here, which is what its later merge step reads
            partials.append(tuple(rows[0][i] for i in idx))
        else:
            partials.append(perform_reduce_block(rows, in_cols, idx, out_cols, graph))
    if not partials:
        raise ValueError("empty collection")
    acc = partials[0]
    for nxt in partials[1:]:   # RDD.reduce with reducePairBlock: the same graph on a 2-row block
        acc = perform_reduce_block([acc, nxt], red_cols, list(range(len(red_cols))), out_cols, graph)
    return acc


# ---- aggregate -----------------------------------------------------------------------------
class TensorFlowUDAF:
    """Buffer of at most bufferSize+1 rows per key, compacted by a block reduce."""

    def __init__(self, red_cols: Sequence[Col], out_cols: Sequence[Col], graph: bytes, buffer_size: int = 10):
        self.red_cols, self.out_cols, self.graph, self.buffer_size = red_cols, out_cols, graph, buffer_size

    def initialize(self):
        return {"count": 0, "rows": [None] * (self.buffer_size + 1)}

    def _compact2(self, rows):
        live = [r for r in rows if r is not None]
        return perform_reduce_block(live, self.red_cols, list(range(len(self.red_cols))), self.out_cols, self.graph)

    def _compact(self, buf):
        row = self._compact2(buf["rows"])
        buf["rows"] = [None] * (self.buffer_size + 1)
        buf["rows"][0] = row
        buf["count"] = 1

    def update(self, buf, row):
        c = buf["count"]
        buf["count"] = c + 1
        buf["rows"][c] = tuple(row)
        if c >= self.buffer_size:
            self._compact(buf)

    def merge(self, buf, other):
        initial, oc = buf["count"], other["count"]
        if initial + oc >= self.buffer_size:
            self._compact(buf)
            current, total = 1, 1 + oc
        else:
            current, total = initial, initial + oc
        buf["count"] = total
        for i in range(oc):
            buf["rows"][current + i] = other["rows"][i]

    def evaluate(self, buf):
        assert buf["count"] >= 1
        if buf["count"] == 1:
            return buf["rows"][0]
        return self._compact2(buf["rows"])


def aggregate(partitions, cols: Sequence[Col], key_idx: Sequence[int], graph: bytes, outputs: Sequence[str]):
    """Partial aggregation per partition, merge per key, evaluate.  Returns rows (keys..., outputs...)."""
    idx, _in_cols, out_cols, red_cols = _reduce_cols(cols, outputs)
    udaf = TensorFlowUDAF(red_cols, out_cols, graph)
    partial_maps = []
    for rows in partitions:
        m: Dict[tuple, dict] = {}
        for r in rows:
            key = tuple(r[k] for k in key_idx)
            buf = m.get(key)
            if buf is None:
                buf = m[key] = udaf.initialize()
            udaf.update(buf, tuple(r[i] for i in idx))
        partial_maps.append(m)
    final: Dict[tuple, dict] = {}
    for m in partial_maps:
        for key, buf in m.items():
            if key not in final:
                final[key] = udaf.initialize()
            udaf.merge(final[key], buf)
    return [key + tuple(udaf.evaluate(buf)) for key, buf in final.items()]


# ---- row-wise ------------------------------------------------------------------------------
def _cell_array(value, dtype_code):
    return np.asarray(value, dtype=DT[dtype_code])


def map_rows(partitions, cols: Sequence[Col], graph: bytes, feed: Dict[str, str],
             fetch_info: Dict[str, Tuple[int, Tuple[int, ...]]]):
    """feed: placeholder name -> column name.  fetch_info: fetch -> (dtype code, cell shape).
    Returns (column names, partitions of rows): TF columns (sorted by name) first, then the inputs."""
    by_name = {c.name: i for i, c in enumerate(cols)}
    for ph, col in feed.items():
        if col not in by_name:
            raise Exception(f"Graph input {ph} found, but no column to match it.")
    fetches = sorted(fetch_info)
    for f in fetches:
        if f in by_name:
            raise Exception(f"TF graph has an output node called '{f}', but this column already exists.")
    out_parts = []
    for rows in partitions:
        out_rows = []
        for r in rows:                                   # one Session.run per row
            feeds = {ph: _cell_array(r[by_name[col]], cols[by_name[col]].dtype) for ph, col in feed.items()}
            outs = evaluate(graph, feeds, fetches)
            cells = []
            for f, t in zip(fetches, outs):
                n_rows, shape = infer_physical_shape(int(t.size), tuple(fetch_info[f][1]), 1)
                arr = t.reshape(shape)
                cells.append(arr.tolist() if shape else arr.item())
            out_rows.append(tuple(cells) + tuple(r))
        out_parts.append(out_rows)
    return fetches + [c.name for c in cols], out_parts


def perform_reduce_pairwise(rows, cols: Sequence[Col], graph: bytes) -> tuple:
    assert rows, "Cannot provide empty input"
    result = tuple(rows[0])
    names = [c.name for c in cols]
    for r in rows[1:]:
        feeds = {}
        for i, c in enumerate(cols):
            feeds[c.name + "_1"] = _cell_array(result[i], c.dtype)
            feeds[c.name + "_2"] = _cell_array(r[i], c.dtype)
        outs = evaluate(graph, feeds, names)
        result = tuple(o.tolist() if o.ndim else o.item() for o in outs)
    return result


def reduce_rows(partitions, cols: Sequence[Col], graph: bytes) -> tuple:
    partials = [perform_reduce_pairwise(rows, cols, graph) for rows in partitions if rows]
    if not partials:
        raise ValueError("empty collection")
    return perform_reduce_pairwise(partials, cols, graph)
