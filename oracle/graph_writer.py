"""Independent GraphDef WRITER for the oracle (test infrastructure only).

The oracle's known-answer tests used to take their input graphs from the product's own DSL
(tensorframes_b200/dsl.py + proto.py), so an encoding bug there would have been shared by the checker
and the thing checked.  This module restates the reference's graph construction for the op vocabulary of
its own tests -- Placeholder, Const, Identity, Add, Sum / Min / Max -- and shares no code with the
product:

  * node layout and attributes: dsl/DslImpl.scala:77-88 (build: `T` = dtype on every op but Placeholder /
    Const, which carry `dtype`), :90-107 (extractPlaceholder: `dtype` + `shape`), :175-188 (build_reducer:
    second input `<parent>/reduction_indices`, an int32 Const; attrs `Tidx` = DT_INT32, `keep_dims` = false);
  * Const payload: impl/DenseTensor.scala:73-98 (typed `double_val` / `float_val` / `int_val` lists, not
    `tensor_content`); int64 uses `int64_val` (tensor.proto:52);
  * automatic names `Const`, `Const_1`, ... : dsl/Paths.scala, checked against T/dsl/BasicSuite.scala:12-33;
  * wire format: graph.proto:14-112, attr_value.proto:16-53, tensor.proto:13-60, tensor_shape.proto:12-45,
    types.proto:9-57 (proto3: repeated scalars are packed).

`from_spec` turns the graph description of a case of tests/golden/reference_vectors.json into GraphDef
bytes plus the (dtype, shape) of every fetch, the way the reference's DSL infers them
(dsl/DslImpl.scala:109-140 broadcast of Add, :190-200 reduce_shape).
"""
from __future__ import annotations

import struct
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

DT_FLOAT, DT_DOUBLE, DT_INT32, DT_INT64 = 1, 2, 3, 9
CODE = {"float": DT_FLOAT, "double": DT_DOUBLE, "int": DT_INT32, "long": DT_INT64}
NP = {DT_FLOAT: "<f4", DT_DOUBLE: "<f8", DT_INT32: "<i4", DT_INT64: "<i8"}
REDUCERS = {"reduce_sum": "Sum", "reduce_min": "Min", "reduce_max": "Max"}


# ---- protobuf wire format --------------------------------------------------------------------------------
def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1                      # negative int32 / int64 values travel as 10-byte varints
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _tag(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _bytes_field(field: int, payload: bytes) -> bytes:
    return _tag(field, 2) + _varint(len(payload)) + payload


def _varint_field(field: int, v: int) -> bytes:
    return _tag(field, 0) + _varint(v)


# ---- framework messages ------------------------------------------------------------------------------------
def shape_proto(dims: Sequence[int]) -> bytes:
    """TensorShapeProto: one `dim { size }` per dimension, -1 = unknown (tensor_shape.proto:20-38)."""
    out = b""
    for d in dims:
        out += _bytes_field(2, _varint_field(1, int(d)) if d != 0 else b"")
    return out


def tensor_proto(code: int, values: np.ndarray) -> bytes:
    """TensorProto with typed value lists (DenseTensor.scala:73-98)."""
    values = np.asarray(values, dtype=NP[code])
    out = _varint_field(1, code) + _bytes_field(2, shape_proto(values.shape))
    flat = values.reshape(-1)
    if code == DT_DOUBLE:
        out += _bytes_field(6, struct.pack("<%dd" % flat.size, *flat.tolist()))
    elif code == DT_FLOAT:
        out += _bytes_field(5, struct.pack("<%df" % flat.size, *flat.tolist()))
    elif code == DT_INT32:
        out += _bytes_field(7, b"".join(_varint(int(x)) for x in flat.tolist()))
    elif code == DT_INT64:
        out += _bytes_field(10, b"".join(_varint(int(x)) for x in flat.tolist()))
    else:
        raise ValueError("dtype code %d" % code)
    return out


def attr_type(code: int) -> bytes:
    return _varint_field(6, code)


def attr_shape(dims: Sequence[int]) -> bytes:
    return _bytes_field(7, shape_proto(dims))


def attr_tensor(code: int, values: np.ndarray) -> bytes:
    return _bytes_field(8, tensor_proto(code, values))


def attr_bool(v: bool) -> bytes:
    return _varint_field(5, 1 if v else 0)


def node_def(name: str, op: str, inputs: Sequence[str], attrs: Dict[str, bytes]) -> bytes:
    out = _bytes_field(1, name.encode()) + _bytes_field(2, op.encode())
    for i in inputs:
        out += _bytes_field(3, i.encode())
    for key in sorted(attrs):               # map entries: key = 1, value = 2
        out += _bytes_field(5, _bytes_field(1, key.encode()) + _bytes_field(2, attrs[key]))
    return out


# ---- graph construction ------------------------------------------------------------------------------------
class GraphWriter:
    """Nodes in creation order; every node knows its dtype code and its shape (-1 = unknown dimension)."""

    def __init__(self):
        self.nodes: List[bytes] = []
        self.meta: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        self._auto: Dict[str, int] = {}

    def _name(self, requested: Optional[str], op: str) -> str:
        if requested is not None:
            if requested in self.meta:
                raise ValueError("duplicate node name " + requested)
            return requested
        k = self._auto.get(op, 0)
        while True:
            name = op if k == 0 else "%s_%d" % (op, k)
            k += 1
            if name not in self.meta:
                self._auto[op] = k
                return name

    def _add(self, name, op, inputs, attrs, code, shape):
        self.nodes.append(node_def(name, op, inputs, attrs))
        self.meta[name] = (code, tuple(int(d) for d in shape))
        return name

    def placeholder(self, code: int, shape: Sequence[int], name: str) -> str:
        return self._add(self._name(name, "Placeholder"), "Placeholder", [],
                         {"dtype": attr_type(code), "shape": attr_shape(shape)}, code, shape)

    def constant(self, code: int, values, name: Optional[str] = None) -> str:
        values = np.asarray(values, dtype=NP[code])
        return self._add(self._name(name, "Const"), "Const", [],
                         {"dtype": attr_type(code), "value": attr_tensor(code, values)}, code, values.shape)

    def identity(self, x: str, name: Optional[str] = None) -> str:
        code, shape = self.meta[x]
        return self._add(self._name(name, "Identity"), "Identity", [x], {"T": attr_type(code)}, code, shape)

    def add(self, x: str, y: str, name: Optional[str] = None) -> str:
        (cx, sx), (cy, sy) = self.meta[x], self.meta[y]
        if cx != cy:
            raise ValueError("Add of two dtypes")
        return self._add(self._name(name, "Add"), "Add", [x, y], {"T": attr_type(cx)}, cx, broadcast(sx, sy))

    def reducer(self, op: str, x: str, indices: Optional[Sequence[int]], name: Optional[str] = None) -> str:
        code, shape = self.meta[x]
        if indices is None:                 # python's tf.reduce_sum(x): every axis, sent as range(rank) (core_test.py:93-103)
            indices = list(range(len(shape)))
        idx = self.constant(DT_INT32, np.asarray(list(indices), dtype="<i4"), name=x + "/reduction_indices")
        out_shape = () if len(indices) == 0 else tuple(d for i, d in enumerate(shape) if i not in set(indices))
        return self._add(self._name(name, op), op, [x, idx],
                         {"T": attr_type(code), "Tidx": attr_type(DT_INT32), "keep_dims": attr_bool(False)}, code, out_shape)

    def graph(self) -> bytes:
        return b"".join(_bytes_field(1, n) for n in self.nodes)


def broadcast(a: Sequence[int], b: Sequence[int]) -> Tuple[int, ...]:
    """numpy-style broadcast of two shapes with -1 for unknown dimensions (dsl/DslImpl.scala:109-140)."""
    a, b = list(a), list(b)
    while len(a) < len(b):
        a.insert(0, 1)
    while len(b) < len(a):
        b.insert(0, 1)
    out = []
    for x, y in zip(a, b):
        if x == y or y == 1:
            out.append(x)
        elif x == 1:
            out.append(y)
        elif x == -1 or y == -1:
            out.append(max(x, y) if min(x, y) == -1 and max(x, y) > 1 else -1)
        else:
            raise ValueError("shapes %r and %r do not broadcast" % (a, b))
    return tuple(out)


def from_spec(spec: dict, columns: Dict[str, Tuple[int, Tuple[int, ...]]]):
    """(GraphDef bytes, {fetch name: (dtype code, shape)}) for a case's graph description.  `columns` maps a column
    name to its (dtype code, block shape) -- what `block` / `row` placeholders take from the DataFrame's metadata
    (python front end: core.py:338-391, the lead dimension of a block is always unknown)."""
    w = GraphWriter()
    env: Dict[str, str] = {}
    for name, dt, shape in spec.get("placeholders", []):
        env[name] = w.placeholder(CODE[dt], [-1 if d is None else d for d in shape], name)
    for col, tf_name in spec.get("block_placeholders", []):
        code, shape = columns[col]
        env[tf_name] = w.placeholder(code, (-1,) + tuple(shape[1:]), tf_name)
    for col, tf_name in spec.get("row_placeholders", []):
        code, shape = columns[col]
        env[tf_name] = w.placeholder(code, tuple(shape[1:]), tf_name)
    for name, dt, value in spec.get("consts", []):
        named = name in spec["fetches"] or name in spec.get("keep_consts", [])
        env[name] = w.constant(CODE[dt], value, name if named else None)
    for op in spec.get("ops", []):
        name, kind, args = op[0], op[1], op[2:]
        if kind == "identity":
            env[name] = w.identity(env[args[0]], name)
        elif kind == "add":
            env[name] = w.add(env[args[0]], env[args[1]], name)
        elif kind in REDUCERS:
            env[name] = w.reducer(REDUCERS[kind], env[args[0]], args[1], name)
        else:
            raise KeyError(kind)
    return w.graph(), {f: w.meta[env[f]] for f in spec["fetches"]}
