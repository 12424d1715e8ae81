"""Minimal protobuf wire writer for the TF 1.x GraphDef subset the DSL emits.

The reference builds these messages with protoc-generated Java
(src/main/java/org/tensorflow/framework/*, from
src/main/protobuf/tensorflow/core/framework/{graph,attr_value,tensor,tensor_shape,types}.proto).
There is no protoc here, and only five messages matter, so they are written by hand.
Field numbers are those of the .proto files cited in each function.
"""
from __future__ import annotations

from typing import Dict, Iterable, Sequence

import numpy as np

from .shape import ScalarType, Shape, DoubleType, FloatType, IntegerType, LongType


def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
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


def _ld(field: int, payload: bytes) -> bytes:
    return _tag(field, 2) + _varint(len(payload)) + payload


def _vi(field: int, v: int) -> bytes:
    return _tag(field, 0) + _varint(v)


def shape_proto(shape: Shape) -> bytes:
    """TensorShapeProto (tensor_shape.proto:12-45): repeated Dim dim = 2 {int64 size = 1}."""
    out = b""
    for d in shape.dims:
        # proto3 omits zero scalars; a dim of size 0 is an empty Dim message
        out += _ld(2, _vi(1, d) if d != 0 else b"")
    return out


def tensor_proto(dtype: ScalarType, shape: Shape, values: np.ndarray, use_content: bool = False) -> bytes:
    """TensorProto (tensor.proto:13-60).

    Typed value lists as DenseTensor.toTensorProto writes them
    (reference: impl/DenseTensor.scala:73-98), or tensor_content as python TF does.
    """
    flat = np.ascontiguousarray(values, dtype=dtype.np_dtype).reshape(-1)
    out = _vi(1, dtype.tf_code) + _ld(2, shape_proto(shape))
    if use_content:
        return out + _ld(4, flat.tobytes())
    if flat.size == 0:
        return out
    if dtype is DoubleType:
        out += _ld(6, flat.astype("<f8").tobytes())
    elif dtype is FloatType:
        out += _ld(5, flat.astype("<f4").tobytes())
    elif dtype is IntegerType:
        out += _ld(7, b"".join(_varint(int(v)) for v in flat))
    elif dtype is LongType:
        out += _ld(10, b"".join(_varint(int(v)) for v in flat))
    else:
        raise ValueError(f"Cannot convert type {dtype}")
    return out


class AttrValue:
    """AttrValue (attr_value.proto:16-53): only the alternatives the DSL needs."""

    def __init__(self, payload: bytes, text: str):
        self.payload = payload
        self.text = text

    @staticmethod
    def of_type(dtype: ScalarType) -> "AttrValue":
        return AttrValue(_vi(6, dtype.tf_code), f"type:{dtype.name}")

    @staticmethod
    def of_shape(shape: Shape) -> "AttrValue":
        return AttrValue(_ld(7, shape_proto(shape)), f"shape:{shape}")

    @staticmethod
    def of_bool(b: bool) -> "AttrValue":
        # an explicit oneof member is written even when false
        return AttrValue(_vi(5, 1 if b else 0), f"b:{b}")

    @staticmethod
    def of_int(v: int) -> "AttrValue":
        return AttrValue(_vi(3, int(v)), f"i:{int(v)}")

    @staticmethod
    def of_tensor(dtype: ScalarType, shape: Shape, values, use_content: bool = False) -> "AttrValue":
        return AttrValue(_ld(8, tensor_proto(dtype, shape, np.asarray(values), use_content)),
                         f"tensor:{dtype.name}{shape}")

    def __repr__(self) -> str:
        return f"AttrValue({self.text})"


def node_def(name: str, op: str, inputs: Sequence[str], attrs: Dict[str, AttrValue]) -> bytes:
    """NodeDef (graph.proto:60-112): name=1, op=2, input=3, attr=5 (map<string, AttrValue>)."""
    out = _ld(1, name.encode()) + _ld(2, op.encode())
    for i in inputs:
        out += _ld(3, i.encode())
    for key in sorted(attrs):
        entry = _ld(1, key.encode()) + _ld(2, attrs[key].payload)
        out += _ld(5, entry)
    return out


def graph_def(nodes: Iterable[bytes]) -> bytes:
    """GraphDef (graph.proto:14-56): repeated NodeDef node = 1."""
    return b"".join(_ld(1, n) for n in nodes)
