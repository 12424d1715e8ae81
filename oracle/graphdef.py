"""Independent GraphDef reader for the oracle (test infrastructure only).

Deliberately shares no code with the product's C++ parser (csrc/graph.cpp) or its
python writer (tensorframes_b200/proto.py), so the two sides check each other.
Schema: /root/reference/src/main/protobuf/tensorflow/core/framework/
graph.proto:14-112, attr_value.proto:16-53, tensor.proto:13-60,
tensor_shape.proto:12-45, types.proto:9-57.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

DT = {1: np.dtype("<f4"), 2: np.dtype("<f8"), 3: np.dtype("<i4"), 9: np.dtype("<i8")}
DT_NAME = {1: "DT_FLOAT", 2: "DT_DOUBLE", 3: "DT_INT32", 9: "DT_INT64"}


def _fields(buf: bytes):
    """Yield (field_number, wire_type, value) for one message."""
    i, n = 0, len(buf)
    while i < n:
        key, i = _varint(buf, i)
        f, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(buf, i)
        elif wt == 1:
            v, i = buf[i:i + 8], i + 8
        elif wt == 2:
            ln, i = _varint(buf, i)
            v, i = buf[i:i + ln], i + ln
            if len(v) != ln:
                raise ValueError("truncated length-delimited field")
        elif wt == 5:
            v, i = buf[i:i + 4], i + 4
        else:
            raise ValueError(f"wire type {wt}")
        yield f, wt, v


def _varint(buf: bytes, i: int) -> Tuple[int, int]:
    shift = v = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, i
        shift += 7


def _signed(v: int, bits: int = 64) -> int:
    return v - (1 << 64) if v >= (1 << 63) else v


def _packed_varints(buf: bytes) -> List[int]:
    out, i = [], 0
    while i < len(buf):
        v, i = _varint(buf, i)
        out.append(_signed(v))
    return out


def parse_shape(buf: bytes) -> Optional[Tuple[int, ...]]:
    dims, unknown = [], False
    for f, wt, v in _fields(buf):
        if f == 2:
            size = 0
            for f2, _, v2 in _fields(v):
                if f2 == 1:
                    size = _signed(v2)
            dims.append(size)
        elif f == 3:
            unknown = bool(v)
    return None if unknown else tuple(dims)


def parse_tensor(buf: bytes) -> np.ndarray:
    dtype, shape, content = 0, (), b""
    fv: List[float] = []
    dv: List[float] = []
    iv: List[int] = []
    lv: List[int] = []
    for f, wt, v in _fields(buf):
        if f == 1:
            dtype = v
        elif f == 2:
            shape = parse_shape(v)
        elif f == 4:
            content = v
        elif f == 5:
            fv.extend(struct.unpack("<%df" % (len(v) // 4), v))
        elif f == 6:
            dv.extend(struct.unpack("<%dd" % (len(v) // 8), v))
        elif f == 7:
            iv.extend(_packed_varints(v) if wt == 2 else [_signed(v)])
        elif f == 10:
            lv.extend(_packed_varints(v) if wt == 2 else [_signed(v)])
    np_dt = DT[dtype]
    n = int(np.prod(shape, dtype=np.int64))
    if content:
        return np.frombuffer(content, dtype=np_dt).reshape(shape).copy()
    vals = {1: fv, 2: dv, 3: iv, 9: lv}[dtype]
    if not vals:
        return np.zeros(shape, dtype=np_dt)
    if len(vals) < n:  # a short list repeats its last value (tensor.proto:25-28)
        vals = list(vals) + [vals[-1]] * (n - len(vals))
    if dtype == 3:
        vals = [((x + (1 << 31)) % (1 << 32)) - (1 << 31) for x in vals]
    return np.asarray(vals, dtype=np_dt).reshape(shape)


def parse_attr(buf: bytes) -> Any:
    for f, wt, v in _fields(buf):
        if f == 2:
            return ("s", v)
        if f == 3:
            return ("i", _signed(v))
        if f == 4:
            return ("f", struct.unpack("<f", v)[0])
        if f == 5:
            return ("b", bool(v))
        if f == 6:
            return ("type", v)
        if f == 7:
            return ("shape", parse_shape(v))
        if f == 8:
            return ("tensor", parse_tensor(v))
        if f == 1:
            return ("list", v)
    return ("none", None)


@dataclass
class NodeDef:
    name: str = ""
    op: str = ""
    inputs: List[str] = field(default_factory=list)
    attr: Dict[str, Any] = field(default_factory=dict)


def parse_graph(buf: bytes) -> List[NodeDef]:
    nodes = []
    for f, wt, v in _fields(bytes(buf)):
        if f != 1:
            continue
        nd = NodeDef()
        for f2, wt2, v2 in _fields(v):
            if f2 == 1:
                nd.name = v2.decode()
            elif f2 == 2:
                nd.op = v2.decode()
            elif f2 == 3:
                nd.inputs.append(v2.decode())
            elif f2 == 5:
                key, val = None, ("none", None)
                for f3, _, v3 in _fields(v2):
                    if f3 == 1:
                        key = v3.decode()
                    elif f3 == 2:
                        val = parse_attr(v3)
                nd.attr[key] = val
        nodes.append(nd)
    return nodes
