"""numpy evaluator of the TF 1.1 op subset (test infrastructure only).

This restates what `Session.run` computes for the reference
(/root/reference/src/main/scala/org/tensorframes/impl/DebugRowOps.scala:794, 844, 912).
The kernels themselves are in the un-vendored dependency
org.tensorflow:tensorflow:1.1.0-rc1; semantics follow the TF op definitions:

  Add/Sub/Mul        IEEE-754 round-to-nearest per element; integers wrap
  Div / RealDiv      floats: IEEE division; integers: C++ `/` (truncates toward zero),
                     "Integer division by zero" error on a zero divisor
  Neg/Abs/Square/Sqrt/Reciprocal/Relu   per element
  Minimum/Maximum    Eigen scalar forms  (b < a ? b : a)  /  (a < b ? b : a)
  Cast               C++ static_cast (float->int truncates, double->float rounds to nearest)
  Sum/Min/Max        over `reduction_indices`, keep_dims attr; float sums are accumulated in
                     extended precision so the oracle is the *exact* reference for a
                     tolerance check (summation order is not specified by TF)
  MatMul             float32 operands, products accumulated in float64, rounded once
  Fill, Identity, Const, Placeholder
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np

from .graphdef import DT, NodeDef, parse_graph


class OracleError(Exception):
    pass


def _strip(name: str) -> str:
    if name.startswith("^"):
        raise OracleError("control inputs are not part of the subset")
    if ":" in name:
        base, idx = name.rsplit(":", 1)
        if idx != "0":
            raise OracleError(f"only output 0 is addressable: {name}")
        return base
    return name


def _int_div(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    if np.any(b == 0):
        raise OracleError("Integer division by zero")
    q = np.abs(a.astype(object)) // np.abs(b.astype(object))
    sign = np.where((a < 0) != (b < 0), -1, 1)
    bits = a.dtype.itemsize * 8
    out = np.asarray(q * sign, dtype=object)
    out = ((out + (1 << (bits - 1))) % (1 << bits)) - (1 << (bits - 1))
    return np.asarray(out, dtype=object).astype(a.dtype).reshape(np.broadcast(a, b).shape)


def _reduce(op: str, x: np.ndarray, axes: Sequence[int], keep: bool) -> np.ndarray:
    axes = tuple(sorted({a + x.ndim if a < 0 else a for a in axes}))
    if not axes:
        return x.copy()
    if op == "Sum":
        if x.dtype.kind == "f":
            return np.sum(x.astype(np.longdouble), axis=axes, keepdims=keep).astype(x.dtype)
        with np.errstate(over="ignore"):
            return np.sum(x, axis=axes, keepdims=keep, dtype=x.dtype)
    if op == "Min":
        return np.min(x, axis=axes, keepdims=keep, initial=_identity("Min", x.dtype))
    return np.max(x, axis=axes, keepdims=keep, initial=_identity("Max", x.dtype))


def _identity(op: str, dt: np.dtype):
    if dt.kind == "f":
        return np.inf if op == "Min" else -np.inf
    info = np.iinfo(dt)
    return info.max if op == "Min" else info.min


def evaluate(graph_bytes: bytes, feeds: Dict[str, np.ndarray], fetches: Sequence[str]) -> List[np.ndarray]:
    nodes = {n.name: n for n in parse_graph(graph_bytes)}
    cache: Dict[str, np.ndarray] = {}

    def attr(n: NodeDef, key: str, kind: str, default=None):
        v = n.attr.get(key)
        if v is None or v[0] != kind:
            if default is not None:
                return default
            raise OracleError(f"node {n.name}: missing attr {key}")
        return v[1]

    def ev(name: str) -> np.ndarray:
        name = _strip(name)
        if name in cache:
            return cache[name]
        n = nodes.get(name)
        if n is None:
            raise OracleError(f"unknown node {name}")
        op = n.op
        if op == "Placeholder":
            if name not in feeds:
                raise OracleError(f"placeholder {name} is not fed")
            v = np.asarray(feeds[name])
            want = DT[attr(n, "dtype", "type")]
            if v.dtype != want:
                raise OracleError(f"placeholder {name} is {want} but was fed {v.dtype}")
            out = v
        elif op == "Const":
            out = attr(n, "value", "tensor")
        elif op == "Identity":
            out = ev(n.inputs[0])
        elif op in ("Add", "BiasAdd", "Sub", "Mul", "Div", "RealDiv", "Minimum", "Maximum"):
            a, b = ev(n.inputs[0]), ev(n.inputs[1])
            if a.dtype != b.dtype:
                raise OracleError(f"node {name}: operand dtypes differ ({a.dtype}, {b.dtype})")
            with np.errstate(all="ignore"):
                if op in ("Add", "BiasAdd"):
                    out = a + b
                elif op == "Sub":
                    out = a - b
                elif op == "Mul":
                    out = a * b
                elif op in ("Div", "RealDiv"):
                    out = _int_div(a, b) if a.dtype.kind == "i" else (a / b).astype(a.dtype)
                elif op == "Minimum":
                    out = np.where(b < a, b, a)
                else:
                    out = np.where(a < b, b, a)
            out = np.asarray(out, dtype=a.dtype)
        elif op in ("Neg", "Relu", "Square", "Abs", "Sqrt", "Reciprocal", "Inv"):
            a = ev(n.inputs[0])
            with np.errstate(all="ignore"):
                if op == "Neg":
                    out = -a
                elif op == "Relu":
                    z = np.zeros((), dtype=a.dtype)
                    out = np.where(a < z, z, a)
                elif op == "Square":
                    out = a * a
                elif op == "Abs":
                    out = np.abs(a)
                elif op == "Sqrt":
                    out = np.sqrt(a)
                else:
                    out = (np.ones((), dtype=a.dtype) / a)
            out = np.asarray(out, dtype=a.dtype)
        elif op == "Cast":
            a = ev(n.inputs[0])
            dst = DT[attr(n, "DstT", "type")]
            with np.errstate(all="ignore"):
                out = a.astype(dst)
        elif op in ("Sum", "Min", "Max"):
            a = ev(n.inputs[0])
            axes = [int(v) for v in np.asarray(ev(n.inputs[1])).reshape(-1)]
            out = _reduce(op, a, axes, bool(attr(n, "keep_dims", "b", default=False)))
        elif op == "MatMul":
            a, b = ev(n.inputs[0]), ev(n.inputs[1])
            if attr(n, "transpose_a", "b", default=False):
                a = a.T
            if attr(n, "transpose_b", "b", default=False):
                b = b.T
            out = (a.astype(np.float64) @ b.astype(np.float64)).astype(a.dtype)
        elif op == "Fill":
            dims = [int(v) for v in np.asarray(ev(n.inputs[0])).reshape(-1)]
            val = ev(n.inputs[1])
            out = np.full(dims, val, dtype=val.dtype)
        elif op in ("OnesLike", "ZerosLike"):
            a = ev(n.inputs[0])
            out = np.ones_like(a) if op == "OnesLike" else np.zeros_like(a)
        elif op == "Shape":
            a = ev(n.inputs[0])
            out = np.asarray(a.shape, dtype=DT[attr(n, "out_type", "type", default=3)])
        elif op == "StridedSlice":
            # x[b:e:s] / x[i] on vectors, as python TF emits for tf.shape(x)[0]
            a = ev(n.inputs[0])
            b, e, st = (np.asarray(ev(i)).reshape(-1) for i in n.inputs[1:4])
            if a.ndim != 1 or len(b) != 1:
                raise OracleError(f"node {name}: only one-dimensional StridedSlice is in the subset")
            if attr(n, "ellipsis_mask", "i", default=0) or attr(n, "new_axis_mask", "i", default=0):
                raise OracleError(f"node {name}: ellipsis / new-axis masks are not in the subset")
            if attr(n, "shrink_axis_mask", "i", default=0) & 1:
                out = np.asarray(a[int(b[0])])
            else:
                lo = None if attr(n, "begin_mask", "i", default=0) & 1 else int(b[0])
                hi = None if attr(n, "end_mask", "i", default=0) & 1 else int(e[0])
                out = a[lo:hi:int(st[0])]
        elif op == "Pack":
            out = np.stack([ev(i) for i in n.inputs], axis=int(attr(n, "axis", "i", default=0)))
        elif op == "ExpandDims":
            out = np.expand_dims(ev(n.inputs[0]), int(np.asarray(ev(n.inputs[1])).reshape(-1)[0]))
        elif op == "Tile":
            a = ev(n.inputs[0])
            reps = [int(v) for v in np.asarray(ev(n.inputs[1])).reshape(-1)]
            if len(reps) != a.ndim:
                raise OracleError(f"node {name}: multiples must have one entry per dimension")
            out = np.tile(a, reps)
        elif op in ("ArgMin", "ArgMax"):
            a = ev(n.inputs[0])
            ax = int(np.asarray(ev(n.inputs[1])).reshape(-1)[0])
            idx = np.argmin(a, axis=ax) if op == "ArgMin" else np.argmax(a, axis=ax)   # first extreme wins, like Eigen
            out = idx.astype(DT[attr(n, "output_type", "type", default=9)])
        elif op == "UnsortedSegmentSum":
            data, ids = ev(n.inputs[0]), np.asarray(ev(n.inputs[1]))
            k = int(np.asarray(ev(n.inputs[2])).reshape(-1)[0])
            if ids.ndim != 1 or ids.shape[0] != data.shape[0]:
                raise OracleError(f"node {name}: segment_ids must be a vector over the rows of data")
            if ids.size and (ids.min() < 0 or ids.max() >= k):
                raise OracleError(f"node {name}: segment ids out of range")
            out = np.zeros((k,) + data.shape[1:], dtype=data.dtype)
            for i in range(data.shape[0]):          # row order, one accumulation per row: the CPU kernel's order
                out[ids[i]] = out[ids[i]] + data[i]
        else:
            raise OracleError(f"op {op} is not in the subset")
        cache[name] = np.asarray(out)
        return cache[name]

    return [ev(f) for f in fetches]


def exact_sum_f64(x: np.ndarray, axis: int = 0) -> np.ndarray:
    """math.fsum along an axis: the correctly rounded float64 sum (small inputs only)."""
    import math
    x = np.moveaxis(np.asarray(x, dtype=np.float64), axis, 0)
    flat = x.reshape(x.shape[0], -1)
    return np.asarray([math.fsum(flat[:, j]) for j in range(flat.shape[1])]).reshape(x.shape[1:])
