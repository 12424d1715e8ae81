"""Graph-building DSL, mirroring the reference's Scala DSL.

Reference: src/main/scala/org/tensorframes/dsl/package.scala:31-130 (the public
functions), dsl/Operation.scala:15-146 (Node: lazy naming, NodeDef emission),
dsl/DslImpl.scala:38-188 (graph closure, constants, placeholders, reducers,
broadcast rule), dsl/Paths.scala:13-56 (scopes and name counters).

Nodes are named lazily: a node gets its final name when it is frozen (by
`named`, or when a graph is built), from the scope path it was created in plus a
per-graph counter for repeated names ("Const", "Const_1", ...), exactly as
Paths.path does.  `build_graph` returns binary GraphDef bytes.

Beyond the reference's op list (Placeholder, Const, Identity, Add, Div, Fill,
Sum, Min) the north-star subset is exposed too: sub, mul, neg, cast, minimum,
maximum, relu, reduce_max, matmul.
"""
from __future__ import annotations

import contextlib
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from . import proto
from .proto import AttrValue
from .shape import (DoubleType, FloatType, IntegerType, LongType, ScalarType, Shape, ShapeDescription,
                    Unknown, scalar_type_from_numpy)

__all__ = [
    "Unknown", "Node", "scope", "withGraph", "placeholder", "constant", "zeros", "ones", "fill",
    "identity", "add", "sub", "mul", "div", "neg", "cast", "minimum", "maximum", "relu",
    "reduce_sum", "reduce_min", "reduce_max", "matmul", "block", "row", "build_graph", "hints",
    "square", "reciprocal", "inv", "sqrt", "abs_", "ones_like", "zeros_like", "shape", "strided_index", "stack",
    "expand_dims", "tile", "argmin", "argmax", "unsorted_segment_sum",
]


# --- Paths (dsl/Paths.scala:13-56) ------------------------------------------------
class _Paths:
    def __init__(self):
        self.rpath: List[str] = []
        self.counters: Dict[str, int] = {}

    def creation_path(self) -> List[str]:
        return list(self.rpath)

    @staticmethod
    def _join(parts: Sequence[str]) -> str:
        return "/".join(p for p in reversed(parts) if p)

    def path(self, creation_path: List[str], requested: Optional[str], op_name: str) -> str:
        full = list(reversed((requested if requested is not None else op_name).split("/"))) + creation_path
        key = self._join(full)
        c = self.counters.get(key, 0)
        self.counters[key] = c + 1
        return key if c == 0 else f"{key}_{c}"


_paths = _Paths()


@contextlib.contextmanager
def scope(path_elem: str):
    """Name scope (dsl/package.scala:31-33)."""
    _paths.rpath.insert(0, path_elem)
    try:
        yield
    finally:
        _paths.rpath.pop(0)


@contextlib.contextmanager
def withGraph():
    """Fresh name counters for the enclosed block (dsl/package.scala:38)."""
    old = _paths.counters
    _paths.counters = {}
    try:
        yield
    finally:
        _paths.counters = old


# --- Node (dsl/Operation.scala:30-146) ----------------------------------------------
class Node:
    def __init__(self, requested_name: Optional[str], op_name: str, scalar_type: ScalarType, shape: Shape,
                 parents: Sequence["Node"], internal_parents: Callable[[str], Sequence["Node"]],
                 is_op: bool, extra_attr: Dict[str, AttrValue], creation_path: Optional[List[str]] = None):
        self.requestedName = requested_name
        self.creationPath = _paths.creation_path() if creation_path is None else creation_path
        self.opName = op_name
        self.scalarType = scalar_type
        self.shape = shape
        self.parents = list(parents)
        self.internalParents = internal_parents
        self.isOp = is_op
        self.extraAttr = dict(extra_attr)
        self._path: Optional[str] = None
        self._created_parents: List[Node] = []
        # value of small integer nodes that follows from shapes and constants (Unknown where it does not):
        # what lets tile / unsorted_segment_sum give their outputs a static shape, as TF's shape inference does
        self.ivalue: Optional[List[int]] = None

    @property
    def frozen(self) -> bool:
        return self._path is not None

    def freeze(self, everything: bool = False) -> None:
        if not self.frozen:
            p = _paths.path(self.creationPath, self.requestedName, self.opName)
            self._path = p
            diff = "/".join(x for x in reversed(self.creationPath) if x)
            suffix = p[len(diff):]
            ns = list(self.internalParents(suffix))
            for n in ns:
                n.freeze()
            self._created_parents = ns
        if everything:
            for n in self._all_parents():
                n.freeze(everything)

    def _all_parents(self) -> List["Node"]:
        return self.parents + self._created_parents

    @property
    def name(self) -> str:
        if not self.frozen:
            raise RuntimeError(f"The node {self} is not frozen")
        return self._path

    @property
    def dims(self) -> List[int]:
        return list(self.shape.dims)

    def named(self, new_name: str) -> "Node":
        c = Node(new_name, self.opName, self.scalarType, self.shape, self.parents, self.internalParents,
                 self.isOp, self.extraAttr, creation_path=self.creationPath)
        c.ivalue = self.ivalue
        c.freeze()
        return c

    def __getitem__(self, index: int) -> "Node":
        """x[i] on an integer vector such as shape(x): a StridedSlice with shrink_axis_mask = 1, what python TF
        emits for tf.shape(points)[0] (tensorframes_snippets/kmeans_demo.py:26)."""
        return strided_index(self, index)

    def node_defs(self) -> List[bytes]:
        self.freeze()
        attrs = {("T" if self.isOp else "dtype"): AttrValue.of_type(self.scalarType)}
        attrs.update(self.extraAttr)
        out = [proto.node_def(self.name, self.opName, [p.name for p in self._all_parents()], attrs)]
        for p in self._created_parents:
            out.extend(p.node_defs())
        return out

    # operators (dsl/Operation.scala:55-57 has + and /)
    def __add__(self, other): return add(self, _lift(other, self))
    def __radd__(self, other): return add(_lift(other, self), self)
    def __sub__(self, other): return sub(self, _lift(other, self))
    def __rsub__(self, other): return sub(_lift(other, self), self)
    def __mul__(self, other): return mul(self, _lift(other, self))
    def __rmul__(self, other): return mul(_lift(other, self), self)
    def __truediv__(self, other): return div(self, _lift(other, self))
    def __rtruediv__(self, other): return div(_lift(other, self), self)
    def __neg__(self): return neg(self)

    def __repr__(self) -> str:
        f = "frz" if self.frozen else "liv"
        return f"Node({f}, {self.requestedName}, {self.creationPath}, {self.opName}, {self.scalarType}, {self.shape})"


Operation = Node


def _lift(x, like: Node) -> Node:
    """Implicit conversion of python scalars / sequences to constants (dsl/Implicits.scala)."""
    if isinstance(x, Node):
        return x
    if type(x) in (int, float):
        # a bare python number takes the dtype of the tensor it meets, as python TF's operator overloads do
        # (x_float32 * 1.5 is a float32 product); arrays and sequences keep their own dtype
        return constant(like.scalarType.np_dtype.type(x))
    return constant(x)


# --- builders (dsl/DslImpl.scala:143-157) --------------------------------------------
def _common_shape(shapes: Sequence[Shape]) -> Shape:
    assert shapes and all(s == shapes[0] for s in shapes), f"{shapes}"
    return shapes[0]


def _common_type(dtypes: Sequence[ScalarType]) -> ScalarType:
    assert dtypes and all(d == dtypes[0] for d in dtypes), f"All these types should be the same: {dtypes}"
    return dtypes[0]


def broadcast_shape(shapes: Sequence[Shape]) -> Shape:
    """numpy-style broadcast where unknown dims give way to known ones (dsl/DslImpl.scala:118-135)."""
    assert len(shapes) == 2, shapes
    s1, s2 = shapes
    if s1.numDims < s2.numDims:
        return broadcast_shape([s2, s1])
    k = s1.numDims - s2.numDims
    t = list(s1.dims[:k])
    h = []
    for d1, d2 in zip(s1.dims[k:], s2.dims):
        if d1 == Unknown or d1 == 1:
            h.append(d2)
        elif d2 == Unknown or d2 == 1:
            h.append(d1)
        elif d1 == d2:
            h.append(d1)
        else:
            raise Exception(f"Incompatible shapes: {s1} {s2}")
    return Shape(t + h)


def _build(op_name: str, name: Optional[str] = None, parents: Sequence[Node] = (),
           extra_parents: Callable[[str], Sequence[Node]] = lambda _p: [], is_op: bool = True,
           dtype: Optional[ScalarType] = None, shape: Optional[Shape] = None,
           dtype_infer=_common_type, shape_infer=_common_shape,
           extra_attrs: Optional[Dict[str, AttrValue]] = None) -> Node:
    dt = dtype if dtype is not None else dtype_infer([p.scalarType for p in parents])
    sh = shape if shape is not None else shape_infer([p.shape for p in parents])
    return Node(name, op_name, dt, sh, parents, extra_parents, is_op, extra_attrs or {})


# --- public DSL (dsl/package.scala:45-130) ----------------------------------------
_PY_TYPES = {float: DoubleType, int: IntegerType}


def _dense(x, dtype: Optional[ScalarType] = None):
    """python value -> (ScalarType, Shape, ndarray); floats are doubles and ints are int32,
    like the reference's ConvertibleToDenseTensor instances (dsl/ConvertibleToTensor.scala:24-67)."""
    if isinstance(x, np.ndarray) or isinstance(x, np.generic):
        arr = np.asarray(x)
        st = dtype or scalar_type_from_numpy(arr.dtype)
        if st is None:
            raise ValueError(f"Type {arr.dtype} is not supported")
        return st, Shape(arr.shape), arr.astype(st.np_dtype)
    flat = x
    depth = 0
    while isinstance(flat, (list, tuple)):
        depth += 1
        if not flat:
            break
        flat = flat[0]
    if dtype is None:
        if isinstance(flat, bool) or not isinstance(flat, (int, float)):
            if isinstance(flat, (list, tuple)):  # empty sequence
                dtype = IntegerType
            else:
                raise ValueError(f"Type {type(flat)} is not supported")
        else:
            dtype = _PY_TYPES[type(flat)]
    arr = np.asarray(x, dtype=dtype.np_dtype)
    if arr.ndim > 2:
        raise ValueError(f"Shape {Shape(arr.shape)} is too high - only scalars, vectors and matrices convert")
    return dtype, Shape(arr.shape), arr


def placeholder(dtype: ScalarType, *shape: int) -> Node:
    """tf.placeholder (dsl/package.scala:45-48, DslImpl.scala:84-88)."""
    sh = Shape(*shape)
    return _build("Placeholder", shape=sh, dtype=dtype, is_op=False,
                  extra_attrs={"shape": AttrValue.of_shape(sh)})


def constant(x, dtype: Optional[ScalarType] = None) -> Node:
    """tf.constant (dsl/package.scala:50-53, DslImpl.build_constant DslImpl.scala:77-82)."""
    st, sh, arr = _dense(x, dtype)
    n = _build("Const", is_op=False, shape=sh, dtype=st,
               extra_attrs={"value": AttrValue.of_tensor(st, sh, arr)})
    if st in (IntegerType, LongType) and arr.ndim <= 1 and arr.size <= 64:
        n.ivalue = [int(v) for v in arr.reshape(-1)]
    return n


def fill(dims, value) -> Node:
    """tf.fill (dsl/package.scala:67-88)."""
    if not isinstance(dims, Node):
        dims = list(dims)
        if len(dims) == 0:
            dims_n = constant(1)
        elif len(dims) == 1:
            dims_n = constant([int(dims[0])])
        else:
            raise Exception(f"Shape {Shape(dims)} is too high - tensorframes only supports dimensions <= 1 (vectors)")
        out_shape = Shape(dims)
    else:
        dims_n = dims
        out_shape = dims.shape
    value_n = value if isinstance(value, Node) else constant(value)
    assert dims_n.scalarType is IntegerType
    assert value_n.shape.numDims == 0, value_n.shape
    return _build("Fill", shape=out_shape, dtype=value_n.scalarType,
                  extra_parents=lambda p: [dims_n.named(p + "/dims"), value_n.named(p + "/value")])


def zeros(*shape: int, dtype: ScalarType = FloatType) -> Node:
    return fill(shape, constant(dtype.np_dtype.type(0)))


def ones(*shape: int, dtype: ScalarType = FloatType) -> Node:
    return fill(shape, constant(dtype.np_dtype.type(1)))


def identity(op: Node) -> Node:
    return _build("Identity", parents=[op])


def _binary(op_name: str, x, y) -> Node:
    if not isinstance(x, Node):
        x = _lift(x, y)
    if not isinstance(y, Node):
        y = _lift(y, x)
    return _build(op_name, parents=[x, y], shape_infer=broadcast_shape)


def add(x, y) -> Node: return _binary("Add", x, y)
def sub(x, y) -> Node: return _binary("Sub", x, y)
def mul(x, y) -> Node: return _binary("Mul", x, y)
def div(x, y) -> Node: return _binary("Div", x, y)
def minimum(x, y) -> Node: return _binary("Minimum", x, y)
def maximum(x, y) -> Node: return _binary("Maximum", x, y)


def neg(x: Node) -> Node: return _build("Neg", parents=[x])
def relu(x: Node) -> Node: return _build("Relu", parents=[x])


def cast(x: Node, dtype: ScalarType) -> Node:
    n = Node(None, "Cast", dtype, x.shape, [x], lambda _p: [], True,
             {"SrcT": AttrValue.of_type(x.scalarType), "DstT": AttrValue.of_type(dtype)})
    return n


def _reduce_shape(s: Shape, idx: Sequence[int]) -> Shape:
    assert s.numDims >= len(idx)
    if not idx:
        return Shape.empty()
    return Shape([d for i, d in enumerate(s.dims) if i not in idx])


def _reducer(op_name: str, parent: Node, reduction_indices: Optional[Sequence[int]], name: Optional[str] = None) -> Node:
    """Sum/Min/Max with a Const 'reduction_indices' second input (dsl/DslImpl.scala:175-188).

    `None` reduces every axis, as python TF's tf.reduce_sum(x) does (it sends range(rank))."""
    if reduction_indices is None:
        if parent.shape.numDims == 0:
            idx = []
        else:
            idx = list(range(parent.shape.numDims))
    else:
        idx = [int(i) for i in reduction_indices]

    def internal(_p: str) -> List[Node]:
        parent.freeze()
        return [constant(np.asarray(idx, dtype=np.int32)).named(parent.name + "/reduction_indices")]

    return Node(name, op_name, parent.scalarType, _reduce_shape(parent.shape, idx), [parent], internal, True,
                {"Tidx": AttrValue.of_type(IntegerType), "keep_dims": AttrValue.of_bool(False)})


def reduce_sum(input_tensor: Node, reduction_indices: Optional[Sequence[int]] = None) -> Node:
    return _reducer("Sum", input_tensor, reduction_indices)


def reduce_min(input_tensor: Node, reduction_indices: Optional[Sequence[int]] = None) -> Node:
    return _reducer("Min", input_tensor, reduction_indices)


def reduce_max(input_tensor: Node, reduction_indices: Optional[Sequence[int]] = None) -> Node:
    return _reducer("Max", input_tensor, reduction_indices)


def matmul(a: Node, b: Node, transpose_a: bool = False, transpose_b: bool = False) -> Node:
    assert a.shape.numDims == 2 and b.shape.numDims == 2, (a.shape, b.shape)
    m = a.shape.dims[1] if transpose_a else a.shape.dims[0]
    n = b.shape.dims[0] if transpose_b else b.shape.dims[1]
    return _build("MatMul", parents=[a, b], shape=Shape(m, n),
                  extra_attrs={"transpose_a": AttrValue.of_bool(transpose_a),
                               "transpose_b": AttrValue.of_bool(transpose_b)})


# --- the k-means / geometric-mean snippets' vocabulary (tensorframes_snippets/kmeans_demo.py, geom_mean.py) ---------
def square(x: Node) -> Node: return _build("Square", parents=[x])
def reciprocal(x: Node) -> Node: return _build("Reciprocal", parents=[x])
def inv(x: Node) -> Node: return _build("Inv", parents=[x])            # tf.inv, the TF 1.0 name geom_mean.py uses
def sqrt(x: Node) -> Node: return _build("Sqrt", parents=[x])
def abs_(x: Node) -> Node: return _build("Abs", parents=[x])
def ones_like(x: Node) -> Node: return _build("OnesLike", parents=[x])
def zeros_like(x: Node) -> Node: return _build("ZerosLike", parents=[x])


def shape(x: Node) -> Node:
    """tf.shape: the int32 vector of x's dimensions, known once the block's shape is."""
    n = Node(None, "Shape", IntegerType, Shape(x.shape.numDims), [x], lambda _p: [], True,
             {"T": AttrValue.of_type(x.scalarType), "out_type": AttrValue.of_type(IntegerType)})
    n.ivalue = [int(d) for d in x.shape.dims]
    return n


def strided_index(x: Node, index: int) -> Node:
    assert x.shape.numDims == 1 and x.scalarType in (IntegerType, LongType), "only integer vectors can be indexed"
    i = int(index)

    def internal(p: str) -> List[Node]:
        return [constant(np.asarray([i], dtype=np.int32)).named(p + "/stack"),
                constant(np.asarray([i + 1], dtype=np.int32)).named(p + "/stack_1"),
                constant(np.asarray([1], dtype=np.int32)).named(p + "/stack_2")]

    n = Node(None, "StridedSlice", x.scalarType, Shape.empty(), [x], internal, True,
             {"Index": AttrValue.of_type(IntegerType), "begin_mask": AttrValue.of_int(0), "end_mask": AttrValue.of_int(0),
              "ellipsis_mask": AttrValue.of_int(0), "new_axis_mask": AttrValue.of_int(0),
              "shrink_axis_mask": AttrValue.of_int(1)})
    if x.ivalue is not None and -len(x.ivalue) <= i < len(x.ivalue):
        n.ivalue = [x.ivalue[i]]
    return n


def stack(values: Sequence) -> Node:
    """tf.stack of integer scalars (python ints become int32 Consts): a Pack along axis 0."""
    lifted = [v if isinstance(v, Node) else constant(np.int32(v)) for v in values]
    assert lifted and all(v.shape.numDims == 0 for v in lifted), "only scalars can be stacked"
    n = _build("Pack", parents=lifted, shape=Shape(len(lifted)),
               extra_attrs={"N": AttrValue.of_int(len(lifted)), "axis": AttrValue.of_int(0)})
    if all(v.ivalue is not None and len(v.ivalue) == 1 for v in lifted):
        n.ivalue = [v.ivalue[0] for v in lifted]
    return n


def expand_dims(x: Node, dim: int) -> Node:
    r = x.shape.numDims
    ax = dim + r + 1 if dim < 0 else dim
    assert 0 <= ax <= r, (dim, x.shape)
    dims = list(x.shape.dims)
    dims.insert(ax, 1)

    def internal(p: str) -> List[Node]:
        return [constant(np.int32(dim)).named(p + "/dim")]

    return Node(None, "ExpandDims", x.scalarType, Shape(dims), [x], internal, True, {"Tdim": AttrValue.of_type(IntegerType)})


def tile(x: Node, multiples) -> Node:
    m = multiples if isinstance(multiples, Node) else constant(np.asarray(list(multiples), dtype=np.int32))
    assert m.shape.numDims == 1
    known = m.ivalue if m.ivalue is not None and len(m.ivalue) == x.shape.numDims else [Unknown] * x.shape.numDims
    dims = [Unknown if (k == Unknown or d == Unknown) else k * d for k, d in zip(known, x.shape.dims)]
    return _build("Tile", parents=[x, m], shape=Shape(dims), dtype=x.scalarType,
                  extra_attrs={"Tmultiples": AttrValue.of_type(m.scalarType)})


def _arg_reduce(op_name: str, x: Node, dimension: int) -> Node:
    r = x.shape.numDims
    ax = dimension + r if dimension < 0 else dimension
    assert 0 <= ax < r, (dimension, x.shape)

    def internal(p: str) -> List[Node]:
        return [constant(np.int32(dimension)).named(p + "/dimension")]

    return Node(None, op_name, LongType, Shape([d for i, d in enumerate(x.shape.dims) if i != ax]), [x], internal, True,
                {"T": AttrValue.of_type(x.scalarType), "Tidx": AttrValue.of_type(IntegerType),
                 "output_type": AttrValue.of_type(LongType)})


def argmin(x: Node, dimension: int) -> Node: return _arg_reduce("ArgMin", x, dimension)
def argmax(x: Node, dimension: int) -> Node: return _arg_reduce("ArgMax", x, dimension)


def unsorted_segment_sum(data: Node, segment_ids: Node, num_segments) -> Node:
    ns = num_segments if isinstance(num_segments, Node) else constant(np.int32(num_segments))
    assert segment_ids.shape.numDims == 1 and ns.shape.numDims == 0
    k = ns.ivalue[0] if ns.ivalue else Unknown
    return _build("UnsortedSegmentSum", parents=[data, segment_ids, ns], dtype=data.scalarType,
                  shape=Shape([k] + list(data.shape.dims[1:])),
                  extra_attrs={"Tindices": AttrValue.of_type(segment_ids.scalarType),
                               "Tnumsegments": AttrValue.of_type(ns.scalarType)})


def _extract_placeholder(df, col_name: str, tf_name: Optional[str], block_: bool,
                         unknown_lead: bool = False) -> Node:
    """Placeholder typed and shaped from a column's metadata (dsl/DslImpl.scala:90-107)."""
    from .frame import ColumnInformation
    field = next((f for f in df.schema.fields if f.name == col_name), None)
    if field is None:
        raise Exception(f"Cannot find column {col_name}, available columns are {','.join(df.schema.fieldNames)}")
    stf = ColumnInformation.of(field).stf
    if stf is None:
        raise Exception(f"The datatype of column '{col_name}' could not be understood by tensorframes: {field}")
    shape = stf.shape if block_ else stf.shape.tail
    if block_ and unknown_lead and shape.numDims > 0:
        shape = shape.tail.prepend(Unknown)
    return placeholder(stf.dataType, *shape.dims).named(tf_name if tf_name is not None else col_name)


def block(df, col_name: str, tf_name: Optional[str] = None) -> Node:
    return _extract_placeholder(df, col_name, tf_name, True)


def row(df, col_name: str, tf_name: Optional[str] = None) -> Node:
    return _extract_placeholder(df, col_name, tf_name, False)


# --- graph assembly (dsl/DslImpl.scala:38-75, Operation.scala:166-175) ------------
def build_graph(*nodes: Node) -> bytes:
    ns: List[Node] = []
    for n in nodes:
        ns.extend(n if isinstance(n, (list, tuple)) else [n])
    for n in ns:
        n.freeze()
    for n in ns:
        n.freeze(everything=True)
    treated: Dict[str, Node] = {}

    def closure(node: Node) -> None:
        if node.name in treated:
            return
        treated[node.name] = node
        for p in node.parents:
            closure(p)

    for n in ns:
        closure(n)
    defs: List[bytes] = []
    emitted = set()
    for n in treated.values():
        for d in n.node_defs():
            if d not in emitted:
                emitted.add(d)
                defs.append(d)
    return proto.graph_def(defs)


def graph_inputs(nodes: Sequence[Node]) -> List[str]:
    """Names of the Placeholder nodes reachable from `nodes` (test/DslOperations.graphInputs)."""
    seen, out = set(), []

    def walk(n: Node):
        if id(n) in seen:
            return
        seen.add(id(n))
        if n.opName == "Placeholder" and not n.parents:
            out.append(n.name)
        for p in n._all_parents():
            walk(p)

    for n in nodes:
        walk(n)
    return out


def hints(nodes: Sequence[Node]) -> ShapeDescription:
    """Shape hints for the fetches (dsl/Operation.scala:166-175)."""
    for n in nodes:
        n.freeze()
    for n in nodes:
        n.freeze(everything=True)
    ins = graph_inputs(nodes)
    out = {n.name: n.shape for n in nodes}
    # like the python front end (core.py:52-72), also send the shape of every placeholder: TF drops the
    # shape attribute of scalar placeholders, so the hint is the only place it survives
    seen = set()

    def walk(n: Node):
        if id(n) in seen:
            return
        seen.add(id(n))
        if n.opName == "Placeholder" and not n.parents:
            out.setdefault(n.name, n.shape)
        for p in n._all_parents():
            walk(p)

    for n in nodes:
        walk(n)
    return ShapeDescription(out, tuple(n.name for n in nodes), {i: i for i in ins})
