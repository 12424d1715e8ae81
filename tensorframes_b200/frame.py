"""A small partitioned, columnar DataFrame: the stand-in for Spark's DataFrame.

The reference runs on Spark SQL DataFrames (Row objects in RDD partitions).  Spark
is outside the hot path and absent here, so this module carries just what the
block executors need from it: a schema with per-column metadata, partitions, and
the tensor metadata conventions of the reference --

  * metadata keys `org.spartf.shape` / `org.sparktf.type`
    (reference: src/main/scala/org/tensorframes/MetadataConstants.scala:19,27)
  * ColumnInformation: read tensor info from metadata, else guess it from the SQL
    type (ColumnInformation.scala:46-58,124-138); write it back (`merged`, :16-26);
    build an output field (`structField`, :80-92)
  * analyze / deepAnalyzeDataFrame (ExperimentalOperations.scala:35-47,68-157)
  * explain (impl/DebugRowOps.scala:528-545)

Storage is columnar from the start: a partition maps each column name to either a
dense numpy array of shape [n_rows, cell dims...] or, for ragged / non-numeric
cells, a python list with one entry per row.  Dense numeric columns are already
the row-major blocks TFDataOps.convert has to assemble from Row objects
(impl/TFDataOps.scala:27-59), so feeding the device needs no per-row work.
"""
from __future__ import annotations

from dataclasses import dataclass, field as dc_field, replace
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from .shape import (SUPPORTED_TYPES, DoubleType, FloatType, IntegerType, LongType, ScalarType, Shape, Unknown,
                    scalar_type_from_name, scalar_type_from_numpy)

SHAPE_KEY = "org.spartf.shape"
TYPE_KEY = "org.sparktf.type"


# ---- SQL types -----------------------------------------------------------------
@dataclass(frozen=True)
class StringType_:
    name: str = "StringType"
    typeName: str = "string"

    def __repr__(self):
        return self.name


StringType = StringType_()


@dataclass(frozen=True)
class ArrayType:
    elementType: Any
    containsNull: bool = False
    typeName: str = "array"

    def __repr__(self):
        return f"ArrayType({self.elementType},{str(self.containsNull).lower()})"


def sql_type(scalar: ScalarType, cell_shape: Shape):
    """Scalar or nested ArrayType(containsNull=false) by cell rank (ColumnInformation.scala:86-92)."""
    t: Any = scalar
    for _ in range(cell_shape.numDims):
        t = ArrayType(t, False)
    return t


def _base_type(dt) -> Optional[ScalarType]:
    while isinstance(dt, ArrayType):
        dt = dt.elementType
    return dt if isinstance(dt, ScalarType) else None


@dataclass(frozen=True)
class StructField:
    name: str
    dataType: Any
    nullable: bool = False
    metadata: Dict[str, Any] = dc_field(default_factory=dict, compare=False)


@dataclass(frozen=True)
class StructType:
    fields: Tuple[StructField, ...]

    @property
    def fieldNames(self) -> List[str]:
        return [f.name for f in self.fields]

    def __getitem__(self, name: str) -> StructField:
        for f in self.fields:
            if f.name == name:
                return f
        raise KeyError(name)

    def __iter__(self):
        return iter(self.fields)

    def __len__(self):
        return len(self.fields)


@dataclass(frozen=True)
class SparkTFColInfo:
    """Tensor info of a column: block shape (lead dim = rows) and scalar type (Shape.scala:120-123)."""
    shape: Shape
    dataType: ScalarType


class ColumnInformation:
    def __init__(self, field: StructField, stf: Optional[SparkTFColInfo]):
        self.field = field
        self.stf = stf

    @property
    def columnName(self) -> str:
        return self.field.name

    @staticmethod
    def of(field: StructField) -> "ColumnInformation":
        """Metadata first, else derive [?, ?...] from the SQL type (ColumnInformation.scala:46-58)."""
        meta = field.metadata
        stf = None
        if SHAPE_KEY in meta and TYPE_KEY in meta:
            t = scalar_type_from_name(str(meta[TYPE_KEY]))
            if t is not None:
                stf = SparkTFColInfo(Shape(list(meta[SHAPE_KEY])), t)
        if stf is None:
            stf = ColumnInformation._from_sql_type(field.dataType)
        return ColumnInformation(field, stf)

    @staticmethod
    def _from_sql_type(dt) -> Optional[SparkTFColInfo]:
        if isinstance(dt, ArrayType):
            inner = ColumnInformation._from_sql_type(dt.elementType)
            return None if inner is None else SparkTFColInfo(inner.shape.prepend(Unknown), inner.dataType)
        if isinstance(dt, ScalarType):
            return SparkTFColInfo(Shape(Unknown), dt)
        return None

    @property
    def merged(self) -> StructField:
        """The field with the tensor info written into its metadata (ColumnInformation.scala:16-26)."""
        meta = dict(self.field.metadata)
        if self.stf is not None:
            meta[SHAPE_KEY] = list(self.stf.shape.dims)
            meta[TYPE_KEY] = self.stf.dataType.name
        return replace(self.field, metadata=meta)

    @staticmethod
    def structField(name: str, scalar: ScalarType, block_shape: Shape) -> StructField:
        """Output column: non-nullable, SQL type from the cell rank, metadata attached (:80-84)."""
        f = StructField(name, sql_type(scalar, block_shape.tail), nullable=False)
        return ColumnInformation(f, SparkTFColInfo(block_shape, scalar)).merged


# ---- rows ------------------------------------------------------------------------
class Row(tuple):
    """Positional row with optional field names, comparable to plain tuples."""

    def __new__(cls, *values, **named):
        if named:
            assert not values
            r = super().__new__(cls, tuple(named.values()))
            r._fields = tuple(named.keys())
        else:
            r = super().__new__(cls, values)
            r._fields = None
        return r

    @staticmethod
    def of(fields: Sequence[str], values: Sequence[Any]) -> "Row":
        r = tuple.__new__(Row, tuple(values))
        r._fields = tuple(fields)
        return r

    def __getattr__(self, item):
        f = tuple.__getattribute__(self, "_fields")
        if f and item in f:
            return self[f.index(item)]
        raise AttributeError(item)

    def __getitem__(self, item):
        if isinstance(item, str):
            return self[self._fields.index(item)]
        return tuple.__getitem__(self, item)

    def __repr__(self):
        if self._fields:
            return "Row(" + ", ".join(f"{k}={v!r}" for k, v in zip(self._fields, self)) + ")"
        return "Row(" + ", ".join(repr(v) for v in self) + ")"


def _to_python(v):
    if isinstance(v, np.ndarray):
        return v.tolist()
    if isinstance(v, np.generic):
        return v.item()
    return v


# ---- column storage ----------------------------------------------------------------
Column = Union[np.ndarray, list]


def is_device_column(c) -> bool:
    """A dense block that lives in GPU memory (a torch CUDA tensor): TensorFrame.to_device puts columns there and
    map_blocks keeps its outputs there, so a chain of operators does not cross PCIe between steps."""
    return type(c).__module__.split(".")[0] == "torch" and bool(getattr(c, "is_cuda", False))


def _column_len(c: Column) -> int:
    return int(c.shape[0]) if (isinstance(c, np.ndarray) or is_device_column(c)) else len(c)


def _infer_sql_type(value):
    if isinstance(value, (list, tuple, np.ndarray)):
        if len(value) == 0:
            return ArrayType(DoubleType, False)
        return ArrayType(_infer_sql_type(value[0]), False)
    if isinstance(value, np.generic):
        t = scalar_type_from_numpy(value.dtype)
        if t is not None:
            return t
    if isinstance(value, bool):
        raise TypeError("boolean columns are not supported")
    if isinstance(value, float):
        return DoubleType
    if isinstance(value, int):
        return LongType  # what PySpark infers for python ints
    if isinstance(value, str):
        return StringType
    raise TypeError(f"cannot infer a column type from {type(value)}")


def _densify(values: list, dt) -> Column:
    """A dense [n, cell...] array when all cells are numeric and equally shaped, else the list."""
    base = _base_type(dt)
    if base is None:
        return list(values)
    try:
        arr = np.asarray(values, dtype=base.np_dtype)
    except (ValueError, TypeError):
        return list(values)
    rank = 1
    t = dt
    while isinstance(t, ArrayType):
        rank += 1
        t = t.elementType
    if arr.ndim != rank:
        return list(values)
    return np.ascontiguousarray(arr)


class TensorFrame:
    """Schema + partitions of columns."""

    def __init__(self, schema: StructType, partitions: List[Dict[str, Column]]):
        self.schema = schema
        self.partitions = partitions

    # -- construction --
    @staticmethod
    def from_rows(rows: Sequence[Any], names: Sequence[str], num_partitions: int = 1,
                  types: Optional[Sequence[Any]] = None) -> "TensorFrame":
        """Rows (tuples / scalars for one column) split into `num_partitions` contiguous slices."""
        names = list(names)
        norm = []
        for r in rows:
            if len(names) == 1 and not (isinstance(r, tuple) and len(r) == 1):
                r = (r,)
            assert len(r) == len(names), (r, names)
            norm.append(tuple(r))
        if types is None:
            if not norm:
                raise ValueError("cannot infer a schema from zero rows; pass types=")
            types = [_infer_sql_type(v) for v in norm[0]]
        schema = StructType(tuple(StructField(n, t, nullable=False) for n, t in zip(names, types)))
        n = len(norm)
        parts = []
        for p in range(num_partitions):
            lo, hi = (n * p) // num_partitions, (n * (p + 1)) // num_partitions
            part = {}
            for ci, (name, t) in enumerate(zip(names, types)):
                part[name] = _densify([r[ci] for r in norm[lo:hi]], t)
            parts.append(part)
        return TensorFrame(schema, parts)

    @staticmethod
    def from_columns(columns: Dict[str, np.ndarray], num_partitions: int = 1) -> "TensorFrame":
        """Dense numpy columns ([n] or [n, cell...]); rows are split contiguously."""
        fields, n = [], None
        for name, arr in columns.items():
            arr = np.asarray(arr)
            t = scalar_type_from_numpy(arr.dtype)
            if t is None:
                raise TypeError(f"column {name}: dtype {arr.dtype} is not supported")
            n = arr.shape[0] if n is None else n
            assert arr.shape[0] == n, "columns differ in length"
            fields.append(StructField(name, sql_type(t, Shape(arr.shape[1:])), nullable=False))
        parts = []
        for p in range(num_partitions):
            lo, hi = (n * p) // num_partitions, (n * (p + 1)) // num_partitions
            parts.append({name: np.ascontiguousarray(np.asarray(arr)[lo:hi]) for name, arr in columns.items()})
        return TensorFrame(StructType(tuple(fields)), parts)

    @staticmethod
    def from_arrow(data, num_partitions: Optional[int] = None) -> "TensorFrame":
        """Arrow table / record batches -> frame, without copying numeric columns.

        One record batch becomes one partition (Arrow chunk <-> Spark partition), unless `num_partitions`
        asks for a contiguous re-split.  Primitive float64/float32/int32/int64 arrays map to scalar columns,
        (nested) FixedSizeList<primitive> arrays to [n, cell...] blocks viewed over Arrow's own value buffer --
        already the row-major packing TFDataOps.convert produces (impl/TFDataOps.scala:27-59) -- and List<>
        arrays to blocks when every row has the same length, else to ragged per-row cells.  String columns
        (group keys) become Python lists.  Nulls are rejected: the reference's tensors cannot hold them
        (all its columns are declared nullable = false)."""
        import pyarrow as pa
        if isinstance(data, pa.Table):
            batches, arrow_schema = data.to_batches(), data.schema
        elif isinstance(data, pa.RecordBatch):
            batches, arrow_schema = [data], data.schema
        else:
            batches = list(data)
            if not batches:
                raise ValueError("from_arrow needs at least one record batch (or a table, which carries its schema)")
            arrow_schema = batches[0].schema
        if num_partitions:   # contiguous, zero-copy re-split of the combined table
            tbl = pa.Table.from_batches(batches, schema=arrow_schema).combine_chunks()
            empty = pa.RecordBatch.from_arrays([pa.array([], type=f.type) for f in arrow_schema], schema=arrow_schema)
            batches = []
            for p in range(num_partitions):
                lo, hi = (tbl.num_rows * p) // num_partitions, (tbl.num_rows * (p + 1)) // num_partitions
                piece = tbl.slice(lo, hi - lo).to_batches() if hi > lo else []
                batches.append(piece[0] if piece else empty)

        def sql_of(t):
            if pa.types.is_fixed_size_list(t) or pa.types.is_list(t) or pa.types.is_large_list(t):
                return ArrayType(sql_of(t.value_type), False)
            if pa.types.is_string(t) or pa.types.is_large_string(t):
                return StringType
            try:
                st = scalar_type_from_numpy(np.dtype(t.to_pandas_dtype()))
            except (NotImplementedError, TypeError):
                st = None
            if st is None:
                raise TypeError(f"Arrow type {t} is not supported (float64, float32, int32, int64, lists of them, string)")
            return st

        def column_of(arr, name):
            if arr.null_count:
                raise ValueError(f"column {name} holds nulls; tensor columns cannot")
            t = arr.type
            if pa.types.is_string(t) or pa.types.is_large_string(t):
                return arr.to_pylist()
            dims = []
            while pa.types.is_fixed_size_list(t):
                dims.append(t.list_size)
                arr = arr.flatten()
                t = arr.type
                if arr.null_count:
                    raise ValueError(f"column {name} holds nulls; tensor columns cannot")
            if pa.types.is_list(t) or pa.types.is_large_list(t):
                if dims:
                    raise TypeError(f"column {name}: variable-length lists inside fixed-size lists are not supported")
                offsets = arr.offsets.to_numpy(zero_copy_only=True)
                values = arr.values
                if pa.types.is_list(values.type) or pa.types.is_fixed_size_list(values.type) or values.null_count:
                    return [np.asarray(v) for v in arr.to_pylist()]
                flat = values.to_numpy(zero_copy_only=True)
                lens = np.diff(offsets)
                if len(lens) and np.all(lens == lens[0]):
                    return flat[offsets[0]:offsets[-1]].reshape(len(lens), int(lens[0]))
                return [flat[offsets[i]:offsets[i + 1]] for i in range(len(lens))]
            flat = arr.to_numpy(zero_copy_only=True)
            n = len(flat) // int(np.prod(dims, dtype=np.int64)) if dims else len(flat)
            return flat.reshape((n, *dims)) if dims else flat

        def field_of(f):
            # a FixedSizeList type states the cell shape: such columns arrive analysed (metadata [?, dims...])
            t, dims = f.type, []
            while pa.types.is_fixed_size_list(t):
                dims.append(t.list_size)
                t = t.value_type
            if dims and not (pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_string(t) or pa.types.is_large_string(t)):
                return ColumnInformation.structField(f.name, sql_of(t), Shape([Unknown] + dims))
            return StructField(f.name, sql_of(f.type), nullable=False)

        fields = tuple(field_of(f) for f in arrow_schema)
        parts = [{f.name: column_of(b.column(i), f.name) for i, f in enumerate(arrow_schema)} for b in batches]
        return TensorFrame(StructType(fields), parts)

    def to_arrow(self):
        """The frame as an Arrow table, one record batch per partition; dense [n, cell...] blocks become nested
        FixedSizeList columns over the same memory, ragged cells List columns."""
        import pyarrow as pa

        def array_of(col, dt):
            if isinstance(col, np.ndarray) and col.dtype != object:
                flat = pa.array(np.ascontiguousarray(col).reshape(-1))
                for d in reversed(col.shape[1:]):
                    flat = pa.FixedSizeListArray.from_arrays(flat, int(d))
                return flat
            return pa.array([_to_python(v) for v in col])

        names = self.schema.fieldNames
        batches = [pa.RecordBatch.from_arrays([array_of(part[n], self.schema[n].dataType) for n in names], names=names)
                   for part in self.partitions]
        return pa.Table.from_batches(batches) if batches else pa.table({n: [] for n in names})

    # -- basic relational surface --
    @property
    def columns(self) -> List[str]:
        return self.schema.fieldNames

    def count(self) -> int:
        return sum(self._part_len(p) for p in self.partitions)

    def _part_len(self, part: Dict[str, Column]) -> int:
        for name in self.schema.fieldNames:
            return _column_len(part[name])
        return 0

    def select(self, *names: str) -> "TensorFrame":
        fields = tuple(self.schema[n] for n in names)
        return TensorFrame(StructType(fields), [{n: p[n] for n in names} for p in self.partitions])

    def withColumnRenamed(self, old: str, new: str) -> "TensorFrame":
        fields = tuple(replace(f, name=new) if f.name == old else f for f in self.schema.fields)
        parts = [{(new if k == old else k): v for k, v in p.items()} for p in self.partitions]
        return TensorFrame(StructType(fields), parts)

    def withColumn(self, name: str, source: str) -> "TensorFrame":
        """Alias of an existing column under a new name (df.select(df.y, df.y.alias("z")))."""
        f = replace(self.schema[source], name=name)
        return TensorFrame(StructType(self.schema.fields + (f,)),
                           [dict(p, **{name: p[source]}) for p in self.partitions])

    def repartition(self, n: int) -> "TensorFrame":
        """Round-robin redistribution, like Spark's repartition: sizes differ by at most one."""
        cols = {name: [] for name in self.columns}
        for p in self.partitions:
            for name in self.columns:
                c = p[name]
                cols[name].extend(list(c) if not isinstance(c, np.ndarray) else [c[i] for i in range(c.shape[0])])
        total = len(next(iter(cols.values()))) if cols else 0
        parts = []
        for k in range(n):
            idx = list(range(k, total, n))
            part = {}
            for f in self.schema.fields:
                part[f.name] = _densify([cols[f.name][i] for i in idx], f.dataType)
            parts.append(part)
        return TensorFrame(self.schema, parts)

    def to_device(self, device: int = 0) -> "TensorFrame":
        """Dense numeric columns move to GPU memory (torch owns the allocations); ragged and string columns stay.
        Operators keep device-resident inputs and outputs where they are: 180 GB of HBM hold a lot of partitions."""
        import torch
        parts = [{n: (torch.from_numpy(np.ascontiguousarray(c)).to(f"cuda:{device}")
                      if isinstance(c, np.ndarray) and scalar_type_from_numpy(c.dtype) is not None else c)
                  for n, c in p.items()} for p in self.partitions]
        return TensorFrame(self.schema, parts)

    def to_host(self) -> "TensorFrame":
        parts = [{n: (c.cpu().numpy() if is_device_column(c) else c) for n, c in p.items()} for p in self.partitions]
        return TensorFrame(self.schema, parts)

    def collect(self) -> List[Row]:
        names = self.columns
        out = []
        for p in self.partitions:
            n = self._part_len(p)
            cols = [p[name].cpu().numpy() if is_device_column(p[name]) else p[name] for name in names]
            for i in range(n):
                out.append(Row.of(names, [_to_python(c[i]) for c in cols]))
        return out

    def first(self) -> Row:
        return self.collect()[0]

    def groupBy(self, *cols: str) -> "GroupedData":
        for c in cols:
            self.schema[c]
        return GroupedData(self, list(cols))

    def printSchema(self) -> None:
        print(explain(self))

    # reference-style method sugar (dsl/Implicits.scala:25-116)
    def analyze(self) -> "TensorFrame":
        return analyze(self)

    def block(self, col_name: str, tf_name: Optional[str] = None):
        from . import dsl
        return dsl.block(self, col_name, tf_name)

    def row(self, col_name: str, tf_name: Optional[str] = None):
        from . import dsl
        return dsl.row(self, col_name, tf_name)

    def mapBlocks(self, *fetches):
        from . import ops
        return ops.default_ops().mapBlocks(self, *fetches)

    def mapBlocksTrimmed(self, *fetches):
        from . import ops
        return ops.default_ops().mapBlocksTrimmed(self, *fetches)

    def reduceBlocks(self, *fetches):
        from . import ops
        return ops.default_ops().reduceBlocks(self, *fetches)

    def mapRows(self, *fetches, feed_dict=None):
        from . import ops
        return ops.default_ops().mapRows(self, *fetches, feed_dict=feed_dict)

    def reduceRows(self, *fetches):
        from . import ops
        return ops.default_ops().reduceRows(self, *fetches)


class GroupedData:
    def __init__(self, df: TensorFrame, keys: List[str]):
        self.df = df
        self.keys = keys

    def aggregate(self, *fetches):
        from . import ops
        return ops.default_ops().aggregate(self, *fetches)


# ---- analyze (ExperimentalOperations.scala:35-157) -------------------------------------
def _merge(s1: Optional[Shape], s2: Optional[Shape]) -> Optional[Shape]:
    if s1 is None or s2 is None or s1.numDims != s2.numDims:
        return None
    return Shape([a if a == b else Unknown for a, b in zip(s1.dims, s2.dims)])


def analyze_data(x) -> Optional[Shape]:
    """Shape of one cell (ExperimentalOperations.scala:119-131)."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return Shape(x.shape) if scalar_type_from_numpy(x.dtype) is not None else None
    if isinstance(x, (list, tuple)):
        if len(x) == 0:
            return Shape(0)
        merged = analyze_data(x[0])
        for e in x[1:]:
            merged = _merge(merged, analyze_data(e))
        return None if merged is None else merged.prepend(len(x))
    if isinstance(x, bool):
        return None
    if isinstance(x, (int, float, np.integer, np.floating)):
        return Shape.empty()
    return None


def _partition_shape(col: Column) -> Optional[Shape]:
    """Block shape of one partition's column: merged cell shape with the row count in front."""
    n = _column_len(col)
    if is_device_column(col):
        return Shape(tuple(col.shape))
    if isinstance(col, np.ndarray):
        return Shape(col.shape) if scalar_type_from_numpy(col.dtype) is not None else None
    merged = analyze_data(col[0])
    for cell in col[1:]:
        merged = _merge(merged, analyze_data(cell))
    return None if merged is None else merged.prepend(n)


def analyze(df: TensorFrame) -> TensorFrame:
    """One pass over the data computing each column's block shape, merged over partitions
    ("equal -> keep, different -> Unknown"), written to the column metadata.  Empty partitions
    do not take part (ExperimentalOperations.scala:68-111)."""
    per_col: Dict[str, Optional[Shape]] = {}
    seen = False
    for p in df.partitions:
        if df._part_len(p) == 0:
            continue
        for f in df.schema.fields:
            s = _partition_shape(p[f.name])
            per_col[f.name] = s if not seen else _merge(per_col[f.name], s)
        seen = True
    fields = []
    for f in df.schema.fields:
        shape = per_col.get(f.name) if seen else None
        base = _base_type(f.dataType)
        if shape is not None and base is not None:
            fields.append(ColumnInformation(f, SparkTFColInfo(shape, base)).merged)
        else:
            fields.append(f)
    return TensorFrame(StructType(tuple(fields)), df.partitions)


def explain(df: TensorFrame) -> str:
    """Schema with tensor info, in the reference's text format (impl/DebugRowOps.scala:528-545)."""
    out = ["root"]
    for f in df.schema.fields:
        ci = ColumnInformation.of(f)
        line = f" |-- {f.name}: {f.dataType.typeName} (nullable = {str(f.nullable).lower()})"
        if ci.stf is not None:
            line += f" {ci.stf.dataType.typeName}{ci.stf.shape}"
        else:
            line += " <no tensor info>"
        out.append(line)
    return "\n".join(out) + "\n"
