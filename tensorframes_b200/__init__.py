"""tensorframes_b200 -- B200-native block executors for TensorFrames' DataFrame tensor path.

Host side (this package): schema/metadata model, graph DSL, validation, the operator
interface of the reference.  Device side: libtfsb200.so (csrc/), reached through the C ABI
declared in include/tfs_b200.h.  There is no CPU execution path.
"""
from .shape import (DoubleType, FloatType, IntegerType, LongType, Shape, ShapeDescription, Unknown)
from .frame import Row, TensorFrame, ArrayType, StringType, StructField, StructType, ColumnInformation
from .core import (aggregate, analyze, block, map_blocks, map_rows, print_schema, reduce_blocks, reduce_rows, row)
from . import dsl

__all__ = [
    "DoubleType", "FloatType", "IntegerType", "LongType", "Shape", "ShapeDescription", "Unknown",
    "Row", "TensorFrame", "ArrayType", "StringType", "StructField", "StructType", "ColumnInformation",
    "aggregate", "analyze", "block", "map_blocks", "map_rows", "print_schema", "reduce_blocks", "reduce_rows",
    "row", "dsl",
]
