"""Tensor shapes and scalar types, mirroring the reference's metadata model.

Reference: src/main/scala/org/tensorframes/Shape.scala:16-113 (Shape, Unknown = -1),
src/main/scala/org/tensorframes/impl/datatypes.scala:265-324 (the four supported
scalar types and their SQL / TF counterparts).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

Unknown = -1


class Shape:
    """Dimensions of a tensor; -1 marks an unknown dimension (Shape.scala:16-60)."""

    __slots__ = ("dims",)

    def __init__(self, *dims: int):
        if len(dims) == 1 and isinstance(dims[0], (list, tuple)):
            dims = tuple(dims[0])
        for d in dims:
            if int(d) < -1:
                raise ValueError(f"{list(dims)} should not contain values <= -2")
        self.dims: Tuple[int, ...] = tuple(int(d) for d in dims)

    @staticmethod
    def empty() -> "Shape":
        return Shape()

    @property
    def numDims(self) -> int:
        return len(self.dims)

    @property
    def hasUnknown(self) -> bool:
        return Unknown in self.dims

    @property
    def numElements(self) -> Optional[int]:
        if self.hasUnknown:
            return None
        n = 1
        for d in self.dims:
            n *= d
        return n

    def prepend(self, x: int) -> "Shape":
        return Shape((int(x),) + self.dims)

    def dropInner(self) -> "Shape":
        return Shape(self.dims[:-1])

    @property
    def tail(self) -> "Shape":
        if not self.dims:
            raise ValueError("tail of an empty shape")
        return Shape(self.dims[1:])

    def checkMorePreciseThan(self, other: "Shape") -> bool:
        """Same rank, and every dim equal unless `other` leaves it unknown (Shape.scala:54-59)."""
        if len(self.dims) != len(other.dims):
            return False
        return all(b == Unknown or b == a for a, b in zip(self.dims, other.dims))

    def __eq__(self, other) -> bool:
        return isinstance(other, Shape) and other.dims == self.dims

    def __hash__(self) -> int:
        return hash(self.dims)

    def __iter__(self):
        return iter(self.dims)

    def __len__(self) -> int:
        return len(self.dims)

    def __repr__(self) -> str:
        return "[" + ",".join("?" if d == Unknown else str(d) for d in self.dims) + "]"


@dataclass(frozen=True)
class ScalarType:
    """One supported scalar type: SQL name, TF DataType code (types.proto:16-24), numpy dtype."""

    name: str        # 'DoubleType' -- what the reference stores under org.sparktf.type
    typeName: str    # 'double' -- DataType.typeName, used by explain()
    tf_code: int
    np_dtype: np.dtype

    def __repr__(self) -> str:
        return self.name


DoubleType = ScalarType("DoubleType", "double", 2, np.dtype("<f8"))
FloatType = ScalarType("FloatType", "float", 1, np.dtype("<f4"))
IntegerType = ScalarType("IntegerType", "integer", 3, np.dtype("<i4"))
LongType = ScalarType("LongType", "long", 9, np.dtype("<i8"))

SUPPORTED_TYPES = (DoubleType, FloatType, IntegerType, LongType)


def scalar_type_from_numpy(dt) -> Optional[ScalarType]:
    dt = np.dtype(dt)
    for t in SUPPORTED_TYPES:
        if t.np_dtype == dt.newbyteorder("<") or t.np_dtype == dt:
            return t
    return None


def scalar_type_from_tf(code: int) -> ScalarType:
    for t in SUPPORTED_TYPES:
        if t.tf_code == code:
            return t
    raise ValueError(
        f"Type {code} is not supported. Only the following types are supported: "
        + ", ".join(str(t.tf_code) for t in SUPPORTED_TYPES))


def scalar_type_from_name(name: str) -> Optional[ScalarType]:
    for t in SUPPORTED_TYPES:
        if t.name == name:
            return t
    return None


@dataclass(frozen=True)
class ShapeDescription:
    """Shape hints passed next to a graph (ShapeDescription.scala:12-20).

    out: node name -> Shape (overrides what graph analysis infers)
    requestedFetches: names of the nodes to fetch
    inputs: placeholder name -> DataFrame column (only consulted by mapRows)
    """

    out: dict
    requestedFetches: tuple
    inputs: dict

    @staticmethod
    def empty() -> "ShapeDescription":
        return ShapeDescription({}, (), {})
