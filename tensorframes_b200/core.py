"""The python API of the reference, by name: tensorframes/core.py:10-11.

    map_blocks, map_rows, reduce_blocks, reduce_rows, aggregate, analyze, print_schema, block, row

The reference's functions take TensorFlow tensors built in the default python-TF
graph and ship the serialized GraphDef to the JVM (core.py:38-72, 123-135).
TensorFlow is not part of this build; fetches are nodes of `tensorframes_b200.dsl`
(same op vocabulary, same GraphDef bytes), or an explicit `(graph_bytes,
ShapeDescription)` pair for a GraphDef produced elsewhere.
"""
from __future__ import annotations

from typing import Sequence, Union

import numpy as np

from . import dsl
from .frame import GroupedData, TensorFrame, analyze as _analyze, explain as _explain
from .ops import default_ops

__all__ = ["reduce_blocks", "map_blocks", "analyze", "print_schema", "aggregate", "block", "row",
           "map_rows", "reduce_rows"]


def _check_fetches(fetches):
    """core.py:74-78"""
    if not isinstance(fetches, (list, tuple)):
        return [fetches]
    return list(fetches)


def _unique_names(fetches) -> None:
    names = [f.name for f in fetches]
    if len(set(names)) != len(names):
        raise ValueError(f"Could not infer a list of unique names for the columns: {names}")


def map_blocks(fetches, dframe: TensorFrame, trim: bool = False) -> TensorFrame:
    """Transforms a DataFrame block by block, appending one column per fetch (trim=False) or
    keeping only the fetched columns (trim=True) (core.py:213-253)."""
    fetches = _check_fetches(fetches)
    for f in fetches:
        f.freeze()
    _unique_names(fetches)
    ops = default_ops()
    return ops.mapBlocksTrimmed(dframe, *fetches) if trim else ops.mapBlocks(dframe, *fetches)


def reduce_blocks(fetches, dframe: TensorFrame):
    """Block-wise reduction to a single row; fetch `z` reads placeholder `z_input` (core.py:255-291).
    Returns one numpy array (or python scalar) per fetch, a single value if there is one fetch."""
    fetches = _check_fetches(fetches)
    for f in fetches:
        f.freeze()
    _unique_names(fetches)
    row = default_ops().reduceBlocks(dframe, *fetches)

    def unpack(fetch):  # core.py:90-104
        x = row[fetch.name]
        return np.array(x) if fetch.shape.numDims > 0 else x

    out = [unpack(f) for f in fetches]
    return out[0] if len(out) == 1 else out


def aggregate(fetches, grouped_data: GroupedData) -> TensorFrame:
    """Algebraic aggregation over grouped data (core.py:319-336)."""
    fetches = _check_fetches(fetches)
    for f in fetches:
        f.freeze()
    _unique_names(fetches)
    return default_ops().aggregate(grouped_data, *fetches)


def analyze(dframe: TensorFrame) -> TensorFrame:
    """Adds tensor shape metadata by scanning the data (core.py:305-317)."""
    return _analyze(dframe)


def print_schema(dframe: TensorFrame) -> None:
    """core.py:293-303"""
    print(_explain(dframe))


def block(df: TensorFrame, col_name: str, tf_name: str = None):
    """Placeholder for a block of a column; the lead dimension is always left unknown, as the
    python reference does (core.py:338-351, 386-390)."""
    return dsl._extract_placeholder(df, col_name, tf_name, True, unknown_lead=True)


def row(df: TensorFrame, col_name: str, tf_name: str = None):
    """core.py:353-366"""
    return dsl.row(df, col_name, tf_name)


def map_rows(fetches, dframe: TensorFrame, feed_dict=None) -> TensorFrame:
    """Transforms a DataFrame row by row, adding one column per fetch (core.py:175-211).  Placeholders
    have the shape of one cell; `feed_dict` maps a placeholder name to a column name (default: the
    column with the placeholder's name).  Rows may hold vectors of different lengths."""
    fetches = _check_fetches(fetches)
    for f in fetches:
        f.freeze()
    _unique_names(fetches)
    return default_ops().mapRows(dframe, *fetches, feed_dict=feed_dict)


def reduce_rows(fetches, dframe: TensorFrame):
    """Pairwise reduction to a single row: fetch `z` reads placeholders `z_1` and `z_2` (core.py:138-173)."""
    fetches = _check_fetches(fetches)
    for f in fetches:
        f.freeze()
    _unique_names(fetches)
    row = default_ops().reduceRows(dframe, *fetches)
    out = [np.array(row[f.name]) if f.shape.numDims > 0 else row[f.name] for f in fetches]
    return out[0] if len(out) == 1 else out
