"""Turns a case of tests/golden/reference_vectors.json into (a) a product TensorFrame + DSL
fetches and (b) the oracle's rows / Col descriptions.  Used by CPU and GPU tests alike."""
import json
import os

import numpy as np

import tensorframes_b200 as tfs
from tensorframes_b200 import dsl
from tensorframes_b200.frame import ArrayType, ColumnInformation, StringType, TensorFrame

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPES = {"double": tfs.DoubleType, "float": tfs.FloatType, "int": tfs.IntegerType, "long": tfs.LongType}
TF_NAME = {1: "DT_FLOAT", 2: "DT_DOUBLE", 3: "DT_INT32", 9: "DT_INT64"}


def load():
    with open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")) as f:
        return json.load(f)


def _depth(v):
    d = 0
    while isinstance(v, list):
        d += 1
        v = v[0]
    return d


def build_frame(case) -> TensorFrame:
    names = [c[0] for c in case["columns"]]
    rows = [tuple(r) for r in case["rows"]]
    types = []
    for ci, (name, dt) in enumerate(case["columns"]):
        if dt == "string":
            types.append(StringType)
            continue
        t = TYPES[dt]
        for _ in range(_depth(rows[0][ci])):
            t = ArrayType(t, False)
        types.append(t)
    df = TensorFrame.from_rows(rows, names, num_partitions=case.get("partitions", 1), types=types)
    if case.get("repartition"):
        df = df.repartition(case["repartition"])
    if case.get("analyze"):
        df = tfs.analyze(df)
    return df


def build_fetches(case, df):
    """DSL nodes for the case's graph spec, inside a fresh graph scope. Returns the fetch nodes."""
    spec = case["graph"]
    env = {}
    for name, dt, shape in spec.get("placeholders", []):
        env[name] = dsl.placeholder(TYPES[dt], *shape).named(name)
    for col, tf_name in spec.get("block_placeholders", []):
        env[tf_name] = tfs.block(df, col, tf_name=tf_name)
    for col, tf_name in spec.get("row_placeholders", []):
        env[tf_name] = tfs.row(df, col, tf_name=tf_name)
    for name, dt, value in spec.get("consts", []):
        fetch = name in spec["fetches"] or name in spec.get("keep_consts", [])
        c = dsl.constant(np.asarray(value, dtype=TYPES[dt].np_dtype))
        env[name] = c.named(name) if fetch else c
    for op in spec.get("ops", []):
        name, kind, args = op[0], op[1], op[2:]
        if kind == "identity":
            n = dsl.identity(env[args[0]])
        elif kind == "add":
            n = dsl.add(env[args[0]], env[args[1]])
        elif kind in ("reduce_sum", "reduce_min", "reduce_max"):
            n = getattr(dsl, kind)(env[args[0]], args[1])
        else:
            raise KeyError(kind)
        env[name] = n.named(name)
    return [env[f] for f in spec["fetches"]]


def oracle_inputs(case, df):
    """(partitions of row tuples, [oracle Col]) matching the product frame's partitioning/metadata."""
    from oracle.executors import Col
    cols = []
    for f in df.schema.fields:
        stf = ColumnInformation.of(f).stf
        cols.append(Col(f.name, None if stf is None else stf.dataType.tf_code,
                        None if stf is None else tuple(stf.shape.dims)))
    parts = [[tuple(r) for r in TensorFrame(df.schema, [p]).collect()] for p in df.partitions]
    return parts, cols


def normalize(rows):
    """compareRows of the reference: order-insensitive comparison (T/TensorFlossTestSparkContext.scala:37-41)."""
    return sorted([list(r) for r in rows], key=lambda r: json.dumps(r))
