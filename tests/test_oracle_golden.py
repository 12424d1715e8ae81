"""Pins the oracle: every known-answer vector of the reference must come out of the oracle's
restatement exactly.  CPU only."""
import numpy as np
import pytest

import _golden
from oracle import executors as ox
from oracle import graphdef, tf_ops
from tensorframes_b200 import dsl

G = _golden.load()


@pytest.mark.parametrize("case", G["cases"], ids=[c["id"] for c in G["cases"]])
def test_executor_case(case):
    df = _golden.build_frame(case)
    with dsl.withGraph():
        fetches = _golden.build_fetches(case, df)
        graph = dsl.build_graph(*fetches)
        hints = dsl.hints(fetches)
    parts, cols = _golden.oracle_inputs(case, df)
    kind = case["kind"]
    if kind in ("map_blocks", "map_blocks_trimmed"):
        inputs = [n.name for n in graphdef.parse_graph(graph) if n.op == "Placeholder"]
        info = {f.name: (f.scalarType.tf_code, tuple(hints.out[f.name].dims)) for f in fetches}
        names, out_parts = ox.map_blocks(parts, cols, graph, inputs, info, append_input=(kind == "map_blocks"))
        rows = [r for p in out_parts for r in p]
        if case.get("expected_columns"):
            assert names == case["expected_columns"]
        if case.get("select"):
            idx = [names.index(n) for n in case["select"]]
            rows = [[r[i] for i in idx] for r in rows]
        assert _golden.normalize(rows) == _golden.normalize(case["expected"])
    elif kind == "map_rows":
        phs = [n.name for n in graphdef.parse_graph(graph) if n.op == "Placeholder"]
        feed = {p: case.get("feed_dict", {}).get(p, p) for p in phs}
        info = {f.name: (f.scalarType.tf_code, tuple(hints.out[f.name].dims)) for f in fetches}
        names, out_parts = ox.map_rows(parts, cols, graph, feed, info)
        rows = [r for p in out_parts for r in p]
        idx = [names.index(n) for n in case["select"]]
        rows = [[r[i] for i in idx] for r in rows]
        assert _golden.normalize(rows) == _golden.normalize(case["expected"])
    elif kind == "reduce_rows":
        row = ox.reduce_rows(parts, cols, graph)
        assert list(row) == case["expected"]
    elif kind == "reduce_blocks":
        row = ox.reduce_blocks(parts, cols, graph, [f.name for f in fetches])
        assert list(row) == case["expected"]
    elif kind == "aggregate":
        key_idx = [[c.name for c in cols].index(k) for k in case["keys"]]
        rows = ox.aggregate(parts, cols, key_idx, graph, [f.name for f in fetches])
        assert _golden.normalize(rows) == _golden.normalize(case["expected"])
    else:
        raise KeyError(kind)


@pytest.mark.parametrize("case", G["analyze"], ids=[c["id"] for c in G["analyze"]])
def test_analyze_case(case):
    df = _golden.build_frame(dict(case, analyze=False))
    parts, cols = _golden.oracle_inputs(case, df)
    if case["analyze"]:
        shapes = ox.analyze(parts, len(cols))
    else:  # un-analysed columns: [?] for scalars, one more ? per array level (ColumnInformation.scala:124-138)
        shapes = [c.shape for c in cols]
    for name, (dt, shape) in case["expected"].items():
        i = [c.name for c in cols].index(name)
        assert list(shapes[i]) == shape


@pytest.mark.parametrize("name", sorted(G["fixtures"]))
def test_graph_fixture(name):
    fx = G["fixtures"][name]
    nodes = graphdef.parse_graph(bytes.fromhex(fx["hex"]))
    assert [n.name for n in nodes] == [n["name"] for n in fx["nodes"]]
    for got, want in zip(nodes, fx["nodes"]):
        assert got.op == want["op"]
        assert got.inputs == want.get("inputs", [])
        if want["op"] == "Const":
            v = got.attr["value"][1]
            assert list(v.shape) == want["shape"] and v.reshape(-1).tolist() == want["values"]
        if want["op"] == "Placeholder":
            assert list(got.attr["shape"][1]) == want["shape"]
            assert graphdef.DT_NAME[got.attr["dtype"][1]] == want["dtype"]


def test_graph2_fixture_evaluates():
    fx = G["fixtures"]["graph2.pb"]
    a = np.arange(4, dtype=np.float32).reshape(2, 2)
    (out,) = tf_ops.evaluate(bytes.fromhex(fx["hex"]), {"z_1": a, "z_2": a + 1}, ["out"])
    assert out.tolist() == (a + a + 1).tolist()


def test_evaluator_semantics():
    """The unpinned ops follow the TF definitions stated in oracle/tf_ops.py."""
    from tensorframes_b200 import DoubleType, IntegerType, FloatType
    with dsl.withGraph():
        x = dsl.placeholder(IntegerType, -1).named("x")
        y = dsl.placeholder(IntegerType, -1).named("y")
        q = dsl.div(x, y).named("q")
        g = dsl.build_graph(q)
    xs = np.array([7, -7, 7, -7], dtype=np.int32)
    ys = np.array([2, 2, -2, -2], dtype=np.int32)
    (out,) = tf_ops.evaluate(g, {"x": xs, "y": ys}, ["q"])
    assert out.tolist() == [3, -3, -3, 3]          # truncation toward zero
    with pytest.raises(tf_ops.OracleError):
        tf_ops.evaluate(g, {"x": xs, "y": np.zeros(4, np.int32)}, ["q"])
    with dsl.withGraph():
        x = dsl.placeholder(DoubleType, -1).named("x")
        c = dsl.cast(dsl.maximum(x * 1.5 + 0.25, 0.0), FloatType).named("c")
        i = dsl.cast(x, IntegerType).named("i")
        g = dsl.build_graph(c, i)
    xs = np.array([-1.0, 0.3, 2.7, -2.7])
    c_out, i_out = tf_ops.evaluate(g, {"x": xs}, ["c", "i"])
    assert c_out.dtype == np.float32 and c_out.tolist() == np.maximum(xs * 1.5 + 0.25, 0).astype(np.float32).tolist()
    assert i_out.tolist() == [-1, 0, 2, -2]
