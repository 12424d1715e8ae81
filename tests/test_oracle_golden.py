"""Pins the oracle: every known-answer vector of the reference must come out of the oracle's
restatement exactly.  CPU only."""
import numpy as np
import pytest

import _golden
from oracle import executors as ox
from oracle import graph_writer, graphdef, tf_ops
from tensorframes_b200 import dsl

G = _golden.load()


def _case_graph(case, df, cols, builder):
    """(GraphDef bytes, {fetch: (dtype code, dims)}, fetch names) for a case, from one of two independent sources:
    `writer` = oracle/graph_writer.py (the reference's node layout restated, no product code), `dsl` = the product's
    DSL.  The oracle is pinned with the first; the second keeps checking the product's encoder against it."""
    if builder == "writer":
        columns = {c.name: (c.dtype, tuple(c.shape)) for c in cols if c.dtype is not None}
        graph, info = graph_writer.from_spec(case["graph"], columns)
        return graph, info, list(case["graph"]["fetches"])
    with dsl.withGraph():
        fetches = _golden.build_fetches(case, df)
        graph = dsl.build_graph(*fetches)
        hints = dsl.hints(fetches)
    info = {f.name: (f.scalarType.tf_code, tuple(hints.out[f.name].dims)) for f in fetches}
    return graph, info, [f.name for f in fetches]


@pytest.mark.parametrize("builder", ["writer", "dsl"])
@pytest.mark.parametrize("case", G["cases"], ids=[c["id"] for c in G["cases"]])
def test_executor_case(case, builder):
    df = _golden.build_frame(case)
    parts, cols = _golden.oracle_inputs(case, df)
    graph, info, fetch_names = _case_graph(case, df, cols, builder)
    kind = case["kind"]
    if kind in ("map_blocks", "map_blocks_trimmed"):
        inputs = [n.name for n in graphdef.parse_graph(graph) if n.op == "Placeholder"]
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
        names, out_parts = ox.map_rows(parts, cols, graph, feed, info)
        rows = [r for p in out_parts for r in p]
        idx = [names.index(n) for n in case["select"]]
        rows = [[r[i] for i in idx] for r in rows]
        assert _golden.normalize(rows) == _golden.normalize(case["expected"])
    elif kind == "reduce_rows":
        row = ox.reduce_rows(parts, cols, graph)
        assert list(row) == case["expected"]
    elif kind == "reduce_blocks":
        row = ox.reduce_blocks(parts, cols, graph, fetch_names)
        assert list(row) == case["expected"]
    elif kind == "aggregate":
        key_idx = [[c.name for c in cols].index(k) for k in case["keys"]]
        rows = ox.aggregate(parts, cols, key_idx, graph, fetch_names)
        assert _golden.normalize(rows) == _golden.normalize(case["expected"])
    else:
        raise KeyError(kind)


def _tree(nodes, name):
    """A fetch as a nested tuple that does not depend on automatic node names: (op, dtype, value or shape, inputs)."""
    by_name = {n.name: n for n in nodes}
    n = by_name[name.split(":")[0]]
    dtype = (n.attr.get("T") or n.attr.get("dtype") or (None, None))[1]
    extra = None
    if n.op == "Const":
        v = n.attr["value"][1]
        extra = (tuple(v.shape), tuple(v.reshape(-1).tolist()))
    elif n.op == "Placeholder":
        extra = (n.name, tuple(n.attr["shape"][1]) if "shape" in n.attr and n.attr["shape"][1] is not None else None)
    elif n.op in ("Sum", "Min", "Max"):
        extra = (n.attr["Tidx"][1], n.attr["keep_dims"][1])
    return (n.op, dtype, extra, tuple(_tree(nodes, i) for i in n.inputs))


@pytest.mark.parametrize("case", G["cases"], ids=[c["id"] for c in G["cases"]])
def test_product_dsl_and_independent_writer_encode_the_same_graph(case):
    """Every golden case's graph, built by the product's DSL and by oracle/graph_writer.py, read back with the oracle's
    own parser: same fetches, same ops, dtypes, inputs, constants, reduction attributes, same inferred fetch shapes."""
    df = _golden.build_frame(case)
    parts, cols = _golden.oracle_inputs(case, df)
    g_w, info_w, names_w = _case_graph(case, df, cols, "writer")
    g_d, info_d, names_d = _case_graph(case, df, cols, "dsl")
    assert names_w == names_d
    nodes_w, nodes_d = graphdef.parse_graph(g_w), graphdef.parse_graph(g_d)
    for f in names_w:
        assert _tree(nodes_w, f) == _tree(nodes_d, f)
        assert info_w[f][0] == info_d[f][0] and tuple(info_w[f][1]) == tuple(info_d[f][1])


@pytest.mark.parametrize("case", G["cases"], ids=[c["id"] for c in G["cases"]])
def test_native_importer_reads_the_independent_writer(case):
    """The library's GraphDef reader and static analysis (csrc/graph.cpp, tfs_graph_analyze: no GPU involved) on graphs
    it did not write: same inputs, fetches, dtypes and shapes as on the product DSL's encoding of the case, and -- with
    no hints at all -- the fetch shapes the writer inferred."""
    from tensorframes_b200 import _native
    df = _golden.build_frame(case)
    parts, cols = _golden.oracle_inputs(case, df)
    g_w, info_w, names = _case_graph(case, df, cols, "writer")
    g_d, _info_d, _names_d = _case_graph(case, df, cols, "dsl")
    with _native.Graph(g_w) as gw, _native.Graph(g_d) as gd:
        got = {name: (st.tf_code, tuple(shape.dims), is_ph, is_in, is_out) for name, st, shape, is_ph, is_in, is_out in gw.analyze({}, names)}
        ref = {name: (st.tf_code, tuple(shape.dims), is_ph, is_in, is_out) for name, st, shape, is_ph, is_in, is_out in gd.analyze({}, names)}
    assert got == ref
    for f in names:
        assert got[f][0] == info_w[f][0] and got[f][1] == tuple(info_w[f][1]) and got[f][4]


def test_writer_matches_the_reference_nodedef_vectors():
    """The reference's own NodeDef expectations (T/dsl/BasicSuite.scala:12-33, BasicOpsSuite.scala:12-22: constant(3),
    named constants, duplicate names Const / Const_1, add) hold for the independent writer as well."""
    w = graph_writer.GraphWriter()
    a = w.constant(graph_writer.DT_INT32, 3)
    b = w.constant(graph_writer.DT_INT32, 3)
    c = w.constant(graph_writer.DT_DOUBLE, [1.0, 2.0], name="x")
    s = w.add(w.constant(graph_writer.DT_DOUBLE, [3.0, 4.0]), c, name="z")
    assert (a, b, c, s) == ("Const", "Const_1", "x", "z")
    nodes = {n.name: n for n in graphdef.parse_graph(w.graph())}
    assert nodes["Const"].op == "Const" and nodes["Const"].attr["dtype"][1] == 3 and nodes["Const"].attr["value"][1].tolist() == 3
    assert nodes["Const_1"].attr["value"][1].shape == ()
    assert nodes["z"].op == "Add" and nodes["z"].inputs == ["Const_2", "x"] and nodes["z"].attr["T"][1] == 2
    assert nodes["x"].attr["value"][1].tolist() == [1.0, 2.0]
    (out,) = tf_ops.evaluate(w.graph(), {}, ["z"])
    assert out.tolist() == [4.0, 6.0]


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
