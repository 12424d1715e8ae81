"""Host-side logic that needs no GPU: the DSL and its GraphDef bytes, graph analysis through the
C ABI (no CUDA call), schema validation with the reference's error behaviour, analyze/explain."""
import ctypes

import numpy as np
import pytest

import _golden
import tensorframes_b200 as tfs
from oracle import graphdef
from tensorframes_b200 import _native, dsl, ops
from tensorframes_b200.frame import ColumnInformation, TensorFrame, explain
from tensorframes_b200.shape import Shape, ShapeDescription, Unknown

G = _golden.load()


# ---- DSL -> GraphDef, decoded by the oracle's independent reader ------------------------------
def _run_program(program):
    env, out = {}, []
    for step in program:
        kind, name = step[0], step[1]
        if kind == "constant":
            n = dsl.constant(step[2])
        elif kind == "fill":
            n = dsl.fill(step[2], step[3])
        elif kind == "add":
            n = dsl.add(env[step[2]], env[step[3]])
        if name is not None:
            n = n.named(name)
            env[name] = n
        out.append(n)
    return out


@pytest.mark.parametrize("case", G["dsl"], ids=[c["id"] for c in G["dsl"]])
def test_dsl_matches_python_tf_nodedefs(case):
    with dsl.withGraph():
        nodes = _run_program(case["program"])
        last = [nodes[-1]] if case["id"] == "dsl_add" else nodes
        g = dsl.build_graph(*last)
    got = {n.name: n for n in graphdef.parse_graph(g)}
    assert sorted(got) == sorted(w["name"] for w in case["expected"])
    for want in case["expected"]:
        n = got[want["name"]]
        assert n.op == want["op"] and n.inputs == want["inputs"]
        if "dtype" in want:
            assert graphdef.DT_NAME[n.attr["dtype"][1]] == want["dtype"]
        if "T" in want:
            assert graphdef.DT_NAME[n.attr["T"][1]] == want["T"]
        if "value" in want:
            v = n.attr["value"][1]
            assert graphdef.DT_NAME[{np.dtype("<i4"): 3, np.dtype("<f8"): 2}[v.dtype]] == want["value"]["dtype"]
            assert list(v.shape) == want["value"]["shape"] and v.reshape(-1).tolist() == want["value"]["values"]


@pytest.mark.parametrize("case", G["dense"], ids=[c["id"] for c in G["dense"]])
def test_dense_tensor_endianness(case):
    with dsl.withGraph():
        g = dsl.build_graph(dsl.constant(case["value"]).named("c"))
    (n,) = graphdef.parse_graph(g)
    assert n.attr["value"][1].item() == case["value"]


def test_scopes_and_counters():
    with dsl.withGraph():
        with dsl.scope("outer"):
            a = dsl.constant(1.0)
            with dsl.scope("inner"):
                b = dsl.constant(2.0).named("b")
            c = dsl.constant(3.0)
        g = dsl.build_graph(a, b, c)
    assert sorted(n.name for n in graphdef.parse_graph(g)) == ["outer/Const", "outer/Const_1", "outer/inner/b"]


def test_reducer_emits_reduction_indices_const():
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, Unknown, 2).named("x_input")
        s = dsl.reduce_sum(x, [0]).named("x")
        g = dsl.build_graph(s)
    nodes = {n.name: n for n in graphdef.parse_graph(g)}
    assert nodes["x"].op == "Sum" and nodes["x"].inputs == ["x_input", "x_input/reduction_indices"]
    assert nodes["x"].attr["keep_dims"] == ("b", False) and nodes["x"].attr["Tidx"] == ("type", 3)
    assert nodes["x_input/reduction_indices"].attr["value"][1].tolist() == [0]
    assert s.shape == Shape(2)


# ---- graph import + analysis through the C ABI (no device needed) ------------------------------
@pytest.mark.parametrize("name", sorted(G["fixtures"]))
def test_native_import_of_reference_fixtures(name):
    fx = G["fixtures"][name]
    with _native.Graph(bytes.fromhex(fx["hex"])) as g:
        assert [n for n, _ in g.nodes()] == [n["name"] for n in fx["nodes"]]
        for want in fx["nodes"]:
            dtype, shape, is_ph, n_in = g.node_info(want["name"])
            assert graphdef.DT_NAME[dtype.tf_code] == want["dtype"]
            assert list(shape.dims) == want["shape"]
            assert is_ph == (want["op"] == "Placeholder")
            assert n_in == len(want.get("inputs", []))


def test_native_static_shape_inference():
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, Unknown, 2).named("x")
        y = (x * 2.0 + np.asarray([1.0, 2.0])).named("y")
        s = dsl.reduce_sum(y, [0]).named("s")
        a = dsl.reduce_max(y, None).named("a")
        c = dsl.cast(s, tfs.FloatType).named("c")
        g = dsl.build_graph(s, a, c)
    with _native.Graph(g) as gr:
        assert gr.node_info("y")[:2] == (tfs.DoubleType, Shape(Unknown, 2))
        assert gr.node_info("s:0")[:2] == (tfs.DoubleType, Shape(2))
        assert gr.node_info("a")[:2] == (tfs.DoubleType, Shape())
        assert gr.node_info("c")[:2] == (tfs.FloatType, Shape(2))
        with pytest.raises(_native.NativeError):
            gr.node_info("s:1")
        with pytest.raises(_native.NativeError):
            gr.node_info("nope")


def test_native_scalar_placeholder_is_unconstrained():
    # TF 1.x: a Placeholder whose shape attr has no dims has unknown shape, which is why the
    # reference ships shape hints for every input (core.py:52-72)
    with dsl.withGraph():
        g = dsl.build_graph(dsl.identity(dsl.placeholder(tfs.DoubleType).named("x")).named("y"))
    with _native.Graph(g) as gr:
        assert gr.node_info("x")[1] is None


def test_native_rejects_bad_graphs():
    with pytest.raises(_native.NativeError) as e:
        _native.Graph(b"\x0a\x05\x0a\x01x\x12")        # truncated
    assert e.value.code == _native.TFS_ERR_INVALID
    from tensorframes_b200 import proto
    bad = proto.graph_def([proto.node_def("w", "Conv2D", [], {"T": proto.AttrValue.of_type(tfs.FloatType)})])
    with pytest.raises(_native.UnsupportedError) as e:
        _native.Graph(bad)
    assert "Conv2D" in str(e.value) and "'w'" in str(e.value)
    dangling = proto.graph_def([proto.node_def("y", "Identity", ["x"], {"T": proto.AttrValue.of_type(tfs.FloatType)})])
    with pytest.raises(_native.NativeError):
        _native.Graph(dangling)
    mixed = proto.graph_def([
        proto.node_def("a", "Placeholder", [], {"dtype": proto.AttrValue.of_type(tfs.FloatType)}),
        proto.node_def("b", "Placeholder", [], {"dtype": proto.AttrValue.of_type(tfs.DoubleType)}),
        proto.node_def("c", "Add", ["a", "b"], {"T": proto.AttrValue.of_type(tfs.FloatType)})])
    with pytest.raises(_native.NativeError):
        _native.Graph(mixed)


def test_damaged_graph_bytes_end_in_a_status_code():
    """2000 byte-level mutations (bit flips, truncations, insertions) of valid graphs: import either succeeds, and
    every node can then be described, or fails with a status; nothing may crash the process."""
    import _mutate
    graphs = _mutate.seed_graphs()
    rng = np.random.default_rng(11)
    imported = 0
    for it in range(2000):
        g = _mutate.mutate(rng, graphs[it % len(graphs)], structural=True)
        try:
            G = _native.Graph(g)
        except _native.NativeError as e:
            assert str(e)
            continue
        imported += 1
        h, lib = G.handle, _native.lib()
        for k in range(lib.tfs_graph_num_nodes(h)):
            info = _native.NodeInfo()
            lib.tfs_graph_node_info(h, lib.tfs_graph_node_name(h, k), ctypes.byref(info))
        G.close()
    assert 50 < imported < 1500


def test_analyze_graph_hints_override():
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, Unknown).named("x")
        z = (x + 3.0).named("z")
        g = dsl.build_graph(z)
    hints = ShapeDescription({"z": Shape(10)}, ("z:0",), {})
    with _native.Graph(g) as gr:
        summary = {s.name: s for s in ops.analyze_graph(gr, hints)}
    assert summary["z"].isOutput and summary["z"].shape == Shape(10)
    assert summary["x"].isInput and summary["x"].isPlaceholder and summary["x"].shape == Shape(Unknown)


# ---- analyze / explain ---------------------------------------------------------------------------
@pytest.mark.parametrize("case", G["analyze"], ids=[c["id"] for c in G["analyze"]])
def test_analyze_matches_reference_tests(case):
    df = _golden.build_frame(case)
    for name, (dt, shape) in case["expected"].items():
        stf = ColumnInformation.of(df.schema[name]).stf
        assert stf.dataType == _golden.TYPES[dt]
        assert list(stf.shape.dims) == shape


def test_explain_format():
    df = TensorFrame.from_rows([([float(i), float(-i)], "k") for i in range(10)], ["y", "key"])
    assert explain(df) == ("root\n |-- y: array (nullable = false) double[?,?]\n"
                           " |-- key: string (nullable = false) <no tensor info>\n")
    assert " |-- y: array (nullable = false) double[10,2]\n" in explain(tfs.analyze(df))


def test_metadata_keys_and_output_field():
    f = ColumnInformation.structField("z", tfs.FloatType, Shape(Unknown, 3))
    assert f.metadata == {"org.spartf.shape": [-1, 3], "org.sparktf.type": "FloatType"}
    assert not f.nullable and repr(f.dataType) == "ArrayType(FloatType,false)"


def test_infer_physical_shape():
    assert ops.infer_physical_shape(6, Shape(3), 2) == (2, Shape(3))
    assert ops.infer_physical_shape(6, Shape(Unknown), 2) == (2, Shape(3))
    assert ops.infer_physical_shape(6, Shape(2), None) == (3, Shape(2))
    with pytest.raises(ValueError):
        ops.infer_physical_shape(6, Shape(Unknown), None)
    with pytest.raises(ValueError):
        ops.infer_physical_shape(7, Shape(3), 2)


# ---- validation errors are raised on the host, before any device work -----------------------------
def _df():
    return TensorFrame.from_rows([(1.0, 1, "a"), (2.0, 2, "b")], ["x", "i", "s"])


def test_map_blocks_validation_errors():
    df = _df()
    o = ops.B200Ops()
    with dsl.withGraph():
        with pytest.raises(Exception, match="no column to match it"):
            o.mapBlocks(df, dsl.identity(dsl.placeholder(tfs.DoubleType, Unknown).named("nope")).named("out"))
    with dsl.withGraph():
        with pytest.raises(Exception, match="not compatible with the data type"):
            o.mapBlocks(df, dsl.identity(dsl.placeholder(tfs.DoubleType, Unknown).named("i")).named("out"))
    with dsl.withGraph():
        with pytest.raises(Exception, match="not compatible\\) with shape"):
            o.mapBlocks(df, dsl.identity(dsl.placeholder(tfs.DoubleType, Unknown, 2).named("x")).named("out"))
    with dsl.withGraph():
        with pytest.raises(Exception, match="but this column already exists"):
            o.mapBlocks(df, dsl.identity(dsl.placeholder(tfs.DoubleType, Unknown).named("x")).named("i"))
    with dsl.withGraph():
        with pytest.raises(Exception, match="no column to match it"):
            o.mapBlocks(df, dsl.identity(dsl.placeholder(tfs.DoubleType, Unknown).named("s2")).named("out"))


def test_reduce_blocks_validation_errors():
    df = _df()
    o = ops.B200Ops()
    with dsl.withGraph():
        xi = dsl.placeholder(tfs.DoubleType, Unknown).named("x_input")
        with pytest.raises(Exception, match="some inputs are missing: y"):
            o.reduceBlocks(df, dsl.reduce_sum(xi, [0]).named("y"))
    with dsl.withGraph():
        xi = dsl.placeholder(tfs.DoubleType, Unknown).named("x_in")
        with pytest.raises(Exception, match="Extra graph inputs have been found: x_in"):
            o.reduceBlocks(df, dsl.reduce_sum(xi, [0]).named("x"))
    with dsl.withGraph():
        xi = dsl.placeholder(tfs.IntegerType, Unknown).named("x_input")
        with pytest.raises(Exception, match="has type IntegerType but the column type is DoubleType"):
            o.reduceBlocks(df, dsl.reduce_sum(xi, [0]).named("x"))
    with dsl.withGraph():
        xi = dsl.placeholder(tfs.DoubleType, Unknown).named("x_input")
        with pytest.raises(Exception, match="not compatible with the shape"):
            o.reduceBlocks(df, dsl.identity(xi).named("x"))


def test_unanalyzed_vector_column_is_rejected():
    df = TensorFrame.from_rows([([1.0, 2.0],), ([3.0, 4.0],)], ["v"])
    o = ops.B200Ops()
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, Unknown, Unknown).named("v")
        with pytest.raises(Exception):   # cell shape unknown: impl/datatypes.scala:80-83
            o.mapBlocks(df, dsl.identity(p).named("out"))


def test_python_api_names():
    # tensorframes/core.py:10-11
    for name in ["reduce_rows", "map_rows", "reduce_blocks", "map_blocks", "analyze", "print_schema", "aggregate",
                 "block", "row"]:
        assert callable(getattr(tfs, name))


# ---- Arrow ingestion / egress (columnar hand-over without per-row work) ----------
def test_from_arrow_is_zero_copy_and_keeps_chunks_as_partitions():
    pa = pytest.importorskip("pyarrow")
    from tensorframes_b200.frame import TensorFrame, ArrayType
    n = 10
    x = np.arange(n, dtype=np.float64)
    v = np.arange(n * 6, dtype=np.float32).reshape(n, 2, 3)
    fsl = pa.FixedSizeListArray.from_arrays(pa.FixedSizeListArray.from_arrays(pa.array(v.reshape(-1)), 3), 2)
    key = pa.array([f"k{i % 3}" for i in range(n)])
    tbl = pa.table({"x": pa.array(x), "v": fsl, "key": key})
    two = pa.Table.from_batches(tbl.slice(0, 4).to_batches() + tbl.slice(4).to_batches())
    df = TensorFrame.from_arrow(two)
    assert len(df.partitions) == 2 and df.count() == n
    assert df.schema["x"].dataType == tfs.DoubleType
    assert isinstance(df.schema["v"].dataType, ArrayType) and isinstance(df.schema["v"].dataType.elementType, ArrayType)
    got = np.concatenate([p["v"] for p in df.partitions])
    assert got.dtype == np.float32 and got.shape == (n, 2, 3) and np.array_equal(got, v)
    # the blocks are views over Arrow's own value buffer
    values_buf = two.column("v").chunk(1).flatten().flatten().buffers()[1]
    lo = values_buf.address
    assert lo <= df.partitions[1]["v"].ctypes.data < lo + values_buf.size
    assert df.partitions[0]["key"] == ["k0", "k1", "k2", "k0"]
    # contiguous re-split
    df3 = TensorFrame.from_arrow(two, num_partitions=3)
    assert [len(p["x"]) for p in df3.partitions] == [3, 3, 4]
    assert np.array_equal(np.concatenate([p["x"] for p in df3.partitions]), x)
    # and back
    back = df.to_arrow()
    assert back.num_rows == n and back.column("v").type == fsl.type
    assert back.column("x").to_pylist() == x.tolist() and back.column("key").to_pylist() == key.to_pylist()
    assert np.array_equal(np.asarray(back.column("v").to_pylist(), dtype=np.float32), v)


def test_from_arrow_lists_and_rejections():
    pa = pytest.importorskip("pyarrow")
    from tensorframes_b200.frame import TensorFrame
    same = pa.array([[1, 2], [3, 4], [5, 6]], type=pa.list_(pa.int32()))
    ragged = pa.array([[1.0], [2.0, 3.0], []], type=pa.list_(pa.float64()))
    df = TensorFrame.from_arrow(pa.table({"a": same, "b": ragged}))
    a, b = df.partitions[0]["a"], df.partitions[0]["b"]
    assert isinstance(a, np.ndarray) and a.shape == (3, 2) and a.dtype == np.int32
    assert isinstance(b, list) and [c.tolist() for c in b] == [[1.0], [2.0, 3.0], []]
    with pytest.raises(ValueError, match="nulls"):
        TensorFrame.from_arrow(pa.table({"x": pa.array([1.0, None])}))
    with pytest.raises(TypeError, match="not supported"):
        TensorFrame.from_arrow(pa.table({"x": pa.array([True, False])}))


def test_python_scalars_take_the_tensor_dtype():
    """x_float32 * 1.5 stays float32 (python TF's operator overloads convert the scalar); arrays keep their dtype."""
    with dsl.withGraph():
        x = dsl.placeholder(tfs.FloatType, Unknown).named("x")
        z = (2 - x * 1.5).named("z")
        g = dsl.build_graph(z)
    assert z.scalarType is tfs.FloatType
    nodes = {n.name: n for n in graphdef.parse_graph(g)}
    consts = [n for n in nodes.values() if n.op == "Const"]
    assert len(consts) == 2 and all(n.attr["dtype"] == ("type", 1) for n in consts)
    with dsl.withGraph():
        i = dsl.placeholder(tfs.LongType, Unknown).named("i")
        assert (i + 3).scalarType is tfs.LongType
        with pytest.raises(AssertionError, match="should be the same"):
            i + np.float64(3.0)           # a typed numpy scalar is not converted
