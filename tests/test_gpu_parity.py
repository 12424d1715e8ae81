"""Parity tests proper: the CUDA path (through the C ABI, and through the python operators that
sit on it) against the oracle and the reference's golden vectors.  Needs a B200.

Bars: integer, index and key results bit-exact; element-wise floating point bit-exact (same IEEE
operations in the same order as TF-CPU, no FMA contraction); floating-point reductions within
1e-12 (f64) / 1e-5 (f32) of the exact sum, measured normwise (|gpu - exact| <= tol * sum|x|):
reduction order is the only permitted difference.
"""
import numpy as np
import pytest

import _golden
import tensorframes_b200 as tfs
from oracle import tf_ops
from tensorframes_b200 import _native, dsl
from tensorframes_b200.frame import TensorFrame

pytestmark = pytest.mark.gpu

G = _golden.load()
T = {"f32": tfs.FloatType, "f64": tfs.DoubleType, "i32": tfs.IntegerType, "i64": tfs.LongType}
TOL = {np.dtype("float64"): 1e-12, np.dtype("float32"): 1e-5}


def run_native(graph_bytes, feeds, fetches):
    with _native.Graph(graph_bytes) as g, _native.Session(g, 0) as s:
        return s.run(feeds, fetches)


def rand(rng, st, shape, lo=-100, hi=100, nonzero=False):
    if st.np_dtype.kind == "f":
        a = rng.uniform(lo, hi, shape).astype(st.np_dtype)
        if nonzero:
            a = np.where(np.abs(a) < 0.5, a + np.sign(a + 0.1) * 1.0, a).astype(st.np_dtype)
    else:
        a = rng.integers(lo, hi, shape).astype(st.np_dtype)
        if nonzero:
            a = np.where(a == 0, 7, a).astype(st.np_dtype)
    return a


def assert_same(got, want):
    assert got.dtype == want.dtype and got.shape == want.shape, (got.dtype, want.dtype, got.shape, want.shape)
    assert got.tobytes() == want.tobytes(), np.argwhere(got != want)[:5]


# ---- graphs the product did not write: the golden cases' graphs from the oracle's own writer, run as one block ------
def _block_feeds(case, cols, parts, graph):
    """Dense [n, cell...] arrays for the block placeholders of a case's graph, or None when a column is ragged."""
    from oracle import graphdef
    rows = [r for p in parts for r in p]
    spec = case["graph"]
    source = {name: name for name, _dt, _shape in spec.get("placeholders", [])}
    source.update({tf_name: col for col, tf_name in spec.get("block_placeholders", [])})
    names = [c.name for c in cols]
    feeds = {}
    for n in graphdef.parse_graph(graph):
        if n.op != "Placeholder":
            continue
        col = source[n.name] if source[n.name] in names else source[n.name].replace("_input", "")
        c = cols[names.index(col)]
        try:
            feeds[n.name] = np.asarray([r[names.index(col)] for r in rows], dtype=graphdef.DT[c.dtype])
        except ValueError:
            return None
    return feeds


@pytest.mark.parametrize("case", [c for c in G["cases"] if c["kind"] in ("map_blocks", "map_blocks_trimmed", "reduce_blocks", "aggregate")],
                         ids=lambda c: c["id"])
def test_library_runs_graphs_from_the_independent_writer(case):
    """tfs_session_run on GraphDef bytes written by oracle/graph_writer.py (no product encoder involved), the whole
    frame as one block, against the oracle's evaluator on the same bytes: bit-exact."""
    from oracle import graph_writer, tf_ops
    df = _golden.build_frame(case)
    parts, cols = _golden.oracle_inputs(case, df)
    columns = {c.name: (c.dtype, tuple(c.shape)) for c in cols if c.dtype is not None}
    graph, _info = graph_writer.from_spec(case["graph"], columns)
    feeds = _block_feeds(case, cols, parts, graph)
    if feeds is None or any(v.size == 0 for v in feeds.values()):
        pytest.skip("ragged or empty column: not one dense block")
    fetches = list(case["graph"]["fetches"])
    want = tf_ops.evaluate(graph, feeds, fetches)
    with _native.Graph(graph) as gr, _native.Session(gr, 0) as s:
        got = s.run(feeds, fetches)
    for g_, w_ in zip(got, want):
        assert_same(g_, np.asarray(w_))


# ---- the reference's golden vectors through the product's operators --------------------------------
@pytest.mark.parametrize("case", G["cases"], ids=[c["id"] for c in G["cases"]])
def test_golden_through_operators(case):
    df = _golden.build_frame(case)
    kind = case["kind"]
    with dsl.withGraph():
        fetches = _golden.build_fetches(case, df)
        if kind == "map_blocks":
            out = tfs.map_blocks(fetches, df)
        elif kind == "map_blocks_trimmed":
            out = tfs.map_blocks(fetches, df, trim=True)
        elif kind == "reduce_blocks":
            res = tfs.reduce_blocks(fetches, df)
            res = res if isinstance(res, list) else [res]
            got = [r.tolist() if isinstance(r, np.ndarray) else r for r in res]
            assert got == case["expected"]
            return
        elif kind == "aggregate":
            out = tfs.aggregate(fetches, df.groupBy(*case["keys"]))
        elif kind == "map_rows":
            out = tfs.map_rows(fetches, df, feed_dict=case.get("feed_dict"))
        elif kind == "reduce_rows":
            res = tfs.reduce_rows(fetches, df)
            res = res if isinstance(res, list) else [res]
            got = [r.tolist() if isinstance(r, np.ndarray) else r for r in res]
            assert got == case["expected"]
            return
    if case.get("expected_columns"):
        assert out.columns == case["expected_columns"]
    if kind == "map_blocks" and not case.get("select"):
        # direct executor test of the reference: TF columns first, then the inputs
        assert out.columns[0] == case["graph"]["fetches"][0]
    sel = out.select(*case["select"]) if case.get("select") else out
    assert _golden.normalize(sel.collect()) == _golden.normalize(case["expected"])


# ---- element-wise: every op x dtype x ragged sizes, bit-exact --------------------------------------
SIZES = [1, 2, 3, 5, 1023, 1024, 1025, 4099, 100003]


@pytest.mark.parametrize("dt", list(T))
@pytest.mark.parametrize("n", SIZES)
def test_binary_ops_bit_exact(dt, n):
    st = T[dt]
    rng = np.random.default_rng(n)
    a, b = rand(rng, st, n), rand(rng, st, n, nonzero=True)
    with dsl.withGraph():
        pa = dsl.placeholder(st, dsl.Unknown).named("a")
        pb = dsl.placeholder(st, dsl.Unknown).named("b")
        outs = [dsl.add(pa, pb).named("o_add"), dsl.sub(pa, pb).named("o_sub"), dsl.mul(pa, pb).named("o_mul"),
                dsl.div(pa, pb).named("o_div"), dsl.minimum(pa, pb).named("o_min"), dsl.maximum(pa, pb).named("o_max")]
        g = dsl.build_graph(*outs)
    names = [o.name for o in outs]
    got = run_native(g, {"a": a, "b": b}, names)
    want = tf_ops.evaluate(g, {"a": a, "b": b}, names)
    for gg, ww in zip(got, want):
        assert_same(gg, ww)


@pytest.mark.parametrize("dt", list(T))
@pytest.mark.parametrize("n", [1, 7, 1025, 100003])
def test_unary_and_cast_bit_exact(dt, n):
    st = T[dt]
    rng = np.random.default_rng(n + 1)
    a = rand(rng, st, n, nonzero=True)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown).named("a")
        outs = [dsl.neg(p).named("o_neg"), dsl.relu(p).named("o_relu"), dsl.identity(p).named("o_id"),
                dsl._build("Square", parents=[p]).named("o_sq"), dsl._build("Abs", parents=[p]).named("o_abs")]
        if st.np_dtype.kind == "f":
            outs += [dsl._build("Sqrt", parents=[dsl._build("Abs", parents=[p])]).named("o_sqrt"),
                     dsl._build("Reciprocal", parents=[p]).named("o_rcp")]
        for k, dst in T.items():
            outs.append(dsl.cast(p, dst).named("o_cast_" + k))
        g = dsl.build_graph(*outs)
    names = [o.name for o in outs]
    got = run_native(g, {"a": a}, names)
    want = tf_ops.evaluate(g, {"a": a}, names)
    for name, gg, ww in zip(names, got, want):
        assert_same(gg, ww)


def test_config3_chain_bit_exact():
    """BASELINE config 3 at a size the oracle finishes in seconds: cast_f32(max(x*a+b, 0))."""
    rng = np.random.default_rng(3)
    x = rng.uniform(-1, 1, 3_000_001)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        z = dsl.cast(dsl.maximum(p * 1.5 + 0.25, 0.0), tfs.FloatType).named("z")
        g = dsl.build_graph(z)
    (got,) = run_native(g, {"x": x}, ["z"])
    (want,) = tf_ops.evaluate(g, {"x": x}, ["z"])
    assert_same(got, want)
    # no FMA contraction: x*a+b must round twice
    assert_same(got, np.maximum((x * 1.5) + 0.25, 0.0).astype(np.float32))


def test_dag_programs():
    """Shared sub-expressions, both operands computed, x+x, several fetches, row broadcast."""
    rng = np.random.default_rng(11)
    n = 5003
    x, y = rng.uniform(-4, 4, (n, 3)), rng.uniform(1, 2, (n, 3))
    w = np.asarray([0.5, -2.0, 3.0])
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 3).named("x")
        py = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 3).named("y")
        s = (px + py)                       # used three times
        d = (px - py) * (px + px)           # both sides computed; x + x
        o1 = (s * s).named("o1")
        o2 = ((s / py) - d).named("o2")
        o3 = (dsl.constant(w) * px + dsl.constant(w)).named("o3")   # [3] against [n,3]
        o4 = dsl.cast(dsl.minimum(d, s), tfs.LongType).named("o4")
        o5 = (2.0 - px).named("o5")         # constant on the left of a non-commutative op
        o6 = (py / (px * px + 1.0)).named("o6")
        g = dsl.build_graph(o1, o2, o3, o4, o5, o6)
    names = ["o1", "o2", "o3", "o4", "o5", "o6"]
    got = run_native(g, {"x": x, "y": y}, names)
    want = tf_ops.evaluate(g, {"x": x, "y": y}, names)
    for gg, ww in zip(got, want):
        assert_same(gg, ww)


def test_mixed_width_program():
    rng = np.random.default_rng(12)
    n = 70001
    a = rng.integers(-1000, 1000, n).astype(np.int32)
    b = rng.uniform(-10, 10, n).astype(np.float32)
    with dsl.withGraph():
        pa = dsl.placeholder(tfs.IntegerType, dsl.Unknown).named("a")
        pb = dsl.placeholder(tfs.FloatType, dsl.Unknown).named("b")
        o1 = (dsl.cast(pa, tfs.DoubleType) * dsl.cast(pb, tfs.DoubleType)).named("o1")
        o2 = dsl.cast(dsl.cast(pa, tfs.LongType) * dsl.constant(np.int64(1 << 33)), tfs.FloatType).named("o2")
        o3 = (pa * pa).named("o3")
        g = dsl.build_graph(o1, o2, o3)
    got = run_native(g, {"a": a, "b": b}, ["o1", "o2", "o3"])
    want = tf_ops.evaluate(g, {"a": a, "b": b}, ["o1", "o2", "o3"])
    for gg, ww in zip(got, want):
        assert_same(gg, ww)


def test_integer_wraparound_and_division():
    a = np.array([2**31 - 1, -2**31, 7, -7, 7, -7], dtype=np.int32)
    b = np.array([1, -1, 2, 2, -2, -2], dtype=np.int32)
    with dsl.withGraph():
        pa = dsl.placeholder(tfs.IntegerType, dsl.Unknown).named("a")
        pb = dsl.placeholder(tfs.IntegerType, dsl.Unknown).named("b")
        g = dsl.build_graph((pa + pb).named("s"), (pa * pb).named("m"), dsl.div(pa, pb).named("q"))
    s, m, q = run_native(g, {"a": a, "b": b}, ["s", "m", "q"])
    assert s.tolist() == [-2**31, 2**31 - 1, 9, -5, 5, -9]
    assert q.tolist()[2:] == [3, -3, -3, 3]
    want = tf_ops.evaluate(g, {"a": a[2:], "b": b[2:]}, ["s", "m", "q"])
    assert m[2:].tolist() == want[1].tolist()


def test_integer_division_by_zero_is_an_error():
    with dsl.withGraph():
        pa = dsl.placeholder(tfs.LongType, dsl.Unknown).named("a")
        pb = dsl.placeholder(tfs.LongType, dsl.Unknown).named("b")
        g = dsl.build_graph(dsl.div(pa, pb).named("q"))
    a = np.arange(1, 1001, dtype=np.int64)
    b = np.ones(1000, dtype=np.int64)
    bad = b.copy()
    bad[777] = 0
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.NativeError) as e:
            s.run({"a": a, "b": bad}, ["q"])
        assert e.value.code == _native.TFS_ERR_RUNTIME and "Integer division by zero" in str(e.value)
        (q,) = s.run({"a": a, "b": b}, ["q"])       # the session stays usable
        assert q.tolist() == a.tolist()


# ---- block reductions ----------------------------------------------------------------------------
RED = {"sum": dsl.reduce_sum, "min": dsl.reduce_min, "max": dsl.reduce_max}


def check_reduction(got, want, x, axis):
    if x.dtype.kind == "f" and want.dtype.kind == "f":
        scale = np.abs(x.astype(np.float64)).sum(axis=axis)
        assert got.dtype == want.dtype and got.shape == want.shape
        err = np.abs(got.astype(np.float64) - want.astype(np.float64))
        assert np.all(err <= TOL[x.dtype] * np.maximum(scale, 1e-300)), float((err / np.maximum(scale, 1e-300)).max())
    else:
        assert_same(got, want)


@pytest.mark.parametrize("dt", list(T))
@pytest.mark.parametrize("op", list(RED))
@pytest.mark.parametrize("shape", [(1,), (2,), (1000,), (100003,), (7, 2), (100003, 2), (1001, 3), (4097, 5),
                                   (3001, 64), (501, 100), (37, 1000), (1, 2), (300001, 1), (129, 4, 2)])
def test_axis0_reductions(dt, op, shape):
    st = T[dt]
    rng = np.random.default_rng(sum(shape))
    x = rand(rng, st, shape)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, *shape[1:]).named("x_input")
        g = dsl.build_graph(RED[op](p, [0]).named("x"))
    (got,) = run_native(g, {"x_input": x}, ["x"])
    (want,) = tf_ops.evaluate(g, {"x_input": x}, ["x"])
    if op == "sum":
        check_reduction(got, want, x, 0)
    else:
        assert_same(got, want)


@pytest.mark.parametrize("dt", ["f64", "i32"])
def test_all_axes_and_trailing_axis_reductions(dt):
    st = T[dt]
    rng = np.random.default_rng(5)
    x = rand(rng, st, (1031, 7))
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, 7).named("x")
        outs = [dsl.reduce_sum(p, None).named("all"), dsl.reduce_max(p, [1]).named("rows"),
                dsl.reduce_sum(p, [1]).named("rowsum"), dsl.reduce_min(p, [0, 1]).named("allmin")]
        g = dsl.build_graph(*outs)
    names = ["all", "rows", "rowsum", "allmin"]
    got = run_native(g, {"x": x}, names)
    want = tf_ops.evaluate(g, {"x": x}, names)
    check_reduction(got[0], want[0], x, None)
    assert_same(got[1], want[1])
    check_reduction(got[2], want[2], x, 1)
    assert_same(got[3], want[3])


def test_exact_sum_reference_f64():
    """The f64 block sum against math.fsum (the correctly rounded sum), config-2 shape."""
    rng = np.random.default_rng(2)
    u = rng.uniform(0, 1, 200_000)
    y = np.stack([u, -u], axis=1)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
        g = dsl.build_graph(dsl.reduce_sum(p, [0]).named("y"), dsl.reduce_min(p, [0]).named("z"))
    s, m = run_native(g, {"y_input": y}, ["y", "z"])
    exact = tf_ops.exact_sum_f64(y, 0)
    assert np.all(np.abs(s - exact) <= 1e-12 * np.abs(y).sum(axis=0))
    assert m.tolist() == y.min(axis=0).tolist()


def test_reduction_is_deterministic():
    rng = np.random.default_rng(9)
    x = rng.normal(size=(1_000_003, 2))
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("x_input")
        g = dsl.build_graph(dsl.reduce_sum(p, [0]).named("x"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        first = s.run({"x_input": x}, ["x"])[0]
        for _ in range(5):
            assert s.run({"x_input": x}, ["x"])[0].tobytes() == first.tobytes()


def test_reduce_then_map_in_one_graph():
    """A fetch computed from a reduction (scalar operand produced on the device, second fused group)."""
    rng = np.random.default_rng(21)
    x = rng.uniform(1, 2, (4001, 2))
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("x")
        centered = (p - dsl.reduce_min(p, [0])).named("centered")
        frac = (p / dsl.reduce_max(p, None)).named("frac")
        g = dsl.build_graph(centered, frac)
    got = run_native(g, {"x": x}, ["centered", "frac"])
    want = tf_ops.evaluate(g, {"x": x}, ["centered", "frac"])
    for gg, ww in zip(got, want):
        assert_same(gg, ww)


def test_constants_fill_and_zero_rows():
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g = dsl.build_graph(dsl.fill([5], 2.5).named("f"), dsl.constant([[1.0, 2.0]]).named("c"),
                            dsl.reduce_sum(dsl.constant([1.0, 1.0]), [0]).named("s"), (p + 1.0).named("y"),
                            dsl.reduce_sum(p, [0]).named("e"))
    f, c, s, y, e = run_native(g, {"x": np.zeros(0)}, ["f", "c", "s", "y", "e"])
    assert f.tolist() == [2.5] * 5 and c.tolist() == [[1.0, 2.0]] and s.item() == 2.0
    assert y.shape == (0,) and e.item() == 0.0


# ---- device-resident buffers (no host copies), unaligned pointers ---------------------------------
def test_device_pointer_path_matches_host_path():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(31)
    n = 1_000_001
    x = rng.uniform(-1, 1, n + 1)
    y = rng.uniform(0, 1, (n, 2))
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        py = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
        g = dsl.build_graph(dsl.cast(dsl.maximum(px * 1.5 + 0.25, 0.0), tfs.FloatType).named("z"),
                            dsl.reduce_sum(py, [0]).named("y"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        host = s.run({"x": x[1:], "y_input": y}, ["z", "y"])
        dx = torch.from_numpy(x).cuda()[1:]        # 8-byte aligned only: exercises the scalar variant
        dy = torch.from_numpy(y).cuda()
        dz = torch.empty(n, dtype=torch.float32, device="cuda")
        ds = torch.empty(2, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"x": _native.DeviceArray.from_torch(dx.contiguous()) if dx.data_ptr() % 16 == 0 else
                    _native.DeviceArray(dx.data_ptr(), tfs.DoubleType, (n,), owner=dx),
                    "y_input": _native.DeviceArray.from_torch(dy)},
                   {"z": _native.DeviceArray.from_torch(dz), "y": _native.DeviceArray.from_torch(ds)})
        s.sync()
        assert dx.data_ptr() % 16 == 8
        assert dz.cpu().numpy().tobytes() == host[0].tobytes()
        assert ds.cpu().numpy().tobytes() == host[1].tobytes()


def test_full_size_properties_on_device():
    """BASELINE-scale blocks (1e8 rows) checked through closed forms that need no CPU pass."""
    torch = pytest.importorskip("torch")
    n = 100_000_000
    i = torch.arange(n, dtype=torch.float64, device="cuda")
    u = (i % 1024.0)
    y = torch.stack([u, -u], dim=1).contiguous()
    del i
    with dsl.withGraph():
        py = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
        pz = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("z_input")
        g = dsl.build_graph(dsl.reduce_sum(py, [0]).named("y"), dsl.reduce_min(pz, [0]).named("z"))
    ds = torch.empty(2, dtype=torch.float64, device="cuda")
    dm = torch.empty(2, dtype=torch.float64, device="cuda")
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"y_input": _native.DeviceArray.from_torch(y), "z_input": _native.DeviceArray.from_torch(y)},
                   {"y": _native.DeviceArray.from_torch(ds), "z": _native.DeviceArray.from_torch(dm)})
        s.sync()
    full, rem = divmod(n, 1024)
    exact = full * (1023 * 1024 // 2) + (rem - 1) * rem // 2     # integers: every partial sum is exact in f64
    assert ds.tolist() == [float(exact), float(-exact)]
    assert dm.tolist() == [0.0, -1023.0]
    del y, u
    torch.cuda.empty_cache()
    # map chain at BASELINE config 2's size (1e9 rows): linear in x, so z(x) on a known ramp has a closed form per element
    n = 1_000_000_000
    x = torch.empty(n, dtype=torch.float64, device="cuda")
    for a in range(0, n, 1 << 27):
        b = min(n, a + (1 << 27))
        x[a:b] = (torch.arange(a, b, dtype=torch.float64, device="cuda") % 1024.0) * (1.0 / 512.0) - 1.0
    dz = torch.empty(n, dtype=torch.float32, device="cuda")
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g = dsl.build_graph(dsl.cast(dsl.maximum(px * 1.5 + 0.25, 0.0), tfs.FloatType).named("z"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"x": _native.DeviceArray.from_torch(x)}, {"z": _native.DeviceArray.from_torch(dz)})
        s.sync()
    for a in range(0, n, 1 << 27):
        b = min(n, a + (1 << 27))
        assert torch.equal(dz[a:b], torch.clamp_min(x[a:b] * 1.5 + 0.25, 0.0).to(torch.float32))
    # the ramp has period 1024: every period of the output must equal the first one
    assert torch.equal(dz[:1024 * 976_562].view(-1, 1024), dz[:1024].expand(976_562, 1024))
    # reduce_blocks on the same 1e9-row scalar column (the other half of the north-star target): exact in f64
    with dsl.withGraph():
        ps = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("s_input")
        pm = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("m_input")
        g = dsl.build_graph(dsl.reduce_sum(ps, [0]).named("s"), dsl.reduce_max(pm, [0]).named("m"))
    ds1 = torch.empty((), dtype=torch.float64, device="cuda")
    dm1 = torch.empty((), dtype=torch.float64, device="cuda")
    k = (x + 1.0) * 512.0                      # back to the integers 0..1023: their sum is exact in f64
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"s_input": _native.DeviceArray.from_torch(k), "m_input": _native.DeviceArray.from_torch(x)},
                   {"s": _native.DeviceArray.from_torch(ds1), "m": _native.DeviceArray.from_torch(dm1)})
        s.sync()
    full, rem = divmod(n, 1024)
    assert ds1.item() == float(full * (1023 * 1024 // 2) + (rem - 1) * rem // 2)
    assert dm1.item() == 1023.0 / 512.0 - 1.0


def test_full_size_aggregate_closed_form():
    """BASELINE config 4 at its configured size (5e8 rows, 1e6 int64 keys, FloatType[?,64]; 132 GB of input, reduced in
    two 64 GB row ranges whose partial groups are merged): small-integer payloads make every group sum exact in
    float32, so the result has a closed form that needs no CPU pass."""
    torch = pytest.importorskip("torch")
    torch.cuda.empty_cache()
    n, n_keys, cell = 500_000_000, 1_000_000, 64
    i = torch.arange(n, dtype=torch.int64, device="cuda")
    gid = (i * 7919) % n_keys                                   # 7919 is coprime to 1e6: every key gets exactly n / n_keys rows
    keys = gid * 0x9E3779B97F4A7C15 + 12345                     # wrapping multiply by an odd constant: a bijection on int64
    v = torch.empty((n, cell), dtype=torch.float32, device="cuda")
    for a in range(0, n, 1 << 25):                              # chunked: no second 128 GB temporary
        v[a:a + (1 << 25)] = (i[a:a + (1 << 25)] % 5).to(torch.float32)[:, None] + torch.arange(cell, dtype=torch.float32, device="cuda")[None, :] % 3
    per = n // n_keys
    g = aggregate_graph(tfs.FloatType, (cell,), "sum")
    out = torch.empty((n_keys, cell), dtype=torch.float32, device="cuda")
    out_keys = torch.empty(n_keys, dtype=torch.int64, device="cuda")
    DA = _native.DeviceArray.from_torch
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        n_groups = s.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(out)}, DA(out_keys))
        again = torch.empty_like(out)
        again_keys = torch.empty_like(out_keys)
        assert s.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(again)}, DA(again_keys)) == n_keys
    assert n_groups == n_keys
    assert torch.equal(out_keys, again_keys) and torch.equal(out, again)          # run-to-run deterministic, order included
    want_keys = torch.sort(torch.arange(n_keys, dtype=torch.int64, device="cuda") * 0x9E3779B97F4A7C15 + 12345).values
    assert torch.equal(out_keys, want_keys)                                        # groups come out in key order
    # rows of key g are i = g * inv(7919) + t * n_keys; n_keys is a multiple of 5, so (i % 5) is the same for all of them
    first = torch.full((n_keys,), -1, dtype=torch.int64, device="cuda")
    first[gid[:n_keys]] = i[:n_keys] % 5                       # the first n_keys rows hit every key once
    want = (first[:, None].to(torch.float32) + (torch.arange(cell, device="cuda") % 3).to(torch.float32)[None, :]) * per
    inv = torch.empty(n_keys, dtype=torch.int64, device="cuda")
    got_gid = ((out_keys - 12345) * pow(0x9E3779B97F4A7C15, -1, 1 << 64)) % n_keys
    inv[got_gid] = torch.arange(n_keys, device="cuda")
    assert torch.equal(out[inv], want)
    # checksum of checksums: the group sums add up to the column sums
    col = torch.zeros(cell, dtype=torch.float64, device="cuda")
    for a in range(0, n, 25_000_000):                          # chunked: the f64 copy of the whole payload would not fit
        col += v[a:a + 25_000_000].sum(0, dtype=torch.float64)
    assert torch.equal(out.sum(0, dtype=torch.float64), col)


def test_full_size_matmul_exact_on_small_integers():
    """BASELINE config 5 at full size (2^24 rows): integers below 2^10 are exact in TF32 and their dot products in the
    fp32 accumulator, so the tensor-core path must be bit-exact against integer arithmetic."""
    torch = pytest.importorskip("torch")
    n, k, m = 1 << 24, 256, 256
    rng = np.random.default_rng(55)
    w = rng.integers(-8, 9, (k, m)).astype(np.float32)
    b = rng.integers(-100, 101, m).astype(np.float32)
    i = torch.arange(n, dtype=torch.int32, device="cuda")
    x = (((i[:, None] * 31 + torch.arange(k, dtype=torch.int32, device="cuda")[None, :] * 17) % 23) - 11).to(torch.float32).contiguous()
    y = torch.empty((n, m), dtype=torch.float32, device="cuda")
    with dsl.withGraph():
        px = dsl.placeholder(tfs.FloatType, dsl.Unknown, k).named("x")
        g = dsl.build_graph((dsl.matmul(px, dsl.constant(w)) + dsl.constant(b)).named("y"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"x": _native.DeviceArray.from_torch(x)}, {"y": _native.DeviceArray.from_torch(y)})
        s.sync()
    wt, bt = torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda()
    for lo in (0, n // 2 - 100_000, n - 300_000):              # slices at both ends and the middle (incl. the last partial tile)
        hi = min(n, lo + 300_000)
        want = (x[lo:hi].double() @ wt.double() + bt.double()).to(torch.float32)
        assert torch.equal(y[lo:hi], want)
    # linearity over the whole block: column sums of y follow from the column sums of x
    xs = torch.zeros(k, dtype=torch.float64, device="cuda")
    ys = torch.zeros(m, dtype=torch.float64, device="cuda")
    for a in range(0, n, 1 << 21):
        xs += x[a:a + (1 << 21)].sum(0, dtype=torch.float64)
        ys += y[a:a + (1 << 21)].sum(0, dtype=torch.float64)
    assert torch.equal(ys, (xs @ wt.double()) + n * bt.double())


# ---- operators: partitions, empties, trimming -------------------------------------------------------
def test_map_blocks_partitions_and_empty_partition():
    df = TensorFrame.from_columns({"x": np.arange(10, dtype=np.float64)}, num_partitions=3)
    df.partitions.insert(1, {"x": np.zeros(0)})            # an empty partition never reaches the device
    with dsl.withGraph():
        z = (tfs.block(df, "x") + 3.0).named("z")
        out = tfs.map_blocks(z, df)
    assert out.columns == ["z", "x"]
    assert [r.z for r in out.collect()] == [float(i) + 3.0 for i in range(10)]
    assert len(out.partitions) == 3


def test_reduce_blocks_over_partitions_matches_oracle():
    rng = np.random.default_rng(41)
    y = rng.uniform(0, 1, (10007, 2))
    df = tfs.analyze(TensorFrame.from_columns({"y": y, "pad": np.arange(10007)}, num_partitions=4))
    df = df.withColumn("z", "y")
    with dsl.withGraph():
        yi, zi = tfs.block(df, "y", "y_input"), tfs.block(df, "z", "z_input")
        s, m = tfs.reduce_blocks([dsl.reduce_sum(yi, [0]).named("y"), dsl.reduce_min(zi, [0]).named("z")], df)
    assert np.all(np.abs(s - tf_ops.exact_sum_f64(y, 0)) <= 1e-12 * np.abs(y).sum(axis=0))
    assert m.tolist() == y.min(axis=0).tolist()


# ---- keyed aggregate (K3) through the C ABI -----------------------------------------------------------
def numpy_group_reduce(keys, v, op):
    uniq, inv = np.unique(keys, return_inverse=True)
    out = np.empty((len(uniq),) + v.shape[1:], dtype=v.dtype)
    absum = np.empty((len(uniq),) + v.shape[1:], dtype=np.float64)
    order = np.argsort(inv, kind="stable")
    bounds = np.searchsorted(inv[order], np.arange(len(uniq) + 1))
    for g in range(len(uniq)):
        rows = v[order[bounds[g]:bounds[g + 1]]]
        if op == "sum":
            if v.dtype.kind == "f":
                out[g] = rows.astype(np.longdouble).sum(axis=0).astype(v.dtype)
            else:
                out[g] = rows.sum(axis=0, dtype=v.dtype)
        elif op == "min":
            out[g] = rows.min(axis=0)
        else:
            out[g] = rows.max(axis=0)
        absum[g] = np.abs(rows.astype(np.float64)).sum(axis=0)
    return uniq, out, absum


def aggregate_graph(st, cell_shape, op):
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, *cell_shape).named("v_input")
        return dsl.build_graph(RED[op](p, [0]).named("v"))


@pytest.mark.parametrize("dt", list(T))
@pytest.mark.parametrize("op", list(RED))
@pytest.mark.parametrize("cell", [(), (1,), (3,), (64,), (100,), (8, 8), (257,)])
def test_aggregate_matches_numpy_groupby(dt, op, cell):
    st = T[dt]
    rng = np.random.default_rng(len(cell) * 7 + 1)
    n, n_keys = 20011, 257
    table = rng.integers(-2**62, 2**62, n_keys)           # sparse 64-bit keys: the hash is really exercised
    table[0] = -2**63                                       # the value the table uses as its empty marker
    keys = table[rng.integers(0, n_keys, n)].astype(np.int64)
    v = rand(rng, st, (n,) + cell)
    g = aggregate_graph(st, cell, op)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got_keys, (got,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n_keys)
    want_keys, want, absum = numpy_group_reduce(keys, v, op)
    order = np.argsort(got_keys)
    assert got_keys[order].tolist() == want_keys.tolist()       # keys bit-exact, each exactly once
    got = got[order]
    if op == "sum" and st.np_dtype.kind == "f":
        err = np.abs(got.astype(np.float64) - want.astype(np.float64))
        assert np.all(err <= TOL[st.np_dtype] * np.maximum(absum, 1e-300))
    else:
        assert_same(got, want)


def test_aggregate_config4_shape_and_group_overflow():
    """BASELINE config 4 at reduced scale: int64 keys, FloatType[?,64] payload, reduce_sum."""
    rng = np.random.default_rng(4)
    n, n_keys = 400_003, 10_000
    table = rng.integers(-2**63, 2**63 - 1, n_keys, dtype=np.int64)
    keys = table[rng.integers(0, n_keys, n)]
    v = rng.uniform(0, 1, (n, 64)).astype(np.float32)
    g = aggregate_graph(tfs.FloatType, (64,), "sum")
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got_keys, (got,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n_keys)
        with pytest.raises(_native.NativeError, match="more than max_groups"):
            s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n_keys // 2)
        again_keys, (again,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n)   # session still usable
    want_keys, want, absum = numpy_group_reduce(keys, v, "sum")
    for k_, o_ in ((got_keys, got), (again_keys, again)):
        order = np.argsort(k_)
        assert k_[order].tolist() == want_keys.tolist()
        assert np.all(np.abs(o_[order].astype(np.float64) - want.astype(np.float64)) <= 1e-5 * absum)


@pytest.mark.timeout(120)
def test_aggregate_far_more_keys_than_the_table_holds():
    """A block with far more distinct keys than max_groups fills the whole hash table (1024 slots for max_groups = 100):
    every warp then has a lane probing a full table.  The call must come back with the overflow error, not spin."""
    n = 1_000_000
    keys = (np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(7)).view(np.int64)      # all distinct (odd multiplier)
    v = np.ones((n, 4), dtype=np.float32)
    g = aggregate_graph(tfs.FloatType, (4,), "sum")
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        for mg in (100, 1000, 5000):
            with pytest.raises(_native.NativeError, match="more than max_groups"):
                s.aggregate(keys, {"v_input": v}, ["v"], max_groups=mg)
        got_keys, (got,) = s.aggregate(keys[:5000], {"v_input": v[:5000]}, ["v"], max_groups=5000)    # session still usable
    assert np.array_equal(np.sort(got_keys), np.sort(keys[:5000])) and np.all(got == 1.0)


def test_aggregate_cut_into_row_ranges_matches_one_pass(monkeypatch):
    """Blocks whose payload spans more than 64 GB are reduced range by range and the partial groups merged (the gather
    pass slows down past that span).  TFS_AGG_CHUNK_MB lowers the threshold so that a small block takes the same path:
    keys and integer results must be identical, float sums within tolerance, also for two fetches of one placeholder."""
    rng = np.random.default_rng(91)
    n, n_keys = 300_007, 777
    table = rng.integers(-2**62, 2**62, n_keys)
    keys = table[rng.integers(0, n_keys, n)].astype(np.int64)
    v = rng.uniform(0, 1, (n, 16)).astype(np.float32)
    w = rng.integers(-1000, 1000, (n, 3)).astype(np.int64)
    with dsl.withGraph():
        pv = dsl.placeholder(tfs.FloatType, dsl.Unknown, 16).named("v_input")
        pw = dsl.placeholder(tfs.LongType, dsl.Unknown, 3).named("w_input")
        g = dsl.build_graph(dsl.reduce_sum(pv, [0]).named("v"), dsl.reduce_max(pw, [0]).named("w"),
                            dsl.reduce_min(pw, [0]).named("w_lo"))
    feeds, names = {"v_input": v, "w_input": w}, ["v", "w", "w_lo"]
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        k1, one = s.aggregate(keys, feeds, names, max_groups=1024)
        monkeypatch.setenv("TFS_AGG_CHUNK_MB", "4")                 # 19 MB of float payload: five ranges
        k2, cut = s.aggregate(keys, feeds, names, max_groups=1024)
        with pytest.raises(_native.NativeError, match="more than max_groups"):
            s.aggregate(keys, feeds, names, max_groups=500)
    assert np.array_equal(k1, k2) and np.array_equal(k1, np.unique(keys))
    assert_same(one[1], cut[1])
    assert_same(one[2], cut[2])
    want_keys, want, absum = numpy_group_reduce(keys, v, "sum")
    assert np.all(np.abs(cut[0].astype(np.float64) - want.astype(np.float64)) <= 1e-5 * absum)


def test_aggregate_general_reducer_graphs(monkeypatch):
    """The reference hands each key's rows to the user's graph (TensorFlowUDAF.compact -> performReduceBlock,
    impl/DebugRowOps.scala:644-656), so a reducer need not be a bare Sum / Min / Max of its placeholder.  Element-wise
    operations before and after the one reduction are taken as a map over the block, the keyed reduction, and a map
    over the reduced rows."""
    rng = np.random.default_rng(123)
    n, n_keys = 50_021, 301
    table = rng.integers(-2**62, 2**62, n_keys)
    keys = table[rng.integers(0, n_keys, n)].astype(np.int64)
    x = rng.uniform(-3, 3, (n, 5))
    y = rng.uniform(0, 1, (n, 7)).astype(np.float32)
    z = rng.integers(-50, 50, n).astype(np.int32)
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 5).named("x_input")
        py = dsl.placeholder(tfs.FloatType, dsl.Unknown, 7).named("y_input")
        pz = dsl.placeholder(tfs.IntegerType, dsl.Unknown).named("z_input")
        fx = dsl.neg(dsl.reduce_max(dsl.neg(px), [0])).named("x")                                   # min, the long way round
        fy = dsl.cast(dsl.reduce_sum(dsl.cast(py, tfs.DoubleType), [0]), tfs.FloatType).named("y")  # float column summed in double
        fz = dsl.reduce_sum(dsl._build("Abs", parents=[pz]), [0]).named("z")                        # pre-map only
        g = dsl.build_graph(fx, fy, fz)
    feeds, names = {"x_input": x, "y_input": y, "z_input": z}, ["x", "y", "z"]
    uniq, inv = np.unique(keys, return_inverse=True)
    want_x = np.full((len(uniq), 5), np.inf)
    np.minimum.at(want_x, inv, x)
    want_y = np.zeros((len(uniq), 7))
    np.add.at(want_y, inv, y.astype(np.float64))
    want_z = np.zeros(len(uniq), dtype=np.int64)
    np.add.at(want_z, inv, np.abs(z).astype(np.int64))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        for chunk_mb in (None, "1"):                # one pass, and cut into row ranges whose partial rows go through the graph again
            if chunk_mb:
                monkeypatch.setenv("TFS_AGG_CHUNK_MB", chunk_mb)
            got_keys, (gx, gy, gz) = s.aggregate(keys, feeds, names, max_groups=512)
            assert np.array_equal(got_keys, uniq)
            assert gx.dtype == np.float64 and np.array_equal(gx, want_x)
            assert gy.dtype == np.float32 and np.array_equal(gy, want_y.astype(np.float32)) or chunk_mb      # one pass: the double sum rounded once
            assert np.all(np.abs(gy.astype(np.float64) - want_y) <= 1e-6 * want_y)
            assert gz.dtype == np.int32 and np.array_equal(gz, want_z.astype(np.int32))
        # what cannot be taken apart is named
        with dsl.withGraph():
            p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("m_input")
            bad = dsl.build_graph(dsl.div(dsl.reduce_sum(p, [0]), dsl.reduce_max(p, [0])).named("m"))
    with _native.Graph(bad) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.UnsupportedError, match="combines two reduced values"):
            s.aggregate(keys, {"m_input": x[:, 0].copy()}, ["m"], max_groups=512)


def test_aggregate_operator_with_a_general_reducer_matches_the_oracle_udaf():
    """tfs.aggregate on a partitioned frame with a reducer that is not a bare reduction, against the oracle's
    restatement of TensorFlowUDAF (10-row buffers, the graph re-applied to compacted rows)."""
    from oracle import executors as ox
    rng = np.random.default_rng(321)
    n = 3000
    key = rng.integers(0, 40, n).astype(np.int64)
    x = rng.integers(-9, 9, (n, 3)).astype(np.float64)
    rows = [(int(key[i]), x[i].tolist()) for i in range(n)]
    df = tfs.analyze(TensorFrame.from_rows(rows, ["key", "x"], num_partitions=3))
    with dsl.withGraph():
        xi = tfs.block(df, "x", tf_name="x_input")
        fetch = dsl.neg(dsl.reduce_max(dsl.neg(xi), [0])).named("x")
        out = tfs.aggregate(fetch, df.groupBy("key"))
        g = dsl.build_graph(fetch)
    kparts, kcols = _golden.oracle_inputs(None, df)
    want = ox.aggregate(kparts, kcols, [0], g, ["x"])
    assert _golden.normalize(out.collect()) == _golden.normalize(want)


def test_aggregate_operator_host_time_is_not_per_row_python(monkeypatch):
    """tfs.aggregate on an int64 key column: the key column goes to the library as it is and the group capacity is
    guessed, not counted on the host.  Host time outside the C call stays far below a per-row Python loop's
    (which took ~1 us per row: 20 s for this block)."""
    import time
    n = 20_000_000
    rng = np.random.default_rng(9)
    df = tfs.analyze(TensorFrame.from_columns({"k": rng.integers(0, 100_000, n).astype(np.int64), "x": np.ones(n, dtype=np.float32)},
                                               num_partitions=2))
    inside = [0.0]
    real = _native.Session.aggregate

    def timed(self, *a, **kw):
        t0 = time.perf_counter()
        try:
            return real(self, *a, **kw)
        finally:
            inside[0] += time.perf_counter() - t0

    monkeypatch.setattr(_native.Session, "aggregate", timed)
    with dsl.withGraph():
        xi = tfs.block(df, "x", tf_name="x_input")
        t0 = time.perf_counter()
        out = tfs.aggregate(dsl.reduce_sum(xi, [0]).named("x"), df.groupBy("k"))
        total = time.perf_counter() - t0
    assert total - inside[0] < 1.0, (total, inside[0])
    part = out.partitions[0]
    assert len(part["k"]) == 100_000 and float(np.sum(part["x"])) == float(n)


def test_aggregate_rejects_non_reducer_graphs():
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x_input")
        g = dsl.build_graph(dsl.identity(p).named("x"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.UnsupportedError, match="must be a Sum, Min or Max"):
            s.aggregate(np.zeros(4, np.int64), {"x_input": np.zeros(4)}, ["x"], max_groups=4)


# ---- linear-chain fast path (kernels_chain.cu): single-fetch graphs, one run per graph -----------------
def _chain_graphs(st):
    """(name, builder) pairs; each builder returns the single fetch node from placeholders a, b."""
    two = st.np_dtype.type(2)
    three = st.np_dtype.type(3)
    gs = [
        ("add_imm", lambda a, b: a + dsl.constant(three)),
        ("rsub_imm", lambda a, b: dsl.constant(three) - a),
        ("sub_imm", lambda a, b: a - dsl.constant(three)),
        ("mul_imm", lambda a, b: a * dsl.constant(two)),
        ("div_imm", lambda a, b: dsl.div(a, dsl.constant(three))),
        ("rdiv_imm", lambda a, b: dsl.div(dsl.constant(three), a)),
        ("min_imm", lambda a, b: dsl.minimum(a, dsl.constant(two))),
        ("rmax_imm", lambda a, b: dsl.maximum(dsl.constant(two), a)),
        ("two_columns", lambda a, b: (a + b) * b - a),
        ("rdiv_column", lambda a, b: dsl.div(b, a)),
        ("unaries", lambda a, b: dsl._build("Abs", parents=[dsl.neg(dsl.relu(dsl._build("Square", parents=[a])))])),
        ("affine_relu", lambda a, b: dsl.maximum(a * dsl.constant(two) + dsl.constant(three), dsl.constant(st.np_dtype.type(0)))),
        ("long_chain", lambda a, b: ((((((a + b) - b) + b) - b) + b) - b) + b),      # 7 slots: interpreter fallback
        ("identity", lambda a, b: dsl.identity(a)),
    ]
    for k, dst in T.items():
        gs.append(("cast_" + k, lambda a, b, dst=dst: dsl.cast(a * dsl.constant(two), dst)))
    if st.np_dtype.kind == "f":
        gs.append(("sqrt_recip", lambda a, b: dsl._build("Reciprocal", parents=[dsl._build("Sqrt", parents=[dsl._build("Abs", parents=[a])])])))
    return gs


@pytest.mark.parametrize("dt", list(T))
def test_chain_fast_path_bit_exact(dt):
    st = T[dt]
    for n in (1, 5, 2049, 100003):
        rng = np.random.default_rng(n)
        a, b = rand(rng, st, n, nonzero=True), rand(rng, st, n, nonzero=True)
        for name, build in _chain_graphs(st):
            with dsl.withGraph():
                pa = dsl.placeholder(st, dsl.Unknown).named("a")
                pb = dsl.placeholder(st, dsl.Unknown).named("b")
                g = dsl.build_graph(build(pa, pb).named("out"))
            feeds = {"a": a, "b": b}
            used = {nd.name for nd in __import__("oracle.graphdef", fromlist=["x"]).parse_graph(g) if nd.op == "Placeholder"}
            feeds = {k: v for k, v in feeds.items() if k in used}
            (got,) = run_native(g, feeds, ["out"])
            (want,) = tf_ops.evaluate(g, feeds, ["out"])
            assert got.dtype == want.dtype and got.tobytes() == want.tobytes(), (dt, n, name)


@pytest.mark.parametrize("dt,cell", [("f4", 64), ("f4", 4), ("f8", 16), ("f8", 2), ("f4", 6), ("f8", 3), ("i4", 8), ("f4", 1024), ("f4", 2048)])
def test_row_vector_operands_bit_exact(dt, cell):
    """x * w + b with [cell] row vectors against a [n, cell] block (per-feature scaling): the chain kernel takes the
    vectors as periodic operands when the cell is a multiple of the vector width and divides its tile, the
    interpreter otherwise.  Both must give numpy's separately rounded result bit for bit."""
    st = {"f4": tfs.FloatType, "f8": tfs.DoubleType, "i4": tfs.IntegerType}[dt]
    npdt = np.dtype("<" + dt)
    rng = np.random.default_rng(cell)
    n = 5003
    if dt == "i4":
        x = rng.integers(-1000, 1000, (n, cell)).astype(npdt)
        w = rng.integers(-50, 50, cell).astype(npdt)
        b = rng.integers(-1000, 1000, cell).astype(npdt)
    else:
        x = rng.uniform(-2, 2, (n, cell)).astype(npdt)
        w = rng.uniform(-2, 2, cell).astype(npdt)
        b = rng.uniform(-2, 2, cell).astype(npdt)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, cell).named("x")
        z = dsl.relu(p * dsl.constant(w) + dsl.constant(b)).named("z")
        r = (dsl.constant(b) - p).named("r")                   # vector on the left: the reversed ops
        g = dsl.build_graph(z, r)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got_z, got_r = s.run({"x": x}, ["z", "r"])
    want_z = np.maximum((x * w).astype(npdt) + b, npdt.type(0)).astype(npdt)
    assert got_z.tobytes() == want_z.tobytes()
    assert got_r.tobytes() == (b - x).astype(npdt).tobytes()


@pytest.mark.parametrize("dt", ["f4", "f8", "i4", "i8"])
def test_straight_line_kernel_small_dags_bit_exact(dt):
    """Graphs the general variant of the straight-line kernel takes over from the interpreter: a saved tile
    ((x+y)*(x-y)), a reload with a second fetch (y2 = x*2, z = x+3), the accumulator as its own operand (x+x), a
    column on the left of a non-commutative op.  Ragged sizes exercise the tails."""
    st = {"f4": tfs.FloatType, "f8": tfs.DoubleType, "i4": tfs.IntegerType, "i8": tfs.LongType}[dt]
    npdt = np.dtype("<" + dt)
    rng = np.random.default_rng(7)
    for n in (1, 5, 4099, 70_001):
        if npdt.kind == "i":
            x = rng.integers(-1000, 1000, n).astype(npdt)
            y = rng.integers(1, 1000, n).astype(npdt)
        else:
            x = rng.uniform(-2, 2, n).astype(npdt)
            y = rng.uniform(1, 2, n).astype(npdt)
        two, three = npdt.type(2), npdt.type(3)
        with dsl.withGraph():
            px = dsl.placeholder(st, dsl.Unknown).named("x")
            py = dsl.placeholder(st, dsl.Unknown).named("y")
            c2, c3 = dsl.constant(two), dsl.constant(three)
            fetches = [((px + py) * (px - py)).named("a"), (px * c2).named("b"), (px + c3).named("c"), (px + px).named("d"),
                       (py - px).named("e"), ((px - py) * (px - py)).named("f")]
            g = dsl.build_graph(*fetches)
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            got = s.run({"x": x, "y": y}, [f.name for f in fetches])
        with np.errstate(over="ignore"):
            want = [((x + y).astype(npdt) * (x - y).astype(npdt)).astype(npdt), (x * two).astype(npdt), (x + three).astype(npdt),
                    (x + x).astype(npdt), (y - x).astype(npdt), ((x - y).astype(npdt) * (x - y).astype(npdt)).astype(npdt)]
        for f, w, gt in zip(fetches, want, got):
            assert gt.tobytes() == w.tobytes(), (f.name, n)


def test_chain_fast_path_unaligned_device_pointers():
    torch = pytest.importorskip("torch")
    n = 50001
    x = np.random.default_rng(0).uniform(-1, 1, n + 1)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g = dsl.build_graph(dsl.cast(dsl.maximum(p * 1.5 + 0.25, 0.0), tfs.FloatType).named("z"))
    dx = torch.from_numpy(x).cuda()
    dz = torch.empty(n + 1, dtype=torch.float32, device="cuda")
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s.run_into({"x": _native.DeviceArray(dx.data_ptr() + 8, tfs.DoubleType, (n,), owner=dx)},
                   {"z": _native.DeviceArray(dz.data_ptr() + 4, tfs.FloatType, (n,), owner=dz)})
        s.sync()
    want = np.maximum(x[1:] * 1.5 + 0.25, 0).astype(np.float32)
    assert dz.cpu().numpy()[1:].tobytes() == want.tobytes()


# ---- MatMul (K4) ---------------------------------------------------------------------------------------
def check_matmul(got, a, b, bias=None):
    """float32 result within 1e-5, normwise: |got - exact| <= 1e-5 * sum_k |a_ik||b_kj| (+|bias|)."""
    exact = a.astype(np.float64) @ b.astype(np.float64)
    scale = np.abs(a).astype(np.float64) @ np.abs(b).astype(np.float64)
    if bias is not None:
        exact = exact + bias.astype(np.float64)
        scale = scale + np.abs(bias).astype(np.float64)
    assert got.dtype == np.float32 and got.shape == exact.shape
    err = np.abs(got.astype(np.float64) - exact)
    assert np.all(err <= 1e-5 * np.maximum(scale, 1e-30)), float((err / np.maximum(scale, 1e-30)).max())


@pytest.mark.parametrize("shape", [(1, 1, 1), (7, 5, 3), (33, 100, 10), (1000, 256, 256), (257, 64, 48), (4099, 256, 256)])
def test_matmul_against_float64(shape):
    m, k, n = shape
    rng = np.random.default_rng(m + k + n)
    a = rng.uniform(0, 1, (m, k)).astype(np.float32)
    w = rng.uniform(-1 / 16, 1 / 16, (k, n)).astype(np.float32)
    bias = rng.uniform(-1, 1, n).astype(np.float32)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, dsl.Unknown, k).named("x")
        mm = dsl.matmul(p, dsl.constant(w)).named("mm")
        y = (dsl.matmul(p, dsl.constant(w)) + dsl.constant(bias)).named("y")
        g = dsl.build_graph(mm, y)
    got_mm, got_y = run_native(g, {"x": a}, ["mm", "y"])
    check_matmul(got_mm, a, w)
    check_matmul(got_y, a, w, bias)
    want = tf_ops.evaluate(g, {"x": a}, ["mm", "y"])
    check_matmul(want[0], a, w)          # the oracle itself meets the same bar


def test_matmul_transposes_and_two_columns():
    rng = np.random.default_rng(8)
    a = rng.normal(size=(50, 20)).astype(np.float32)
    b = rng.normal(size=(30, 20)).astype(np.float32)
    with dsl.withGraph():
        pa = dsl.placeholder(tfs.FloatType, dsl.Unknown, 20).named("a")
        pb = dsl.placeholder(tfs.FloatType, dsl.Unknown, 20).named("b")
        g = dsl.build_graph(dsl.matmul(pa, pb, transpose_b=True).named("abt"),
                            dsl.matmul(pa, pa, transpose_a=True).named("ata"))
    abt, ata = run_native(g, {"a": a, "b": b}, ["abt", "ata"])
    check_matmul(abt, a, b.T)
    check_matmul(ata, a.T, a)


def test_matmul_tensor_core_path_is_fp32_accurate():
    """The 3xTF32 split must behave like a float32 GEMM (error ~1e-7 normwise), far inside the 1e-5 bar;
    a single TF32 pass, or an operand rounded instead of truncated by the tensor core, would be ~1e-4."""
    rng = np.random.default_rng(55)
    m, k, n = 8192, 256, 256
    a = rng.normal(size=(m, k)).astype(np.float32)             # zero-mean data: the hard case for splitting
    w = rng.normal(size=(k, n)).astype(np.float32)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, dsl.Unknown, k).named("x")
        g = dsl.build_graph(dsl.matmul(p, dsl.constant(w)).named("y"))
    (got,) = run_native(g, {"x": a}, ["y"])
    exact = a.astype(np.float64) @ w.astype(np.float64)
    scale = np.abs(a).astype(np.float64) @ np.abs(w).astype(np.float64)
    worst = float((np.abs(got.astype(np.float64) - exact) / scale).max())
    assert worst <= 2e-6, worst


# ---- row mode (mapRows / reduceRows) ---------------------------------------------------------------------
def test_row_mode_cells_reductions_and_constants():
    """tfs_session_run_rows: per-row graphs over vector cells -- element-wise with a constant vector,
    reductions over the whole cell -- against the oracle evaluated row by row."""
    rng = np.random.default_rng(61)
    n = 4001
    x = rng.uniform(-2, 2, (n, 5))
    w = np.asarray([1.0, -2.0, 0.5, 3.0, 0.25])
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, 5).named("x")            # ONE cell
        outs = [(p * dsl.constant(w) + 1.0).named("affine"),
                dsl.reduce_sum(p * p, [0]).named("sumsq"),
                dsl.reduce_min(p, None).named("lo"),
                dsl.cast(dsl.reduce_max(p, [0]), tfs.FloatType).named("hi32")]
        g = dsl.build_graph(*outs)
    names = ["affine", "sumsq", "lo", "hi32"]
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got = s.run({"x": x}, names, rows=True)
    assert [a.shape for a in got] == [(n, 5), (n,), (n,), (n,)]
    for i in (0, 1, n // 2, n - 1):
        want = tf_ops.evaluate(g, {"x": x[i]}, names)
        assert_same(got[0][i], want[0])
        assert abs(got[1][i] - want[1]) <= 1e-12 * float((x[i] * x[i]).sum())
        assert got[2][i] == want[2] and got[3][i] == want[3]
    assert np.all(np.abs(got[1] - (x * x).sum(axis=1)) <= 1e-12 * (x * x).sum(axis=1))


def test_map_rows_operator_with_ragged_cells_matches_oracle():
    rng = np.random.default_rng(62)
    rows = [(rng.uniform(-1, 1, int(k)).tolist(),) for k in rng.integers(1, 6, 200)]
    df = tfs.analyze(TensorFrame.from_rows(rows, ["v"], num_partitions=3))
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("v")
        out = tfs.map_rows([(p * 2.0 + 1.0).named("y"), dsl.reduce_sum(p, [0]).named("s")], df)
    assert out.columns == ["s", "y", "v"]
    for r in out.collect():
        v = np.asarray(r.v)
        assert r.y == (v * 2.0 + 1.0).tolist()
        assert abs(r.s - v.sum()) <= 1e-12 * np.abs(v).sum()


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 1001])
def test_reduce_rows_tree_matches_fold(n):
    rng = np.random.default_rng(n)
    xi = rng.integers(-1000, 1000, (n, 3)).astype(np.int64)
    xf = rng.uniform(0, 1, n)
    df = tfs.analyze(TensorFrame.from_columns({"a": xi, "b": xf}, num_partitions=min(n, 4)))
    # analyze records a lead dimension only when all partitions agree; the cells are what matters here
    with dsl.withGraph():
        a1, a2 = tfs.row(df, "a", "a_1"), tfs.row(df, "a", "a_2")
        b1, b2 = tfs.row(df, "b", "b_1"), tfs.row(df, "b", "b_2")
        a, b = tfs.reduce_rows([(a1 + a2).named("a"), dsl.maximum(b1, b2).named("b")], df)
    assert np.asarray(a).tolist() == xi.sum(axis=0).tolist()
    assert b == xf.max()


def test_row_mode_per_row_operands_broadcast_inside_the_row():
    """A per-row value smaller than the cell (a scalar cell next to a vector cell, a reduction of the cell) broadcasts
    inside its own row -- also when the number of rows equals a cell dimension, where matching the batched operand
    against the trailing dimensions of the block would silently pick s[j] instead of s[i]."""
    rng = np.random.default_rng(63)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType).named("x")
        q = dsl.placeholder(tfs.DoubleType, 3).named("y")
        g = dsl.build_graph((p + q).named("z"))          # scalar cell combined with a vector cell
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        for n in (3, 4, 1000):                            # n == 3: the row count equals the cell size
            x, y = rng.uniform(-1, 1, n), rng.uniform(-1, 1, (n, 3))
            (z,) = s.run({"x": x, "y": y}, ["z"], rows=True)
            assert_same(z, x[:, None] + y)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, 128).named("x")
        m = dsl.placeholder(tfs.FloatType, 2, 128).named("m")
        outs = [dsl.div(p, dsl.reduce_sum(p, [0])).named("norm"),                 # per-row scalar against the cell
                (m - p).named("centered"),                                          # per-row vector against a matrix cell
                (m * dsl.reduce_max(m, None)).named("scaled")]
        g = dsl.build_graph(*outs)
    names = ["norm", "centered", "scaled"]
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        for n in (128, 2, 77):                            # 128 rows of 128-wide cells; 2 rows of [2, 128] cells
            x = rng.uniform(1, 2, (n, 128)).astype(np.float32)
            mm = rng.uniform(-1, 1, (n, 2, 128)).astype(np.float32)
            got = s.run({"x": x, "m": mm}, names, rows=True)
            for i in (0, n // 2, n - 1):
                want = tf_ops.evaluate(g, {"x": x[i], "m": mm[i]}, names)
                assert_same(got[1][i], want[1])
                assert_same(got[2][i], want[2])
                # the row sum differs from the oracle's by summation order only
                assert np.allclose(got[0][i], want[0], rtol=1e-5, atol=0)
            assert np.allclose(got[0].sum(axis=1), 1.0, rtol=1e-5)


def test_row_mode_rejects_what_it_cannot_batch():
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, 2, 1).named("x")
        q = dsl.placeholder(tfs.DoubleType, 2, 3).named("y")
        g = dsl.build_graph((p + q).named("z"))          # a [2, 1] cell against a [2, 3] cell: a middle-dimension broadcast
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.UnsupportedError, match="broadcasting"):
            s.run({"x": np.zeros((4, 2, 1)), "y": np.zeros((4, 2, 3))}, ["z"], rows=True)


# ---- host ingest pipeline (K0): pinned host blocks, chunked H2D / kernel / D2H ------------------------------
def test_pipelined_host_map_matches_one_shot():
    n = 20_000_003                                   # 160 MB of f64: well past the pipelining threshold, ragged tail
    rng = np.random.default_rng(71)
    xa = _native.PinnedBuffer((n,), np.float64)
    ya = _native.PinnedBuffer((n,), np.float64)
    za = _native.PinnedBuffer((n,), np.float32)
    wa = _native.PinnedBuffer((n,), np.float64)
    try:
        xa.array[:] = rng.uniform(-1, 1, n)
        ya.array[:] = rng.uniform(1, 2, n)
        with dsl.withGraph():
            px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
            py = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("y")
            z = dsl.cast(dsl.maximum(px * 1.5 + 0.25, 0.0), tfs.FloatType).named("z")      # chain kernel
            w = ((px + py) * (px - py)).named("w")                                          # interpreter (DAG)
            g = dsl.build_graph(z, w)
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            s.run_into({"x": xa.array, "y": ya.array}, {"z": za.array, "w": wa.array})
            launches = s.launch_count
        assert launches > 2                              # the block really went through in chunks
        x, y = xa.array, ya.array
        assert za.array.tobytes() == np.maximum(x * 1.5 + 0.25, 0.0).astype(np.float32).tobytes()
        assert wa.array.tobytes() == ((x + y) * (x - y)).tobytes()
    finally:
        for b in (xa, ya, za, wa):
            b.free()


def test_pipelined_host_matmul_matches_device_run():
    """A MatMul (+ bias) map over a pinned host block goes through the same chunked pipeline (rows are
    independent); the result must equal the one-shot run on device-resident data bit for bit."""
    torch = pytest.importorskip("torch")
    n, k, m = 700_003, 64, 48                        # 179 MB in: past the pipelining threshold, ragged tail
    rng = np.random.default_rng(72)
    w = rng.uniform(-0.25, 0.25, (k, m)).astype(np.float32)
    b = rng.uniform(-1, 1, m).astype(np.float32)
    xa = _native.PinnedBuffer((n, k), np.float32)
    ya = _native.PinnedBuffer((n, m), np.float32)
    try:
        xa.array[:] = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        with dsl.withGraph():
            px = dsl.placeholder(tfs.FloatType, dsl.Unknown, k).named("x")
            g = dsl.build_graph((dsl.matmul(px, dsl.constant(w)) + dsl.constant(b)).named("y"))
        xd = torch.from_numpy(xa.array).cuda()
        yd = torch.empty((n, m), dtype=torch.float32, device="cuda")
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            s.run_into({"x": xa.array}, {"y": ya.array})
            launches = s.launch_count
            s.run_into({"x": _native.DeviceArray.from_torch(xd)}, {"y": _native.DeviceArray.from_torch(yd)})
            s.sync()
        assert launches > 3                              # weight split + several chunks
        assert ya.array.tobytes() == yd.cpu().numpy().tobytes()
        exact = xa.array[:4096].astype(np.float64) @ w.astype(np.float64) + b
        scale = np.abs(xa.array[:4096]).astype(np.float64) @ np.abs(w).astype(np.float64) + np.abs(b)
        assert np.all(np.abs(ya.array[:4096] - exact) <= 1e-5 * scale)
    finally:
        xa.free()
        ya.free()


def test_arrow_columns_through_map_blocks_and_aggregate():
    """Arrow record batches in, Arrow table out: FixedSizeList blocks go to the device as they are."""
    pa = pytest.importorskip("pyarrow")
    from tensorframes_b200.frame import TensorFrame
    n = 1000
    rng = np.random.default_rng(73)
    v = rng.uniform(-1, 1, (n, 4))
    key = rng.integers(0, 7, n)
    tbl = pa.table({"key": pa.array(key), "v": pa.FixedSizeListArray.from_arrays(pa.array(v.reshape(-1)), 4)})
    df = TensorFrame.from_arrow(pa.Table.from_batches(tbl.slice(0, 300).to_batches() + tbl.slice(300).to_batches()))
    with dsl.withGraph():
        p = tfs.block(df, "v")
        out = tfs.map_blocks((p * 2.0 + 1.0).named("w"), df).to_arrow()
    assert np.asarray(out.column("w").to_pylist()).tobytes() == (v * 2.0 + 1.0).tobytes()
    with dsl.withGraph():
        p = tfs.block(df, "v", tf_name="v_input")
        agg = tfs.aggregate(dsl.reduce_sum(p, [0]).named("v"), df.groupBy("key")).to_arrow()
    got = {k: np.asarray(r) for k, r in zip(agg.column("key").to_pylist(), agg.column("v").to_pylist())}
    for k in range(7):
        np.testing.assert_allclose(got[k], v[key == k].sum(axis=0), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("m_rows,k,n_cols", [(1, 32, 16), (129, 64, 48), (257, 32, 256), (1000, 96, 80), (4099, 256, 256)])
def test_tensor_core_matmul_stays_inside_its_output(m_rows, k, n_cols):
    """Ragged last tiles and partial column boxes: the TMA stores must clip exactly at [M, N].  The output sits
    between two guard bands that have to come back untouched."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(m_rows)
    w = rng.uniform(-0.5, 0.5, (k, n_cols)).astype(np.float32)
    x = torch.from_numpy(rng.uniform(-1, 1, (m_rows, k)).astype(np.float32)).cuda()
    guard = 4096
    buf = torch.full((guard + m_rows * n_cols + guard,), 12345.0, dtype=torch.float32, device="cuda")
    y = buf[guard:guard + m_rows * n_cols].view(m_rows, n_cols)
    with dsl.withGraph():
        px = dsl.placeholder(tfs.FloatType, dsl.Unknown, k).named("x")
        g = dsl.build_graph(dsl.matmul(px, dsl.constant(w)).named("y"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"x": _native.DeviceArray.from_torch(x)}, {"y": _native.DeviceArray(y.data_ptr(), tfs.FloatType, (m_rows, n_cols), owner=buf)})
        s.sync()
    assert torch.all(buf[:guard] == 12345.0) and torch.all(buf[guard + m_rows * n_cols:] == 12345.0)
    exact = x.double() @ torch.from_numpy(w).cuda().double()
    scale = x.double().abs() @ torch.from_numpy(w).cuda().double().abs()
    assert torch.all((y.double() - exact).abs() <= 1e-5 * scale.clamp_min(1e-30))


def test_keyed_and_shape_steps_stay_inside_their_outputs():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(99)
    n, cell, n_keys = 10_007, 12, 37
    keys = torch.from_numpy(rng.integers(0, n_keys, n)).cuda()
    v = torch.from_numpy(rng.integers(-5, 6, (n, cell)).astype(np.float32)).cuda()
    guard = 1024
    obuf = torch.full((guard + n_keys * cell + guard,), -7.0, dtype=torch.float32, device="cuda")
    kbuf = torch.full((guard + n_keys + guard,), -7, dtype=torch.int64, device="cuda")
    DA = _native.DeviceArray
    g = aggregate_graph(tfs.FloatType, (cell,), "sum")
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        ng = s.aggregate_into(DA.from_torch(keys), {"v_input": DA.from_torch(v)},
                              {"v": DA(obuf[guard:].data_ptr(), tfs.FloatType, (n_keys, cell), owner=obuf)},
                              DA(kbuf[guard:].data_ptr(), tfs.LongType, (n_keys,), owner=kbuf))
    assert ng == n_keys
    assert torch.all(obuf[:guard] == -7.0) and torch.all(obuf[guard + n_keys * cell:] == -7.0)
    assert torch.all(kbuf[:guard] == -7) and torch.all(kbuf[guard + n_keys:] == -7)
    assert kbuf[guard:guard + n_keys].tolist() == list(range(n_keys))
    want = torch.zeros((n_keys, cell), dtype=torch.float32, device="cuda").index_add_(0, keys, v)
    assert torch.equal(obuf[guard:guard + n_keys * cell].view(n_keys, cell), want)        # small integers: exact
    # tile / argmin / segment sum into guarded buffers
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, dsl.Unknown, cell).named("v")
        ids = dsl.placeholder(tfs.LongType, dsl.Unknown).named("ids")
        fetches = [dsl.tile(p, [1, 3]).named("t"), dsl.argmin(p, 1).named("a"), dsl.unsorted_segment_sum(p, ids, n_keys).named("s")]
        g2 = dsl.build_graph(*fetches)
    shapes = {"t": ((n, 3 * cell), torch.float32, tfs.FloatType), "a": ((n,), torch.int64, tfs.LongType),
              "s": ((n_keys, cell), torch.float32, tfs.FloatType)}
    bufs, outs = {}, {}
    for nm, (sh, tdt, st) in shapes.items():
        numel = int(np.prod(sh))
        b = torch.full((guard + numel + guard,), 99, dtype=tdt, device="cuda")
        bufs[nm] = (b, numel)
        outs[nm] = DA(b[guard:].data_ptr(), st, sh, owner=b)
    with _native.Graph(g2) as gr, _native.Session(gr, 0) as s:
        torch.cuda.synchronize()      # inputs were produced on torch's stream; the session runs on its own (a NULL handle selects it)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.run_into({"v": DA.from_torch(v), "ids": DA.from_torch(keys)}, outs)
        s.sync()
    for nm, (b, numel) in bufs.items():
        assert torch.all(b[:guard] == 99) and torch.all(b[guard + numel:] == 99), nm
    assert torch.equal(bufs["t"][0][guard:guard + n * 3 * cell].view(n, 3 * cell), v.repeat(1, 3))
    assert torch.equal(bufs["a"][0][guard:guard + n], v.argmin(dim=1))
    assert torch.equal(bufs["s"][0][guard:guard + n_keys * cell].view(n_keys, cell), want)


def test_sessions_on_concurrent_threads():
    """The reference runs N task threads in one JVM (local[N]), each with its own Graph + Session per partition:
    distinct handles must work concurrently, creation / destruction included (the per-device pools are shared)."""
    import threading
    rng = np.random.default_rng(17)
    n_threads, rounds = 6, 12
    xs = [rng.uniform(-1, 1, 50_000 + 37 * t) for t in range(n_threads)]
    keys = [rng.integers(0, 50, len(x)) for x in xs]
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g_map = dsl.build_graph(((px + 1.0) * (px - 1.0)).named("z"), dsl.reduce_sum(px, [0]).named("s"))
    g_agg = aggregate_graph(tfs.DoubleType, (), "sum")
    errors = []

    def work(t):
        try:
            x, k = xs[t], keys[t]
            want_z = (x + 1.0) * (x - 1.0)
            for _ in range(rounds):
                with _native.Graph(g_map) as gr, _native.Session(gr, 0) as s:
                    z, tot = s.run({"x": x}, ["z", "s"])
                assert z.tobytes() == want_z.tobytes()
                assert abs(tot - x.sum()) <= 1e-12 * np.abs(x).sum()
                with _native.Graph(g_agg) as gr, _native.Session(gr, 0) as s:
                    got_k, (got_v,) = s.aggregate(k, {"v_input": x}, ["v"], max_groups=64)
                assert got_k.tolist() == sorted(set(k.tolist()))
                ref = np.array([x[k == kk].sum() for kk in got_k])
                assert np.all(np.abs(got_v - ref) <= 1e-12 * np.abs(x).sum())
        except Exception as exc:      # surfaced in the main thread below
            errors.append((t, repr(exc)))

    threads = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors


@pytest.mark.skipif(_native.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process():
    """A session is bound to the device it was created for; one process may drive several (pools, function
    attributes and constant uploads are per device)."""
    rng = np.random.default_rng(23)
    x = rng.uniform(-1, 1, (300_000, 64)).astype(np.float32)
    w = rng.uniform(-0.2, 0.2, (64, 32)).astype(np.float32)
    y2 = rng.uniform(0, 1, (2_000_000, 2))
    keys = rng.integers(0, 100, len(x))
    with dsl.withGraph():
        px = dsl.placeholder(tfs.FloatType, dsl.Unknown, 64).named("x")
        g_mm = dsl.build_graph(dsl.relu(dsl.matmul(px, dsl.constant(w))).named("y"))
    with dsl.withGraph():
        py = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
        g_red = dsl.build_graph(dsl.reduce_sum(py, [0]).named("y"))
    g_agg = aggregate_graph(tfs.FloatType, (64,), "max")
    results = []
    for dev in (0, 1, 0):
        with _native.Graph(g_mm) as gr, _native.Session(gr, dev) as s:
            (mm,) = s.run({"x": x}, ["y"])
        with _native.Graph(g_red) as gr, _native.Session(gr, dev) as s:
            (red,) = s.run({"y_input": y2}, ["y"])
        with _native.Graph(g_agg) as gr, _native.Session(gr, dev) as s:
            k, (v,) = s.aggregate(keys, {"v_input": x}, ["v"], max_groups=128)
        results.append((mm, red, k, v))
    for mm, red, k, v in results[1:]:
        assert mm.tobytes() == results[0][0].tobytes() and red.tobytes() == results[0][1].tobytes()
        assert k.tobytes() == results[0][2].tobytes() and v.tobytes() == results[0][3].tobytes()
    exact = np.maximum(x.astype(np.float64) @ w.astype(np.float64), 0)
    assert np.all(np.abs(results[0][0] - exact) <= 1e-5 * (np.abs(x).astype(np.float64) @ np.abs(w).astype(np.float64)))


@pytest.mark.skipif(_native.device_count() < 2, reason="needs two GPUs in one process")
def test_memory_of_another_device_is_refused():
    """Feeds, fetch buffers and keys must be host memory or memory of the session's own device (header: 'a CUDA device
    pointer on the session's device'); a pointer into a peer GPU is TFS_ERR_INVALID, never dereferenced."""
    torch = pytest.importorskip("torch")
    DA = _native.DeviceArray.from_torch
    x0 = torch.ones((1024, 64), dtype=torch.float32, device="cuda:0")
    x1 = torch.ones((1024, 64), dtype=torch.float32, device="cuda:1")
    y0 = torch.empty((1024, 64), dtype=torch.float32, device="cuda:0")
    y1 = torch.empty((1024, 64), dtype=torch.float32, device="cuda:1")
    k1 = torch.zeros(1024, dtype=torch.int64, device="cuda:1")
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    with dsl.withGraph():
        px = dsl.placeholder(tfs.FloatType, dsl.Unknown, 64).named("x")
        g = dsl.build_graph(dsl.matmul(px, dsl.constant(np.eye(64, dtype=np.float32))).named("y"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.NativeError, match="lives on device 1"):
            s.run_into({"x": DA(x1)}, {"y": DA(y0)})
        with pytest.raises(_native.NativeError, match="lives on device 1"):
            s.run_into({"x": DA(x0)}, {"y": DA(y1)})
        s.run_into({"x": DA(x0)}, {"y": DA(y0)})
        s.sync()
    assert torch.equal(y0, x0)
    with _native.Graph(aggregate_graph(tfs.FloatType, (64,), "sum")) as gr, _native.Session(gr, 0) as s:
        ok = torch.empty(16, dtype=torch.int64, device="cuda:0")
        out = torch.empty((16, 64), dtype=torch.float32, device="cuda:0")
        with pytest.raises(_native.NativeError, match="lives on device 1"):
            s.aggregate_into(DA(k1), {"v_input": DA(x0)}, {"v": DA(out)}, DA(ok))
    import ctypes
    lib = _native.lib()
    status = lib.tfs_reduce_axis0_noalloc(0, None, tfs.FloatType.tf_code, 0, ctypes.c_void_p(x1.data_ptr()), 1024, 64,
                                          ctypes.c_void_p(y0.data_ptr()))
    assert status == _native.TFS_ERR_INVALID and b"lives on device 1" in lib.tfs_last_error()


def test_blocks_beyond_2_31_elements():
    """int64 sizes end to end: the reference caps a block at Int.MaxValue scalars (impl/datatypes.scala:74-78);
    here a 2.2e9-element column is mapped and reduced in one block each (closed-form checks on the device)."""
    torch = pytest.importorskip("torch")
    n = 2_200_000_000
    x = torch.ones(n, dtype=torch.int32, device="cuda")
    x[n - 1] = 7
    x[2**31 + 5] = -3
    y = torch.empty(n, dtype=torch.int32, device="cuda")
    tot = torch.empty((), dtype=torch.int32, device="cuda")
    with dsl.withGraph():
        p = dsl.placeholder(tfs.IntegerType, dsl.Unknown).named("x")
        g = dsl.build_graph((p * 2 + 1).named("y"), dsl.reduce_sum(p, [0]).named("s"), dsl.reduce_min(p, [0]).named("m"))
    mn = torch.empty((), dtype=torch.int32, device="cuda")
    DA = _native.DeviceArray.from_torch
    torch.cuda.synchronize()      # x was filled on torch's stream; the session runs on its own
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s.run_into({"x": DA(x)}, {"y": DA(y), "s": DA(tot), "m": DA(mn)})
        s.sync()
    assert int(y[0]) == 3 and int(y[n - 1]) == 15 and int(y[2**31 + 5]) == -5 and int(y[2**31 + 6]) == 3
    assert bool((y[:1000] == 3).all()) and bool((y[n - 1000:n - 1] == 3).all())
    expect = (n - 2 + 7 - 3) % (1 << 32)
    expect = expect - (1 << 32) if expect >= (1 << 31) else expect      # int32 sums wrap, as in TF
    assert int(tot) == expect and int(mn) == -3


def test_device_resident_frame_keeps_columns_in_hbm():
    """TensorFrame.to_device: blocks live in GPU memory, map_blocks reads and writes them there, reduce_blocks reads
    them there; results are the host path's, bit for bit."""
    torch = pytest.importorskip("torch")
    from tensorframes_b200.frame import is_device_column
    rng = np.random.default_rng(41)
    x = rng.uniform(-1, 1, (100_003, 8))
    key = rng.integers(0, 11, len(x))
    host = tfs.analyze(TensorFrame.from_columns({"key": key, "x": x}, num_partitions=3))
    dev = host.to_device(0)
    assert all(is_device_column(p["x"]) for p in dev.partitions)

    def pipeline(df):
        with dsl.withGraph():
            p = tfs.block(df, "x")
            df2 = tfs.map_blocks([dsl.relu(p * 2.0 - 0.5).named("y"), dsl.reduce_sum(dsl.square(p), [1]).named("n2")], df)
        with dsl.withGraph():
            q = tfs.block(df2, "y")
            df3 = tfs.map_blocks((q + q).named("z"), df2)
        with dsl.withGraph():
            zi = tfs.block(df3, "z", tf_name="z_input")
            ni = tfs.block(df3, "n2", tf_name="n2_input")
            red = tfs.reduce_blocks([dsl.reduce_sum(zi, [0]).named("z"), dsl.reduce_max(ni, [0]).named("n2")], df3)
        with dsl.withGraph():
            zi = tfs.block(df3, "z", tf_name="z_input")
            agg = tfs.aggregate(dsl.reduce_sum(zi, [0]).named("z"), df3.groupBy("key"))
        return df3, red, agg

    h3, h_red, h_agg = pipeline(host)
    d3, d_red, d_agg = pipeline(dev)
    assert all(is_device_column(p[c]) for p in d3.partitions for c in ("x", "y", "n2", "z"))      # nothing came back
    back = d3.to_host()
    for ph, pd in zip(h3.partitions, back.partitions):
        for c in ("z", "y", "n2", "x", "key"):
            assert np.asarray(pd[c]).tobytes() == np.asarray(ph[c]).tobytes(), c
    for a, b in zip(d_red, h_red):
        assert np.asarray(a).tobytes() == np.asarray(b).tobytes()
    rows_d = sorted(d_agg.collect(), key=lambda r: r.key)
    rows_h = sorted(h_agg.collect(), key=lambda r: r.key)
    assert [r.key for r in rows_d] == [r.key for r in rows_h]
    for a, b in zip(rows_d, rows_h):
        assert np.asarray(a.z).tobytes() == np.asarray(b.z).tobytes()
    for a, b in zip(d3.collect()[:3], h3.collect()[:3]):
        assert np.asarray(a.z).tobytes() == np.asarray(b.z).tobytes() and a.key == b.key


@pytest.mark.parametrize("n_keys,cell", [(1, 16), (10, 16), (300, 5), (4000, 64)])
def test_aggregate_few_large_groups_split_across_warps(n_keys, cell):
    """Few, large groups: every segment is cut into pieces so the whole chip works on it, and the pieces are folded
    in order.  Small-integer payloads make the sums exact whatever the association; min is association-free."""
    rng = np.random.default_rng(n_keys)
    n = 3_000_017
    keys = rng.integers(0, n_keys, n)
    v = rng.integers(-8, 9, (n, cell)).astype(np.float32)
    for op in ("sum", "min"):
        g = aggregate_graph(tfs.FloatType, (cell,), op)
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            k1, (o1,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n_keys)
            k2, (o2,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=n_keys)
        assert k1.tolist() == sorted(set(keys.tolist())) and o1.tobytes() == o2.tobytes()
        order = np.argsort(keys, kind="stable")
        bounds = np.searchsorted(keys[order], k1)
        ref = (np.add if op == "sum" else np.minimum).reduceat(v[order].astype(np.float64), bounds, axis=0)
        assert np.array_equal(o1.astype(np.float64), ref)


@pytest.mark.parametrize("n,n_keys,max_groups", [(5_000_011, 1, 100_000), (5_000_011, 3, 100_000), (5_000_011, 70_000, 70_000),
                                                 (6_000_001, 2_400_000, 2_400_000), (1_000_003, 40_000, 50_000),
                                                 (300_007, 299_000, 300_000)])
def test_aggregate_segments_from_the_partition(n, n_keys, max_groups, monkeypatch):
    """Gather-only blocks with a table too large for shared-memory counters: the rows of a slot are not counted at
    insert, the last partition pass records every slot's segment.  Group counts from 1 to 2.4e6, two and three passes;
    the same call with the rows counted at insert must give the same bytes."""
    rng = np.random.default_rng(n_keys)
    universe = rng.integers(-2**62, 2**62, n_keys)
    universe[0] = -1                                    # the table's empty marker is a legal key
    keys = universe[rng.integers(0, n_keys, n)]
    v = rng.integers(-8, 9, (n, 4)).astype(np.float32)
    g = aggregate_graph(tfs.FloatType, (4,), "sum")
    outs = []
    for mode in ("bounds", "insert"):
        monkeypatch.setenv("TFS_AGG_COUNT", mode)
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            k, (o,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=max_groups)
        outs.append((k, o))
    present = np.unique(keys)
    assert np.array_equal(outs[0][0], present)
    order = np.argsort(keys, kind="stable")
    ref = np.add.reduceat(v[order].astype(np.float64), np.searchsorted(keys[order], present), axis=0)
    assert np.array_equal(outs[0][1].astype(np.float64), ref)
    assert outs[0][1].tobytes() == outs[1][1].tobytes() and np.array_equal(outs[0][0], outs[1][0])


def test_one_host_column_fed_to_two_placeholders():
    """reduce_sum and reduce_min of the same column: the X / X_input contract gives every fetch its own placeholder,
    the caller feeds both from one host buffer, the library stages it once.  Same results as from two separate copies."""
    rng = np.random.default_rng(5)
    x = rng.integers(-1000, 1000, 3_000_001).astype(np.float64)
    with dsl.withGraph():
        a = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("s_input")
        b = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("m_input")
        g = dsl.build_graph(dsl.reduce_sum(a, [0]).named("s"), dsl.reduce_min(b, [0]).named("m"), (a + b).named("t"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s1, m1, t1 = s.run({"s_input": x, "m_input": x}, ["s", "m", "t"])
        s2, m2, t2 = s.run({"s_input": x, "m_input": x.copy()}, ["s", "m", "t"])
        y = x[::-1].copy()
        s3, m3, t3 = s.run({"s_input": x, "m_input": y}, ["s", "m", "t"])
    assert s1 == x.sum() and m1 == x.min() and s1.tobytes() == s2.tobytes() and m1.tobytes() == m2.tobytes()
    assert t1.tobytes() == (x + x).tobytes() == t2.tobytes() and t3.tobytes() == (x + y).tobytes()
    assert s3 == x.sum() and m3 == y.min()


def test_large_pageable_blocks_take_the_staged_copies():
    """numpy arrays (pageable memory) of 64 MiB and more go through the pinned staging ring in both directions;
    ragged sizes, two columns in, two out, plus the keyed path."""
    rng = np.random.default_rng(77)
    n = 12_000_017                                      # 96 MB per f64 column, not a multiple of the 8 MiB chunk
    x = rng.uniform(-1, 1, n)
    y = rng.uniform(1, 2, n)
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        py = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("y")
        g = dsl.build_graph(((px + py) * (px - py)).named("a"), (px * 2.0).named("b"), dsl.reduce_sum(py, [0]).named("s"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        a, b, tot = s.run({"x": x, "y": y}, ["a", "b", "s"])
        a2, b2, _ = s.run({"x": x, "y": y}, ["a", "b", "s"])                # the ring is reused
    assert a.tobytes() == ((x + y) * (x - y)).tobytes() and b.tobytes() == (x * 2.0).tobytes()
    assert a2.tobytes() == a.tobytes() and b2.tobytes() == b.tobytes()
    assert abs(tot - y.sum()) <= 1e-12 * np.abs(y).sum()
    keys = rng.integers(0, 1000, n)
    v = rng.integers(-4, 5, (n, 2)).astype(np.float32)              # 96 MB of keys, 96 MB of payload
    gagg = aggregate_graph(tfs.FloatType, (2,), "sum")
    with _native.Graph(gagg) as gr, _native.Session(gr, 0) as s:
        k, (o,) = s.aggregate(keys, {"v_input": v}, ["v"], max_groups=1000)
    assert k.tolist() == list(range(1000))
    ref = np.zeros((1000, 2))
    np.add.at(ref, keys, v.astype(np.float64))
    assert np.array_equal(o.astype(np.float64), ref)


@pytest.mark.parametrize("src,dst", [("i4", "f8"), ("i8", "f8"), ("f4", "f8"), ("f8", "f4"), ("i4", "f4"), ("i8", "f4")])
def test_chains_that_start_with_a_cast_bit_exact(src, dst):
    """tf.to_double(x) followed by arithmetic (the geometric-mean snippet's shape): the straight-line kernel reads the
    column in its own dtype and computes in the cast's target."""
    names = {"f4": tfs.FloatType, "f8": tfs.DoubleType, "i4": tfs.IntegerType, "i8": tfs.LongType}
    s_np, d_np = np.dtype("<" + src), np.dtype("<" + dst)
    rng = np.random.default_rng(5)
    for n in (3, 4097, 100_003):
        x = (rng.integers(1, 1000, n) if s_np.kind == "i" else rng.uniform(1, 2, n)).astype(s_np)
        y = rng.uniform(1, 2, n).astype(d_np)
        with dsl.withGraph():
            px = dsl.placeholder(names[src], dsl.Unknown).named("x")
            py = dsl.placeholder(names[dst], dsl.Unknown).named("y")
            c = dsl.cast(px, names[dst])
            g = dsl.build_graph(dsl.reciprocal(c * 0.5 + 1.0).named("a"), (dsl.cast(px, names[dst]) - py).named("b"))
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            a, b = s.run({"x": x, "y": y}, ["a", "b"])
        xc = x.astype(d_np)
        one = d_np.type(1)
        want_a = (one / ((xc * d_np.type(0.5)).astype(d_np) + one).astype(d_np)).astype(d_np)
        assert a.dtype == d_np and a.tobytes() == want_a.tobytes(), n
        assert b.tobytes() == (xc - y).astype(d_np).tobytes(), n
    # the same with a row vector as the other operand (a [4] vector is two tiles of an 8-byte chain fed by 8-byte
    # data, one widened tile when the data is 4 bytes wide)
    for n, width in ((5, 4), (3001, 8), (20_011, 64)):
        x = (rng.integers(1, 1000, (n, width)) if s_np.kind == "i" else rng.uniform(1, 2, (n, width))).astype(s_np)
        w = rng.uniform(1, 2, width).astype(d_np)
        with dsl.withGraph():
            px = dsl.placeholder(names[src], dsl.Unknown, width).named("x")
            g = dsl.build_graph((dsl.cast(px, names[dst]) * dsl.constant(w) + 1.0).named("a"))
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            (a,) = s.run({"x": x}, ["a"])
        want = ((x.astype(d_np) * w).astype(d_np) + d_np.type(1)).astype(d_np)
        assert a.tobytes() == want.tobytes(), (n, width)


@pytest.mark.parametrize("dt", list(T))
@pytest.mark.parametrize("op", list(RED))
@pytest.mark.parametrize("xcell", [(), (2,)])
def test_aggregate_scalar_payloads_ride_through_the_sort(dt, op, xcell):
    """groupBy(key).sum(x) with at most 8 bytes per row (a scalar, or two 4-byte values): blocks of 2^16 rows and more
    sort the values along with the slots instead of gathering them; a second, wide fetch in the same call still
    takes the gather path."""
    st = T[dt]
    rng = np.random.default_rng(3)
    n, n_keys = 300_007, 5000
    table = rng.integers(-2**62, 2**62, n_keys, dtype=np.int64)
    keys = table[rng.integers(0, n_keys, n)]
    if st.np_dtype.kind == "i":
        x = rng.integers(-1000, 1000, (n,) + xcell).astype(st.np_dtype)
        w = rng.integers(-1000, 1000, (n, 3)).astype(st.np_dtype)
    else:
        x = rng.uniform(-1, 1, (n,) + xcell).astype(st.np_dtype)
        w = rng.uniform(-1, 1, (n, 3)).astype(st.np_dtype)
    with dsl.withGraph():
        px = dsl.placeholder(st, dsl.Unknown, *xcell).named("x_input")
        pw = dsl.placeholder(st, dsl.Unknown, 3).named("w_input")
        g = dsl.build_graph(RED[op](px, [0]).named("x"), RED[op](pw, [0]).named("w"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        k1, (x1, w1) = s.aggregate(keys, {"x_input": x, "w_input": w}, ["x", "w"], max_groups=n_keys)
        k2, (x2, w2) = s.aggregate(keys, {"x_input": x, "w_input": w}, ["x", "w"], max_groups=n_keys)
    assert x1.tobytes() == x2.tobytes() and w1.tobytes() == w2.tobytes() and k1.tolist() == sorted(set(keys.tolist()))
    for got, data in ((x1, x), (w1, w)):
        want_keys, want, absum = numpy_group_reduce(keys, data, op)
        assert k1.tolist() == want_keys.tolist()
        if st.np_dtype.kind == "i" or op != "sum":
            assert np.array_equal(got, want)
        else:
            tol = 1e-12 if st.np_dtype.itemsize == 8 else 1e-5
            assert np.all(np.abs(got.astype(np.float64) - want.astype(np.float64)) <= tol * absum)


def test_staged_copies_from_concurrent_threads():
    """Several task threads pushing large pageable blocks at once: the staging ring is shared per device and taken in
    turn; results must be intact and nothing may deadlock."""
    import threading
    rng = np.random.default_rng(91)
    n = 9_000_011                                          # 72 MB per f64 column: above the staging threshold
    xs = [rng.uniform(-1, 1, n) for _ in range(3)]
    with dsl.withGraph():
        px = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g = dsl.build_graph((px * 3.0 - 1.0).named("z"), dsl.reduce_max(px, [0]).named("m"))
    errors = []

    def work(t):
        try:
            for _ in range(3):
                with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
                    z, m = s.run({"x": xs[t]}, ["z", "m"])
                assert z.tobytes() == (xs[t] * 3.0 - 1.0).tobytes() and m == xs[t].max()
        except Exception as exc:
            errors.append((t, repr(exc)))

    threads = [threading.Thread(target=work, args=(t,)) for t in range(3)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not any(th.is_alive() for th in threads), "a staged copy did not return"
    assert not errors, errors


def _random_expression(rng, leaves, depth):
    """A random element-wise expression over the given leaf nodes (no division: no NaN / inf to argue about)."""
    if depth == 0 or rng.random() < 0.25:
        return leaves[rng.integers(0, len(leaves))]
    kind = rng.integers(0, 9)
    a = _random_expression(rng, leaves, depth - 1)
    if kind == 0:
        return dsl.neg(a)
    if kind == 1:
        return dsl.relu(a)
    if kind == 2:
        return dsl.square(a)
    if kind == 3:
        return dsl.abs_(a)
    b = _random_expression(rng, leaves, depth - 1)
    return [dsl.add, dsl.sub, dsl.mul, dsl.minimum, dsl.maximum][kind % 5](a, b)


@pytest.mark.parametrize("seed", range(160))
def test_random_elementwise_graphs_match_the_oracle(seed):
    """Randomised graphs over two columns, a row vector and scalars, with one to three fetches, optional casts in
    front and behind: whatever mix of straight-line kernel and interpreter the planner picks, the bits must be the
    oracle's.  (Fixed seeds: the suite stays deterministic.)"""
    rng = np.random.default_rng(1000 + seed)
    dts = [("f4", tfs.FloatType), ("f8", tfs.DoubleType), ("i4", tfs.IntegerType), ("i8", tfs.LongType)]
    name, st = dts[seed % 4]
    npdt = np.dtype("<" + name)
    cell = [1, 4, 6, 64][rng.integers(0, 4)]
    n = int(rng.choice([1, 37, 1025, 50_001]))

    def values(shape):
        if npdt.kind == "i":
            return rng.integers(-9, 10, shape).astype(npdt)
        if seed >= 100:                                                  # general floats: every op is separately rounded
            return rng.uniform(-2, 2, shape).astype(npdt)                #   on both sides, so the bits still agree
        return (rng.integers(-64, 65, shape) / 8.0).astype(npdt)        # exactly representable: products stay finite

    x, y, vec = values((n, cell)), values((n, cell)), values((cell,))
    feeds = {"x": x, "y": y}
    with dsl.withGraph():
        px = dsl.placeholder(st, dsl.Unknown, cell).named("x")
        py = dsl.placeholder(st, dsl.Unknown, cell).named("y")
        lead = px
        if rng.random() < 0.3:                                           # a chain that starts with a cast
            src_name, src_st = dts[(seed + 1) % 4]
            ps = dsl.placeholder(src_st, dsl.Unknown, cell).named("s")
            feeds["s"] = rng.integers(-9, 10, (n, cell)).astype(np.dtype("<" + src_name))
            lead = dsl.cast(ps, st)
        leaves = [lead, px, py, dsl.constant(vec), dsl.constant(npdt.type(2)), dsl.constant(npdt.type(-3))]
        fetches = []
        for k in range(int(rng.integers(1, 4))):
            e = _random_expression(rng, leaves, int(rng.integers(1, 5)))
            if e in leaves:
                e = dsl.add(e, lead)
            if rng.random() < 0.3:
                e = dsl.cast(e, dts[int(rng.integers(0, 4))][1])
            fetches.append(e.named(f"out{k}"))
        g = dsl.build_graph(*fetches)
    names = [f.name for f in fetches]
    with np.errstate(all="ignore"):
        want = tf_ops.evaluate(g, feeds, names)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        present = {nm for nm, _ in gr.nodes()}
        got = s.run({k: v for k, v in feeds.items() if k in present}, names)      # a column the expression never picked is not fed
    for nm, w, gt in zip(names, want, got):
        w = np.broadcast_to(w, gt.shape) if w.shape != gt.shape else w
        assert gt.dtype == w.dtype and gt.tobytes() == np.ascontiguousarray(w).tobytes(), (seed, nm)


@pytest.mark.parametrize("seed", range(60))
def test_random_graphs_with_reductions_match_the_oracle(seed):
    """Random expressions feeding Sum / Min / Max over the rows or the cells, and element-wise nodes that read a
    reduction back (x - min(x), a row vector computed at run time): several fused groups and reduction steps in
    one plan.  Small integer values keep every sum exact in all four dtypes."""
    rng = np.random.default_rng(5000 + seed)
    dts = [("f4", tfs.FloatType), ("f8", tfs.DoubleType), ("i4", tfs.IntegerType), ("i8", tfs.LongType)]
    name, st = dts[seed % 4]
    npdt = np.dtype("<" + name)
    cell = [2, 4, 5, 64][rng.integers(0, 4)]
    n = int(rng.choice([1, 33, 2049, 20_001]))
    x = rng.integers(-3, 4, (n, cell)).astype(npdt)
    y = rng.integers(-3, 4, (n, cell)).astype(npdt)
    reducers = [dsl.reduce_sum, dsl.reduce_min, dsl.reduce_max]
    with dsl.withGraph():
        px = dsl.placeholder(st, dsl.Unknown, cell).named("x")
        py = dsl.placeholder(st, dsl.Unknown, cell).named("y")
        leaves = [px, py, dsl.constant(npdt.type(2))]
        fetches = []
        for k in range(int(rng.integers(1, 4))):
            e = _random_expression(rng, leaves, int(rng.integers(1, 3)))
            if e in leaves or e.shape.numDims < 2:          # constants only: give it the block's shape
                e = dsl.add(e, px)
            kind = rng.integers(0, 4)
            if kind == 0:
                e = reducers[rng.integers(0, 3)](e, [0])                          # block reduction -> [cell]
            elif kind == 1:
                e = reducers[rng.integers(0, 3)](e, [1])                          # per-row reduction -> [n]
            elif kind == 2:
                e = dsl.sub(e, reducers[rng.integers(1, 3)](e, [0]))              # read a reduction back, broadcast
            else:
                e = reducers[rng.integers(0, 3)](dsl.mul(e, e), None)             # everything -> scalar
            fetches.append(e.named(f"out{k}"))
        g = dsl.build_graph(*fetches)
    names = [f.name for f in fetches]
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        present = {nm for nm, _ in gr.nodes()}
        feeds = {k: v for k, v in {"x": x, "y": y}.items() if k in present}
        got = s.run(feeds, names)
    want = tf_ops.evaluate(g, feeds, names)
    for nm, w, gt in zip(names, want, got):
        assert gt.dtype == w.dtype and gt.shape == w.shape, (seed, nm, gt.dtype, w.dtype, gt.shape, w.shape)
        assert gt.tobytes() == np.ascontiguousarray(w).tobytes(), (seed, nm, gt.reshape(-1)[:6], np.asarray(w).reshape(-1)[:6])


@pytest.mark.parametrize("seed", range(48))
def test_random_aggregates_match_numpy(seed):
    """Random keyed reductions: block sizes either side of the carried-values and split-groups thresholds, one to
    three fetches of different cells, keys from a handful to one per row, the empty-marker key value included.
    Small integer payloads make sums exact in every dtype."""
    rng = np.random.default_rng(9000 + seed)
    st = list(T.values())[seed % 4]
    npdt = st.np_dtype
    n = int(rng.choice([1, 7, 1000, 70_000, 300_001]))
    n_keys = int(rng.choice([1, 3, 50, 4000, max(1, n)]))
    table = rng.integers(-2**63, 2**63 - 1, n_keys, dtype=np.int64)
    if rng.random() < 0.3:
        table[0] = np.int64(-2**63)                       # the value the hash table uses as its empty marker
    table = np.unique(table)
    keys = table[rng.integers(0, len(table), n)]
    cells = [(), (1,), (2,), (3,), (16,), (64,), (5, 3)]
    ops = list(RED)
    feeds, fetch_nodes, want = {}, [], {}
    with dsl.withGraph():
        for k in range(int(rng.integers(1, 4))):
            cell = cells[rng.integers(0, len(cells))]
            op = ops[rng.integers(0, len(ops))]
            v = rng.integers(-4, 5, (n,) + cell).astype(npdt)
            p = dsl.placeholder(st, dsl.Unknown, *cell).named(f"c{k}_input")
            fetch_nodes.append(RED[op](p, [0]).named(f"c{k}"))
            feeds[f"c{k}_input"] = v
            want[f"c{k}"] = (v, op)
        g = dsl.build_graph(*fetch_nodes)
    names = [f.name for f in fetch_nodes]
    uniq = np.unique(keys)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got_keys, outs = s.aggregate(keys, feeds, names, max_groups=len(uniq))
    assert got_keys.tolist() == uniq.tolist()
    order = np.argsort(keys, kind="stable")
    bounds = np.searchsorted(keys[order], uniq)
    for nm, got in zip(names, outs):
        v, op = want[nm]
        fn = {"sum": np.add, "min": np.minimum, "max": np.maximum}[op]
        ref = fn.reduceat(v[order].astype(np.int64), bounds, axis=0)
        assert got.shape == ref.shape and np.array_equal(got.astype(np.int64), ref), (seed, nm, op, v.shape)


@pytest.mark.parametrize("seed", range(40))
def test_random_row_mode_graphs_match_the_oracle_row_by_row(seed):
    """mapRows semantics: the graph describes ONE row; the device maps it over the block in one call.  The oracle
    evaluates it row by row, as the reference does (impl/DebugRowOps.scala:819-857)."""
    rng = np.random.default_rng(7000 + seed)
    dts = [("f4", tfs.FloatType), ("f8", tfs.DoubleType), ("i4", tfs.IntegerType), ("i8", tfs.LongType)]
    name, st = dts[seed % 4]
    npdt = np.dtype("<" + name)
    cell = [(), (3,), (4,), (2, 3)][rng.integers(0, 4)]
    n = int(rng.choice([1, 5, 37]))
    x = rng.integers(-3, 4, (n,) + cell).astype(npdt)
    y = rng.integers(-3, 4, (n,) + cell).astype(npdt)
    with dsl.withGraph():
        px = dsl.placeholder(st, *cell).named("x")
        py = dsl.placeholder(st, *cell).named("y")
        leaves = [px, py, dsl.constant(npdt.type(2))]
        fetches = []
        for k in range(int(rng.integers(1, 3))):
            e = _random_expression(rng, leaves, int(rng.integers(1, 4)))
            if e in leaves or e.shape.numDims < len(cell):
                e = dsl.add(e, px)
            if cell and rng.random() < 0.5:
                red = [dsl.reduce_sum, dsl.reduce_min, dsl.reduce_max][rng.integers(0, 3)]
                e = red(e, [0]) if rng.random() < 0.5 else red(e, None)
            fetches.append(e.named(f"out{k}"))
        g = dsl.build_graph(*fetches)
    names = [f.name for f in fetches]
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        present = {nm for nm, _ in gr.nodes()}
        feeds = {k: v for k, v in {"x": x, "y": y}.items() if k in present}
        got = s.run(feeds, names, rows=True)
    for i in range(n):
        want = tf_ops.evaluate(g, {k: v[i] for k, v in feeds.items()}, names)
        for nm, w, gt in zip(names, want, got):
            assert gt.shape[0] == n and gt[i].shape == w.shape, (seed, nm, gt.shape, w.shape)
            assert gt[i].tobytes() == np.ascontiguousarray(w).tobytes(), (seed, nm, i)


@pytest.mark.parametrize("dt", list(T))
def test_middle_axis_reductions(dt):
    """Sum / Min / Max over an axis that is neither first nor last (adjacent axes may be reduced together)."""
    st = T[dt]
    rng = np.random.default_rng(12)
    x = rng.integers(-5, 6, (301, 5, 7)).astype(st.np_dtype)
    x4 = rng.integers(-5, 6, (19, 3, 4, 6)).astype(st.np_dtype)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, 5, 7).named("x")
        p4 = dsl.placeholder(st, dsl.Unknown, 3, 4, 6).named("x4")
        g = dsl.build_graph(dsl.reduce_sum(p, [1]).named("s"), dsl.reduce_min(p, [1]).named("mn"),
                            dsl.reduce_max(p4, [1, 2]).named("mx"), dsl.reduce_sum(p4, [2]).named("s2"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got = s.run({"x": x, "x4": x4}, ["s", "mn", "mx", "s2"])
        with pytest.raises(_native.UnsupportedError, match="adjacent"):
            with dsl.withGraph():
                q = dsl.placeholder(st, dsl.Unknown, 3, 4, 6).named("x4")
                bad = dsl.build_graph(dsl.reduce_sum(q, [1, 3]).named("b"))
            with _native.Graph(bad) as gb, _native.Session(gb, 0) as sb:
                sb.run({"x4": x4}, ["b"])
    want = [x.sum(1, dtype=st.np_dtype), x.min(1), x4.max((1, 2)), x4.sum(2, dtype=st.np_dtype)]
    for gt, w in zip(got, want):
        assert gt.shape == w.shape and np.array_equal(gt, w)


@pytest.mark.parametrize("seed", range(48))
def test_random_matmuls_exact_on_small_integers(seed):
    """Random MatMul shapes, transposes, constant or fed right-hand sides, optional bias and relu, float32 and
    float64: tensor-core, pipelined and SIMT paths.  Small integers make every product and sum exact, so all of them
    must agree with integer arithmetic bit for bit."""
    rng = np.random.default_rng(11000 + seed)
    st = [tfs.FloatType, tfs.DoubleType][seed % 2]
    npdt = st.np_dtype
    m = int(rng.choice([1, 2, 127, 129, 1000, 4097]))
    k = int(rng.choice([1, 3, 17, 32, 64, 96, 100, 256]))
    n = int(rng.choice([1, 5, 16, 48, 100, 256]))
    ta, tb = bool(rng.random() < 0.2), bool(rng.random() < 0.3)
    const_b = bool(rng.random() < 0.7)
    a = rng.integers(-3, 4, (k, m) if ta else (m, k)).astype(npdt)
    b = rng.integers(-3, 4, (n, k) if tb else (k, n)).astype(npdt)
    bias = rng.integers(-9, 10, n).astype(npdt)
    use_bias, use_relu = bool(rng.random() < 0.6), bool(rng.random() < 0.3)
    feeds = {"a": a}
    with dsl.withGraph():
        pa = dsl.placeholder(st, *( (k, dsl.Unknown) if ta else (dsl.Unknown, k) )).named("a")
        if const_b:
            nb = dsl.constant(b)
        else:
            nb = dsl.placeholder(st, *b.shape).named("b")
            feeds["b"] = b
        y = dsl.matmul(pa, nb, transpose_a=ta, transpose_b=tb)
        if use_bias:
            y = y + dsl.constant(bias)
        if use_relu:
            y = dsl.relu(y)
        g = dsl.build_graph(y.named("y"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        (got,) = s.run(feeds, ["y"])
    want = (a.T if ta else a).astype(np.int64) @ (b.T if tb else b).astype(np.int64)
    if use_bias:
        want = want + bias.astype(np.int64)
    if use_relu:
        want = np.maximum(want, 0)
    assert got.dtype == npdt and got.shape == want.shape and np.array_equal(got.astype(np.int64), want), (seed, m, k, n, ta, tb, const_b)


@pytest.mark.parametrize("seed", range(6))
def test_random_graphs_through_the_pinned_pipeline(seed):
    """The chunked H2D / kernel / D2H pipeline on random two-column graphs with ragged row counts: results must be
    the one-shot (pageable) run's bits, fetch by fetch."""
    rng = np.random.default_rng(17000 + seed)
    st, npdt = [(tfs.DoubleType, np.float64), (tfs.FloatType, np.float32), (tfs.LongType, np.int64)][seed % 3]
    cell = [1, 4, 16][rng.integers(0, 3)]
    n = (int(rng.integers(9_000_000, 11_000_000)) * 8 // np.dtype(npdt).itemsize) // cell
    bufs = []
    try:
        xa = _native.PinnedBuffer((n, cell), npdt); bufs.append(xa)
        ya = _native.PinnedBuffer((n, cell), npdt); bufs.append(ya)
        xa.array[:] = rng.integers(-50, 51, (n, cell)).astype(npdt)
        ya.array[:] = rng.integers(1, 9, (n, cell)).astype(npdt)
        vec = rng.integers(-3, 4, cell).astype(npdt)
        with dsl.withGraph():
            px = dsl.placeholder(st, dsl.Unknown, cell).named("x")
            py = dsl.placeholder(st, dsl.Unknown, cell).named("y")
            leaves = [px, py, dsl.constant(vec), dsl.constant(npdt(2))]
            fetches = []
            for k in range(int(rng.integers(1, 3))):
                e = _random_expression(rng, leaves, int(rng.integers(1, 4)))
                if e in leaves or e.shape.numDims < 2:
                    e = dsl.add(e, px)
                fetches.append(e.named(f"out{k}"))
            g = dsl.build_graph(*fetches)
        names = [f.name for f in fetches]
        outs = {}
        for nm in names:
            b = _native.PinnedBuffer((n, cell), npdt); bufs.append(b)
            outs[nm] = b.array
        with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
            present = {nm for nm, _ in gr.nodes()}
            pinned_feeds = {k: v for k, v in {"x": xa.array, "y": ya.array}.items() if k in present}
            s.run_into(pinned_feeds, outs)
            launches = s.launch_count
            pageable = s.run({k: np.array(v) for k, v in pinned_feeds.items()}, names)
        assert launches >= 1
        for nm, ref in zip(names, pageable):
            assert outs[nm].tobytes() == ref.tobytes(), (seed, nm)
    finally:
        for b in bufs:
            b.free()


@pytest.mark.parametrize("seed", range(30))
def test_random_frames_through_the_operators_match_the_oracle_executors(seed):
    """Random partitioned frames (empty partitions, a string column along for the ride, scalar or vector cells)
    through map_blocks, reduce_blocks and aggregate, against the row-by-row restatement of the reference's
    executors (oracle/executors.py: performMap, reduceBlocks with its pairwise merge, TensorFlowUDAF)."""
    from oracle import executors as ox
    rng = np.random.default_rng(19000 + seed)
    st = [tfs.DoubleType, tfs.LongType][seed % 2]        # what from_rows infers for python floats / ints
    npdt = st.np_dtype
    cell = [(), (3,), (2, 2)][rng.integers(0, 3)]
    n = int(rng.choice([1, 4, 23, 200]))
    n_parts = int(rng.integers(1, 6))
    x = rng.integers(-4, 5, (n,) + cell).astype(npdt)
    key = rng.integers(0, 4, n).astype(np.int64)
    tag = np.array([f"r{i % 3}" for i in range(n)], dtype=object)
    rows = [(int(key[i]), tag[i], x[i].tolist() if cell else x[i].item()) for i in range(n)]
    df = tfs.analyze(TensorFrame.from_rows(rows, ["key", "tag", "x"], num_partitions=n_parts))
    parts, cols = _golden.oracle_inputs(None, df)
    # map_blocks: z = f(x), appended in front of the input columns
    with dsl.withGraph():
        p = tfs.block(df, "x")
        z = _random_expression(rng, [p, dsl.constant(npdt.type(2))], int(rng.integers(1, 4)))
        if z.shape.numDims < 1 + len(cell) or z is p:
            z = dsl.add(z, p)
        z = z.named("z")
        out = tfs.map_blocks(z, df)
        g = dsl.build_graph(z)
    names, want_parts = ox.map_blocks(parts, cols, g, ["x"], {"z": (st.tf_code, (dsl.Unknown,) + cell)})
    assert out.columns == names
    assert _golden.normalize(out.collect()) == _golden.normalize([r for pr in want_parts for r in pr])
    # reduce_blocks and aggregate on the numeric column
    with dsl.withGraph():
        xi = tfs.block(df, "x", tf_name="x_input")
        red = [dsl.reduce_sum, dsl.reduce_min, dsl.reduce_max][rng.integers(0, 3)](xi, [0]).named("x")
        got_red = tfs.reduce_blocks(red, df.select("x"))
        got_agg = tfs.aggregate(red, df.select("key", "x").groupBy("key"))
        g2 = dsl.build_graph(red)
    sel_parts, sel_cols = _golden.oracle_inputs(None, df.select("x"))
    want_red = ox.reduce_blocks(sel_parts, sel_cols, g2, ["x"])
    assert np.array_equal(np.asarray(got_red), np.asarray(want_red[0]))
    kparts, kcols = _golden.oracle_inputs(None, df.select("key", "x"))
    want_agg = ox.aggregate(kparts, kcols, [0], g2, ["x"])
    assert _golden.normalize(got_agg.collect()) == _golden.normalize(want_agg)


@pytest.mark.parametrize("seed", range(40))
def test_random_reduction_shapes(seed):
    """Block reductions over random [outer, inner...] shapes: narrow and wide cells, outer from 1 to 1e6, every dtype
    and op; the flat, TMA-staged, column and row kernels all get their turn."""
    rng = np.random.default_rng(23000 + seed)
    dt = list(T)[seed % 4]
    st = T[dt]
    op = list(RED)[rng.integers(0, 3)]
    outer = int(rng.choice([1, 2, 31, 1000, 65_537, 1_000_003]))
    inner = int(rng.choice([1, 2, 3, 7, 64, 100, 257, 2048, 5000]))
    if outer * inner > 40_000_000:
        outer = max(1, 40_000_000 // inner)
    x = rng.integers(-3, 4, (outer, inner)).astype(st.np_dtype)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, inner).named("x")
        g = dsl.build_graph(RED[op](p, [0]).named("c"), RED[op](p, [1]).named("r"), RED[op](p, None).named("a"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        c, r, a = s.run({"x": x}, ["c", "r", "a"])
    fn = {"sum": lambda v, ax: v.sum(ax, dtype=st.np_dtype), "min": lambda v, ax: v.min(ax), "max": lambda v, ax: v.max(ax)}[op]
    assert np.array_equal(c, fn(x, 0)) and np.array_equal(r, fn(x, 1)) and np.array_equal(a, fn(x, None)), (seed, dt, op, outer, inner)


@pytest.mark.parametrize("seed", range(4))
def test_damaged_graphs_plan_and_run_without_faults(seed):
    """1500 byte-level mutations per seed of valid graphs (constants, attributes, node names, wire types all get
    hit).  The ones that still import are planned and every node is fetched: an answer or a status are both fine; a
    device fault (which would poison this process) or a crash is not.  A healthy graph must run right afterwards."""
    import _mutate
    graphs = _mutate.seed_graphs()
    rng = np.random.default_rng(31000 + seed)
    ran = refused = 0
    for it in range(1500):
        g = _mutate.mutate(rng, graphs[it % len(graphs)], structural=False)
        try:
            gr = _native.Graph(g)
        except _native.NativeError:
            continue
        try:
            feeds, fetches = {}, []
            for name, op in gr.nodes():
                dt, shape, is_ph, _ = gr.node_info(name)
                if not is_ph:
                    fetches.append(name)
                    continue
                if shape is None:
                    raise ValueError("placeholder of unknown rank")
                dims = [16 if d < 0 else d for d in shape.dims]
                if int(np.prod(dims)) > 1 << 20:
                    raise ValueError("placeholder too large to feed")
                feeds[name] = rng.integers(0, 5, dims).astype(dt.np_dtype)
            with _native.Session(gr, 0) as s:
                for f in fetches:
                    try:
                        s.run(feeds, [f])
                        ran += 1
                    except _native.NativeError:
                        refused += 1
        except (_native.NativeError, ValueError, UnicodeDecodeError, AttributeError):
            refused += 1
        finally:
            gr.close()
    assert ran > 1000 and refused > 0
    import torch
    torch.cuda.synchronize()
    x = np.arange(12, dtype=np.float32).reshape(3, 4)
    with _native.Graph(graphs[2]) as gr, _native.Session(gr, 0) as s:
        (got,) = s.run({"x": x}, ["s"])
    assert np.array_equal(got, x.sum(0))


def _carve(arr, off):
    """Device copy of arr that starts `off` elements past an aligned address: (DeviceArray, whole buffer, view)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(arr).reshape(-1))
    buf = torch.zeros(t.numel() + off + 64, dtype=t.dtype, device="cuda")
    v = buf[off:off + t.numel()]
    v.copy_(t)
    return _native.DeviceArray(v.data_ptr(), _native.scalar_type_from_numpy(arr.dtype), arr.shape, owner=buf), buf, v


@pytest.mark.parametrize("seed", range(40))
def test_device_buffers_that_are_only_element_aligned(seed):
    """Feeds and outputs 0-3 elements past an aligned address (a slice of a caller's larger buffer): the vector and
    TMA paths must step aside, results stay exact and nothing is written outside the output range."""
    import torch
    rng = np.random.default_rng(33000 + seed)
    dt = list(T)[seed % 4]
    st = T[dt]
    n = int(rng.choice([1, 5, 1000, 65537, 300001]))
    inner = int(rng.choice([1, 2, 3, 8, 64, 100]))
    off_out = int(rng.integers(0, 4))
    x = rng.integers(-3, 4, (n, inner)).astype(st.np_dtype)
    y = rng.integers(-3, 4, (n, inner)).astype(st.np_dtype)
    with dsl.withGraph():
        px = dsl.placeholder(st, dsl.Unknown, inner).named("x")
        py = dsl.placeholder(st, dsl.Unknown, inner).named("y")
        g = dsl.build_graph(((px + py) * px - 2).named("z"), dsl.reduce_sum(px, [0]).named("c"),
                            dsl.reduce_max(px, [1]).named("r"), dsl.reduce_min(py, None).named("a"))
    dx = _carve(x, int(rng.integers(0, 4)))[0]
    dy = _carve(y, int(rng.integers(0, 4)))[0]
    want = {"z": (x + y) * x - st.np_dtype.type(2), "c": x.sum(0, dtype=st.np_dtype), "r": x.max(1), "a": np.asarray(y.min())}
    outs, keep = {}, []
    for name, w in want.items():
        d, buf, v = _carve(np.zeros_like(w), off_out)
        outs[name] = d
        keep.append((name, buf, v, w))
    torch.cuda.synchronize()
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s.run_into({"x": dx, "y": dy}, outs)
        s.sync()
    for name, buf, v, w in keep:
        assert np.array_equal(v.cpu().numpy().reshape(w.shape), w), (name, dt, n, inner, off_out)
        b = buf.cpu().numpy()
        assert not b[:off_out].any() and not b[off_out + w.size:].any(), ("wrote outside", name, dt, n, inner, off_out)


@pytest.mark.parametrize("trial", range(8))
def test_matmul_and_aggregate_on_element_aligned_buffers(trial):
    import torch
    rng = np.random.default_rng(34000 + trial)
    off = trial % 4
    n = [128, 1000, 70000][trial % 3]
    x = rng.integers(-2, 3, (n, 256)).astype(np.float32)
    w = rng.integers(-2, 3, (256, 256)).astype(np.float32)
    b = rng.integers(-2, 3, 256).astype(np.float32)
    with dsl.withGraph():
        px = dsl.placeholder(tfs.FloatType, dsl.Unknown, 256).named("x")
        g = dsl.build_graph((dsl.matmul(px, dsl.constant(w)) + dsl.constant(b)).named("y"))
    dx = _carve(x, off)[0]
    dout, _, v = _carve(np.zeros((n, 256), np.float32), (off + 1) % 4)
    torch.cuda.synchronize()
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s.run_into({"x": dx}, {"y": dout})
        s.sync()
    assert np.array_equal(v.cpu().numpy().reshape(n, 256), x @ w + b)

    dt = list(T)[trial % 4]
    st = T[dt]
    n, cell = [1000, 300001][trial % 2], [1, 2, 3, 64][(trial // 2) % 4]
    keys = rng.integers(0, 50, n).astype(np.int64)
    x = rng.integers(-3, 4, (n, cell)).astype(st.np_dtype)
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, cell).named("x_input")
        g = dsl.build_graph(dsl.reduce_sum(p, [0]).named("x"))
    dk = _carve(keys, trial % 2)[0]
    dx = _carve(x, trial % 4)[0]
    dok, _, vk = _carve(np.zeros(64, np.int64), 1)
    dout, _, vo = _carve(np.zeros((64, cell), st.np_dtype), (trial + 1) % 4)
    torch.cuda.synchronize()
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        ng = s.aggregate_into(dk, {"x_input": dx}, {"x": dout}, dok)
        s.sync()
    uk = np.unique(keys)
    assert ng == len(uk) and np.array_equal(vk.cpu().numpy()[:ng], uk)
    want = np.stack([x[keys == k].sum(0, dtype=st.np_dtype) for k in uk])
    assert np.array_equal(vo.cpu().numpy().reshape(64, cell)[:ng], want)
