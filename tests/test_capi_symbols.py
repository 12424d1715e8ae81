"""The C-ABI library loads and exports exactly what include/tfs_b200.h declares. No compute."""
import ctypes
import os
import re

import numpy as np
import pytest

import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "tfs_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tfs_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _declared() == sorted(_native.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_native.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), name


def test_version_and_error_string():
    lib = _native.lib()
    assert lib.tfs_version() >= 100
    h = ctypes.c_void_p()
    assert lib.tfs_graph_import(b"\xff\xff", 2, ctypes.byref(h)) == _native.TFS_ERR_INVALID
    assert b"GraphDef" in lib.tfs_last_error()
    assert not h


def test_resolve_needs_no_device():
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("x_input")
        g = dsl.build_graph(dsl.reduce_sum(x, [0]).named("x"), dsl.cast(x, tfs.FloatType).named("c"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        metas = s.resolve({"x_input": np.zeros((7, 2))}, ["x", "c"])
        assert metas == [(tfs.DoubleType, (2,)), (tfs.FloatType, (7, 2))]
        with pytest.raises(_native.NativeError, match="not compatible with the data type"):
            s.resolve({"x_input": np.zeros((7, 2), np.float32)}, ["x"])
        with pytest.raises(_native.NativeError, match="not compatible\\) with shape"):
            s.resolve({"x_input": np.zeros((7, 3))}, ["x"])
        with pytest.raises(_native.NativeError, match="was not fed"):
            s.resolve({}, ["x"])


@pytest.mark.skipif(_native.device_count() > 0, reason="checks the no-GPU failure mode")
def test_run_without_gpu_fails_loudly():
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        g = dsl.build_graph((x + 3.0).named("z"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.NativeError) as e:
            s.run({"x": np.arange(10.0)}, ["z"])
        assert e.value.code == _native.TFS_ERR_CUDA and "no CPU execution path" in str(e.value)


def test_product_never_touches_the_oracle():
    """oracle/ is the checker: nothing under tensorframes_b200/ (Python or native source) or examples/ may name it,
    and importing the package must not pull it in."""
    import subprocess
    import sys
    offenders = []
    for top in ("tensorframes_b200", "examples"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cpp", ".h", ".inc", ".c")):
                    text = open(os.path.join(dirpath, f), errors="replace").read()
                    if re.search(r"\boracle\b", text):
                        offenders.append(os.path.join(dirpath, f))
    assert not offenders, offenders
    code = "import sys; import tensorframes_b200; from tensorframes_b200 import ops, distributed, frame; " \
           "print([m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')])"
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, check=True).stdout
    assert out.strip() == "[]", out


def test_resolve_rejects_sizes_that_are_not_numbers():
    """Dimensions arrive as int64: a product that overflows (or a rank that would) is refused at planning time
    instead of wrapping into a small allocation."""
    big = (1 << 32) + 1
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown, dsl.Unknown).named("x")
        g = dsl.build_graph((x + 1.0).named("z"), dsl.reduce_sum(x, [0]).named("s"))
    fake = _native.DeviceArray(256, tfs.DoubleType, (big, big))     # resolve never touches the data
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        with pytest.raises(_native.NativeError, match="too many elements"):
            s.resolve({"x": fake}, ["z"])
        assert s.resolve({"x": _native.DeviceArray(256, tfs.DoubleType, (big, 0))}, ["s"]) == [(tfs.DoubleType, (0,))]
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, 1, 1, 1, 1, 1, 1, 2).named("x")
        g = dsl.build_graph((x + dsl.constant(np.ones((1,) * 8))).named("z"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        assert s.resolve({"x": np.zeros((1,) * 6 + (2,))}, ["z"]) == [(tfs.DoubleType, (1,) * 7 + (2,))]
        with pytest.raises(_native.NativeError, match="rank above the supported maximum"):
            s.resolve({"x": np.zeros((5,) + (1,) * 6 + (2,))}, ["z"], rows=True)


def test_graph_analyze_is_analyzeGraphTF_with_hints():
    """tfs_graph_analyze = impl/TensorFlowOps.scala:101-141: inputs are zero-input Placeholders, outputs the requested
    fetches (":0" accepted), a hinted shape overrides the inferred one, an unknown rank without a hint reads as a scalar."""
    from tensorframes_b200.shape import Shape
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("x_input")
        s = dsl.placeholder(tfs.FloatType).named("scalar_in")
        g = dsl.build_graph(dsl.reduce_sum(x, [0]).named("x"), dsl.cast(x, tfs.FloatType).named("c"), dsl.identity(s).named("t"))
    with _native.Graph(g) as gr:
        res = {r[0]: r for r in gr.analyze({"c": [7, 2], "x_input:0": [7, 2]}, ["x:0", "c", "t"])}
        assert set(res) == {"x_input", "scalar_in", "x", "c", "t"}
        name, dt, shape, is_ph, is_in, is_out = res["x_input"]
        assert dt == tfs.DoubleType and shape == Shape([7, 2]) and is_ph and is_in and not is_out      # hinted under "name:0"
        assert res["c"][2] == Shape([7, 2]) and res["c"][5] and not res["c"][4]                          # hint wins over [?, 2]
        assert res["x"][2] == Shape([2]) and res["x"][1] == tfs.DoubleType                                # inferred
        assert res["scalar_in"][2] == Shape([]) and res["t"][2] == Shape([])
        with pytest.raises(_native.NativeError, match="Requested fetches are not in the graph: nope"):
            gr.analyze({}, ["x", "nope"])
        # capacity smaller than the result: the count is still reported
        out = (_native.NodeSummary * 1)()
        n = ctypes.c_int(0)
        f = (ctypes.c_char_p * 1)(b"x")
        assert _native.lib().tfs_graph_analyze(gr.handle, 0, None, 1, f, 1, out, ctypes.byref(n)) == 0
        assert n.value == 3 and out[0].name in (b"x", b"x_input", b"scalar_in")


def test_pack_and_unpack_rows():
    """tfs_pack_rows / tfs_unpack_rows: one array per row <-> the dense [n, cell] block TFDataOps.convert builds."""
    rng = np.random.default_rng(5)
    for dt, n, cell in ((np.float64, 1000, 3), (np.float32, 7, 1), (np.int32, 50_000, 64), (np.int64, 0, 4)):
        rows = [rng.integers(-100, 100, cell).astype(dt) for _ in range(n)]
        block = np.empty((n, cell), dtype=dt)
        _native.pack_rows(rows, block)
        assert n == 0 or np.array_equal(block, np.stack(rows))
        back = [np.zeros(cell, dtype=dt) for _ in range(n)]
        _native.unpack_rows(block, back)
        assert all(np.array_equal(a, b) for a, b in zip(rows, back))
    lib = _native.lib()
    bad = (ctypes.c_void_p * 2)(0, 0)
    assert lib.tfs_pack_rows(tfs.DoubleType.tf_code, 2, 3, bad, ctypes.c_void_p(np.zeros(6).ctypes.data)) == _native.TFS_ERR_INVALID
    assert b"row 0 is null" in lib.tfs_last_error()


def test_comm_entry_points_fail_cleanly_without_a_clique():
    """No GPU here: the communicator entry points must answer with a status, never crash (NCCL itself is only
    loaded when one of them is called)."""
    lib = _native.lib()
    h = ctypes.c_void_p()
    assert lib.tfs_comm_create(0, 2, 5, b"\0" * 128, ctypes.byref(h)) == _native.TFS_ERR_INVALID      # rank outside [0, world)
    assert not h
    assert lib.tfs_comm_world(None) == 0 and lib.tfs_comm_rank(None) == -1
    lib.tfs_comm_free(None)
    if _native.device_count() == 0:
        st = lib.tfs_comm_create(0, 1, 0, b"\0" * 128, ctypes.byref(h))
        assert st != 0 and lib.tfs_last_error()
