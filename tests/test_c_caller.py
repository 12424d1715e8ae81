"""The boundary really is plain C: include/tfs_b200.h compiles as C99 and a C program (tests/c_caller.c) drives the
library the way a JNI / cgo / Panama binding would.  Without a GPU the program stops after import + analysis."""
import os
import shutil
import subprocess

import pytest

import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GCC = shutil.which("gcc")
pytestmark = pytest.mark.skipif(GCC is None, reason="needs gcc")


def _build(tmp_path):
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        graph_bytes = dsl.build_graph((x + 3.0).named("z"))
    pb = tmp_path / "graph.pb"
    pb.write_bytes(graph_bytes)
    exe = tmp_path / "c_caller"
    libdir = os.path.dirname(_native.LIB_PATH)
    subprocess.run([GCC, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c_caller.c"), "-o", str(exe), "-L", libdir, "-ltfsb200",
                    "-Wl,-rpath," + libdir], check=True)
    return exe, pb


def test_header_is_c99():
    subprocess.run([GCC, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c",
                    os.path.join(ROOT, "include", "tfs_b200.h")], check=True)


def test_struct_layouts_a_binding_is_written_against(tmp_path):
    """INTEGRATION.md's Panama MemoryLayouts (and jvm/.../Native.java) hard-code these sizes and offsets; the ctypes
    binding must agree with the C compiler too."""
    import ctypes
    exe = tmp_path / "c_layout"
    subprocess.run([GCC, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c_layout.c"), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines()
    got = {}
    for line in out:
        w = line.split()
        got[w[0]] = {"size": int(w[1]), **{w[i]: int(w[i + 1]) for i in range(2, len(w), 2)}}
    assert got["tfs_tensor"] == {"size": 88, "name": 0, "dtype": 8, "rank": 12, "dims": 16, "data": 80}
    assert got["tfs_node_info"] == {"size": 80, "dtype": 0, "rank": 4, "dims": 8, "is_placeholder": 72, "n_inputs": 76}
    assert got["tfs_shape_hint"] == {"size": 80, "name": 0, "rank": 8, "dims": 16}
    assert got["tfs_node_summary"] == {"size": 96, "name": 0, "dtype": 8, "rank": 12, "dims": 16, "is_placeholder": 80,
                                       "is_input": 84, "is_output": 88}
    for name, cls in (("tfs_tensor", _native.Tensor), ("tfs_node_info", _native.NodeInfo), ("tfs_shape_hint", _native.ShapeHint),
                      ("tfs_node_summary", _native.NodeSummary)):
        assert ctypes.sizeof(cls) == got[name]["size"], name
        for field, _ in cls._fields_:
            if field in got[name]:
                assert getattr(cls, field).offset == got[name][field], (name, field)


@pytest.mark.skipif(_native.device_count() > 0, reason="the GPU variant below covers the full program")
def test_c_program_imports_and_analyses_without_a_gpu(tmp_path):
    exe, pb = _build(tmp_path)
    out = subprocess.run([str(exe), str(pb), "10"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "import and analysis only" in out.stdout


@pytest.mark.gpu
def test_c_program_runs_the_readme_example(tmp_path):
    exe, pb = _build(tmp_path)
    for n in ("10", "1000003"):
        out = subprocess.run([str(exe), str(pb), n], capture_output=True, text=True)
        assert out.returncode == 0, (out.stdout, out.stderr)
        assert out.stdout.startswith(f"ok: {n} rows")
