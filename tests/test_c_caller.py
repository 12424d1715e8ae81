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
