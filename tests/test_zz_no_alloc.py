"""The allocation-free block reduction (tfs_reduce_axis0_noalloc), through the C ABI.

Some callers of libtfsb200 cannot let the library allocate: an executor whose GPU memory is all
owned by a pool, a task that must not trigger the device-wide synchronisation of cudaMalloc /
cudaFree, a process on a fixed memory budget.  For them the reduction's scratch is a fixed array
inside the library.  This file checks that entry point against the regular one
(tfs_session_run on `reduce_*(x_input, [0])`) and, independently, against the oracle's evaluation
of the same graph: same bits for integers and Min/Max, and float sums within the tolerance the
other parity tests use (1e-12 f64 / 1e-5 f32 of sum|x|) -- on two calls in a row, because the
array is shared and keeps the previous call's contents.

The file sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It also passes on its own in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import tensorframes_b200 as tfs
from oracle import tf_ops
from tensorframes_b200 import _native, dsl

pytestmark = pytest.mark.gpu

SUM, MIN, MAX = 0, 1, 2
RED = {SUM: dsl.reduce_sum, MIN: dsl.reduce_min, MAX: dsl.reduce_max}
TOL = {np.dtype("float64"): 1e-12, np.dtype("float32"): 1e-5}


def regular_entry_point(st, x, op):
    """The existing path: a session run of reduce_*(x_input, [0])."""
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, *x.shape[1:]).named("x_input")
        g = dsl.build_graph(RED[op](p, [0]).named("x"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        return s.run({"x_input": x}, ["x"])[0]


def oracle_value(st, x, op):
    """The oracle's numpy evaluation of the same reducer graph (oracle/tf_ops.py) -- a check that does not pass
    through any code of the library."""
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, *x.shape[1:]).named("x_input")
        g = dsl.build_graph(RED[op](p, [0]).named("x"))
    return tf_ops.evaluate(g, {"x_input": x}, ["x"])[0]


def no_alloc_entry_point(torch, st, x, op, stream):
    """in / out are buffers the caller owns (torch allocates them here); the call allocates nothing."""
    d_in = torch.from_numpy(x).cuda()
    inner = int(np.prod(x.shape[1:], dtype=np.int64))
    d_out = torch.empty(inner, dtype=d_in.dtype, device="cuda")
    status = _native.lib().tfs_reduce_axis0_noalloc(0, C.c_void_p(stream.cuda_stream), st.tf_code, op,
                                                    C.c_void_p(d_in.data_ptr()), x.shape[0], inner,
                                                    C.c_void_p(d_out.data_ptr()))
    assert status == 0, _native.lib().tfs_last_error()
    stream.synchronize()
    return d_out.cpu().numpy().reshape(x.shape[1:])


def check(got, want, x, op):
    assert got.dtype == want.dtype and got.shape == want.shape
    if op == SUM and x.dtype.kind == "f":
        scale = np.abs(x.astype(np.float64)).sum(axis=0)
        assert np.all(np.abs(got.astype(np.float64) - want.astype(np.float64)) <= TOL[x.dtype] * scale)
    else:
        assert got.tobytes() == want.tobytes()


def test_no_alloc_matches_regular_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    stream = torch.cuda.Stream()
    rng = np.random.default_rng(7)
    # two calls back to back on the same shared scratch: different ops, shapes and dtypes, the second one
    # smaller than the first so that stale partials of the first call are sitting in the array
    first = (tfs.DoubleType, rng.uniform(-1, 1, (300_007, 2)), SUM)
    second = (tfs.FloatType, rng.uniform(0, 1, (1_003, 64)).astype(np.float32), MIN)
    for st, x, op in (first, second):
        got = no_alloc_entry_point(torch, st, x, op, stream)
        check(got, regular_entry_point(st, x, op), x, op)
        check(got, oracle_value(st, x, op), x, op)
    # and once more around the size limit, every dtype
    for st, dt in ((tfs.IntegerType, np.int32), (tfs.LongType, np.int64), (tfs.DoubleType, np.float64)):
        for shape, op in (((5_001, 256), MAX), ((40_001, 3), SUM), ((7,), SUM)):
            x = rng.integers(-1000, 1000, shape).astype(dt)
            got = no_alloc_entry_point(torch, st, x, op, stream)
            check(got, regular_entry_point(st, x, op), x, op)
            check(got, oracle_value(st, x, op), x, op)


def test_no_alloc_rejects_cells_beyond_its_limit():
    torch = pytest.importorskip("torch")
    d = torch.zeros((4, 257), dtype=torch.float64, device="cuda")
    out = torch.zeros(257, dtype=torch.float64, device="cuda")
    lib = _native.lib()
    status = lib.tfs_reduce_axis0_noalloc(0, None, tfs.DoubleType.tf_code, SUM, C.c_void_p(d.data_ptr()), 4, 257,
                                          C.c_void_p(out.data_ptr()))
    assert status == _native.TFS_ERR_UNSUPPORTED
    msg = lib.tfs_last_error().decode()
    assert "257" in msg and "256" in msg and "tfs_session_run" in msg
    assert lib.tfs_reduce_axis0_noalloc(0, None, 4, SUM, C.c_void_p(d.data_ptr()), 4, 2, C.c_void_p(out.data_ptr())) \
        == _native.TFS_ERR_UNSUPPORTED          # dtype code 4 (uint8) is not one of the four


def test_no_alloc_takes_device_memory_only():
    """`in` and `out` must already be on the device: a host pointer is turned down, not copied (copying would need a
    staging allocation, which is what this entry point exists to avoid)."""
    torch = pytest.importorskip("torch")
    host = np.ones((8, 2))
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    lib = _native.lib()
    status = lib.tfs_reduce_axis0_noalloc(0, None, tfs.DoubleType.tf_code, SUM, C.c_void_p(host.ctypes.data), 8, 2,
                                          C.c_void_p(out.data_ptr()))
    assert status == _native.TFS_ERR_INVALID and b"device pointer" in lib.tfs_last_error()
