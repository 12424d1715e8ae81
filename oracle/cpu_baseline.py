"""ctypes door to oracle/cpu_ref.c (test infrastructure only): the timed CPU restatement of the
reference's per-block algorithm.  Imported by bench.py's cpu_baseline / --impl reference legs and
by tests; never by the product."""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle_cpu.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            raise ImportError(f"{LIB} is missing: run __graft_entry__.build()")
        l = C.CDLL(LIB)
        l.ref_reduce_blocks_sum_min_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int,
                                                    C.c_void_p, C.c_void_p]
        l.ref_reduce_blocks_sum_min_f64.restype = None
        l.ref_map_blocks_chain_f64.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p]
        l.ref_map_blocks_chain_f64.restype = None
        l.ref_aggregate_sum_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int64,
                                            C.c_void_p, C.c_void_p]
        l.ref_aggregate_sum_f32.restype = C.c_int64
        l.ref_map_blocks_matmul_bias_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                                                     C.c_int, C.c_int, C.c_void_p]
        l.ref_map_blocks_matmul_bias_f32.restype = None
        l.ref_max_threads.restype = C.c_int
        _lib = l
    return _lib


def cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reduce_blocks_sum_min(y: np.ndarray, z: np.ndarray, partitions: int, threads: int):
    assert y.dtype == np.float64 and z.dtype == np.float64 and y.shape == z.shape and y.flags.c_contiguous
    cell = int(np.prod(y.shape[1:], dtype=np.int64))
    s, m = np.empty(cell), np.empty(cell)
    lib().ref_reduce_blocks_sum_min_f64(y.ctypes.data, z.ctypes.data, y.shape[0], cell, partitions, threads,
                                        s.ctypes.data, m.ctypes.data)
    return s.reshape(y.shape[1:]), m.reshape(y.shape[1:])


def map_blocks_chain(x: np.ndarray, a: float, b: float, partitions: int, threads: int) -> np.ndarray:
    assert x.dtype == np.float64 and x.ndim == 1 and x.flags.c_contiguous
    out = np.empty(x.shape[0], dtype=np.float32)
    lib().ref_map_blocks_chain_f64(x.ctypes.data, x.shape[0], partitions, threads, a, b, out.ctypes.data)
    return out


def aggregate_sum(keys: np.ndarray, v: np.ndarray, partitions: int, threads: int, max_groups: int):
    assert keys.dtype == np.int64 and v.dtype == np.float32 and v.flags.c_contiguous
    cell = int(np.prod(v.shape[1:], dtype=np.int64))
    ok = np.empty(max_groups, dtype=np.int64)
    out = np.empty((max_groups, cell), dtype=np.float32)
    g = lib().ref_aggregate_sum_f32(keys.ctypes.data, v.ctypes.data, v.shape[0], cell, partitions, threads, max_groups,
                                    ok.ctypes.data, out.ctypes.data)
    if g < 0:
        raise ValueError("max_groups too small")
    return ok[:g].copy(), out[:g].reshape((g,) + v.shape[1:]).copy()


def matmul_bias(x: np.ndarray, w: np.ndarray, bias: np.ndarray, partitions: int, threads: int) -> np.ndarray:
    assert x.dtype == w.dtype == bias.dtype == np.float32
    out = np.empty((x.shape[0], w.shape[1]), dtype=np.float32)
    lib().ref_map_blocks_matmul_bias_f32(x.ctypes.data, w.ctypes.data, bias.ctypes.data, x.shape[0], x.shape[1],
                                         w.shape[1], partitions, threads, out.ctypes.data)
    return out


def timed(fn, *args, repeat: int = 1):
    best = None
    res = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        res = fn(*args)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, res
