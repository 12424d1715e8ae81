"""ctypes binding of libtfsb200.so (include/tfs_b200.h).

This is the only door between the python host code and the device: if the shared
library is missing the import of the operators fails loudly -- there is no
python/numpy execution path in this package.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .shape import ScalarType, Shape, scalar_type_from_numpy, scalar_type_from_tf

TFS_MAX_RANK = 8
LIB_NAME = "libtfsb200.so"
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)

TFS_OK, TFS_ERR_INVALID, TFS_ERR_UNSUPPORTED, TFS_ERR_CUDA, TFS_ERR_RUNTIME = 0, 1, 2, 3, 4


class NativeError(Exception):
    """A non-zero status from the C ABI; the JVM binding would throw the same way."""

    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


class UnsupportedError(NativeError):
    pass


class NodeInfo(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("rank", C.c_int32), ("dims", C.c_int64 * TFS_MAX_RANK),
                ("is_placeholder", C.c_int32), ("n_inputs", C.c_int32)]


class Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("dtype", C.c_int32), ("rank", C.c_int32),
                ("dims", C.c_int64 * TFS_MAX_RANK), ("data", C.c_void_p)]


class ShapeHint(C.Structure):
    _fields_ = [("name", C.c_char_p), ("rank", C.c_int32), ("dims", C.c_int64 * TFS_MAX_RANK)]


class NodeSummary(C.Structure):
    _fields_ = [("name", C.c_char_p), ("dtype", C.c_int32), ("rank", C.c_int32), ("dims", C.c_int64 * TFS_MAX_RANK),
                ("is_placeholder", C.c_int32), ("is_input", C.c_int32), ("is_output", C.c_int32), ("pad_", C.c_int32)]


# every symbol include/tfs_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "tfs_last_error": (C.c_char_p, []),
    "tfs_version": (C.c_int, []),
    "tfs_device_count": (C.c_int, []),
    "tfs_graph_import": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "tfs_graph_free": (None, [C.c_void_p]),
    "tfs_graph_num_nodes": (C.c_int, [C.c_void_p]),
    "tfs_graph_node_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "tfs_graph_node_op": (C.c_char_p, [C.c_void_p, C.c_int]),
    "tfs_graph_node_info": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(NodeInfo)]),
    "tfs_graph_analyze": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(ShapeHint), C.c_int, C.POINTER(C.c_char_p), C.c_int,
                                    C.POINTER(NodeSummary), C.POINTER(C.c_int)]),
    "tfs_session_create": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]),
    "tfs_session_free": (None, [C.c_void_p]),
    "tfs_session_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tfs_session_sync": (C.c_int, [C.c_void_p]),
    "tfs_session_launch_count": (C.c_int64, [C.c_void_p]),
    "tfs_session_resolve": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Tensor), C.c_int, C.POINTER(Tensor)]),
    "tfs_session_run": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Tensor), C.c_int, C.POINTER(Tensor)]),
    "tfs_session_resolve_rows": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Tensor), C.c_int, C.POINTER(Tensor)]),
    "tfs_session_run_rows": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Tensor), C.c_int, C.POINTER(Tensor)]),
    "tfs_session_aggregate": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.POINTER(Tensor), C.c_int,
                                        C.POINTER(Tensor), C.c_int64, C.c_void_p, C.POINTER(C.c_int64)]),
    "tfs_reduce_axis0_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int64,
                                           C.c_void_p]),
    "tfs_comm_unique_id": (C.c_int, [C.c_void_p]),
    "tfs_comm_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "tfs_comm_create_all": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p)]),
    "tfs_comm_free": (None, [C.c_void_p]),
    "tfs_comm_world": (C.c_int, [C.c_void_p]),
    "tfs_comm_rank": (C.c_int, [C.c_void_p]),
    "tfs_session_reduce_combine": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(Tensor), C.POINTER(Tensor), C.c_int]),
    "tfs_session_combine_wait": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tfs_session_aggregate_combine": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.POINTER(Tensor),
                                                C.POINTER(Tensor), C.c_int64, C.c_void_p, C.POINTER(C.c_int64)]),
    "tfs_pack_rows": (C.c_int, [C.c_int32, C.c_int64, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]),
    "tfs_unpack_rows": (C.c_int, [C.c_int32, C.c_int64, C.c_int64, C.c_void_p, C.POINTER(C.c_void_p)]),
    "tfs_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "tfs_host_free": (C.c_int, [C.c_void_p]),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). tensorframes_b200 has no CPU execution path.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def _check(status: int) -> None:
    if status != TFS_OK:
        msg = lib().tfs_last_error().decode("utf-8", "replace")
        cls = UnsupportedError if status == TFS_ERR_UNSUPPORTED else NativeError
        raise cls(status, msg)


def device_count() -> int:
    return int(lib().tfs_device_count())


COMM_ID_BYTES = 128


def comm_unique_id() -> bytes:
    """The 128-byte rendezvous token rank 0 draws and the caller carries to the other ranks (tfs_comm_unique_id)."""
    buf = (C.c_char * COMM_ID_BYTES)()
    _check(lib().tfs_comm_unique_id(buf))
    return bytes(buf.raw)


class Comm:
    """This rank's end of the exchange steps (tfs_comm_create): one partition per GPU, NCCL behind the C ABI."""

    def __init__(self, device: int, world: int, rank: int, ident: bytes, _handle=None):
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
            return
        if len(ident) != COMM_ID_BYTES:
            raise ValueError(f"communicator id must be {COMM_ID_BYTES} bytes")
        _check(lib().tfs_comm_create(device, world, rank, ident, C.byref(self._h)))

    @staticmethod
    def create_all(devices: Sequence[int]) -> List["Comm"]:
        """The communicators of several devices of this process, created by one thread (tfs_comm_create_all)."""
        n = len(devices)
        out = (C.c_void_p * n)()
        _check(lib().tfs_comm_create_all(n, (C.c_int * n)(*devices), out))
        return [Comm(0, 0, 0, b"", _handle=C.c_void_p(out[i])) for i in range(n)]

    @property
    def handle(self):
        return self._h

    @property
    def world(self) -> int:
        return int(lib().tfs_comm_world(self._h))

    @property
    def rank(self) -> int:
        return int(lib().tfs_comm_rank(self._h))

    def close(self) -> None:
        if self._h:
            lib().tfs_comm_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pack_rows(rows: Sequence[np.ndarray], out: np.ndarray) -> None:
    """Dense [n_rows, cell] block from one array per row (tfs_pack_rows)."""
    st = scalar_type_from_numpy(out.dtype)
    n = len(rows)
    cell = int(out.size // n) if n else 0
    keep = [np.ascontiguousarray(r, dtype=out.dtype) for r in rows]
    ptrs = (C.c_void_p * max(n, 1))(*[k.ctypes.data for k in keep])
    _check(lib().tfs_pack_rows(st.tf_code, n, cell, ptrs, C.c_void_p(out.ctypes.data)))


def unpack_rows(block: np.ndarray, rows: Sequence[np.ndarray]) -> None:
    """The inverse of pack_rows: row i of the block is copied into rows[i] (tfs_unpack_rows)."""
    st = scalar_type_from_numpy(block.dtype)
    n = len(rows)
    cell = int(block.size // n) if n else 0
    ptrs = (C.c_void_p * max(n, 1))(*[r.ctypes.data for r in rows])
    _check(lib().tfs_unpack_rows(st.tf_code, n, cell, C.c_void_p(block.ctypes.data), ptrs))


class DeviceArray:
    """A caller-owned device buffer described without any torch type: pointer, dtype, shape."""

    def __init__(self, ptr: int, dtype: ScalarType, shape: Sequence[int], owner=None):
        self.ptr = int(ptr)
        self.dtype = dtype
        self.shape = tuple(int(d) for d in shape)
        self.owner = owner  # keeps the allocation alive

    @staticmethod
    def from_torch(t) -> "DeviceArray":
        import torch
        assert t.is_cuda and t.is_contiguous()
        st = {torch.float64: "f8", torch.float32: "f4", torch.int32: "i4", torch.int64: "i8"}[t.dtype]
        return DeviceArray(t.data_ptr(), scalar_type_from_numpy(np.dtype("<" + st)), tuple(t.shape), owner=t)


def _fill_tensor(t: Tensor, name: str, dtype_code: int, shape: Sequence[int], ptr: Optional[int]) -> None:
    t.name = name.encode()
    t.dtype = dtype_code
    t.rank = len(shape)
    for i, d in enumerate(shape):
        t.dims[i] = int(d)
    t.data = ptr


class Graph:
    """An imported GraphDef (tfs_graph_import)."""

    def __init__(self, graph_bytes: bytes):
        self._h = C.c_void_p()
        buf = bytes(graph_bytes)
        self._bytes = buf
        _check(lib().tfs_graph_import(buf, len(buf), C.byref(self._h)))

    def close(self) -> None:
        if self._h:
            lib().tfs_graph_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    def nodes(self) -> List[Tuple[str, str]]:
        n = lib().tfs_graph_num_nodes(self._h)
        return [(lib().tfs_graph_node_name(self._h, i).decode(), lib().tfs_graph_node_op(self._h, i).decode())
                for i in range(n)]

    def node_info(self, name: str):
        """(ScalarType, Shape or None for unknown rank, is_placeholder, n_inputs)."""
        info = NodeInfo()
        _check(lib().tfs_graph_node_info(self._h, name.encode(), C.byref(info)))
        shape = None if info.rank < 0 else Shape([info.dims[i] for i in range(info.rank)])
        return scalar_type_from_tf(info.dtype), shape, bool(info.is_placeholder), int(info.n_inputs)

    def analyze(self, hints: Dict[str, Sequence[int]], requested_fetches: Sequence[str]):
        """analyzeGraphTF in the library (tfs_graph_analyze): [(name, ScalarType, Shape, is_placeholder, is_input,
        is_output)] for every input and requested fetch, hinted shapes winning over inferred ones."""
        harr = (ShapeHint * max(1, len(hints)))()
        for i, (name, dims) in enumerate(hints.items()):
            harr[i].name = name.encode()
            harr[i].rank = len(dims)
            for d, v in enumerate(dims):
                harr[i].dims[d] = int(v)
        farr = (C.c_char_p * max(1, len(requested_fetches)))(*[f.encode() for f in requested_fetches])
        cap = lib().tfs_graph_num_nodes(self._h)
        out = (NodeSummary * max(1, cap))()
        n = C.c_int(0)
        _check(lib().tfs_graph_analyze(self._h, len(hints), harr, len(requested_fetches), farr, cap, out, C.byref(n)))
        return [(out[i].name.decode(), scalar_type_from_tf(out[i].dtype), Shape([out[i].dims[d] for d in range(out[i].rank)]),
                 bool(out[i].is_placeholder), bool(out[i].is_input), bool(out[i].is_output)) for i in range(n.value)]


class Session:
    """A graph bound to a device (tfs_session_create); `run` is one Session.run of the reference."""

    def __init__(self, graph: Graph, device: int = 0):
        self._graph = graph
        self._h = C.c_void_p()
        _check(lib().tfs_session_create(graph.handle, device, C.byref(self._h)))

    def close(self) -> None:
        if self._h:
            lib().tfs_session_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_stream(self, cuda_stream: Optional[int]) -> None:
        _check(lib().tfs_session_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def sync(self) -> None:
        _check(lib().tfs_session_sync(self._h))

    @property
    def launch_count(self) -> int:
        return int(lib().tfs_session_launch_count(self._h))

    @staticmethod
    def _feed_array(feeds: Dict[str, object]):
        arr = (Tensor * max(1, len(feeds)))()
        keep = []
        for i, (name, v) in enumerate(feeds.items()):
            if isinstance(v, DeviceArray):
                _fill_tensor(arr[i], name, v.dtype.tf_code, v.shape, v.ptr)
            else:
                a = np.ascontiguousarray(v)
                st = scalar_type_from_numpy(a.dtype)
                if st is None:
                    raise TypeError(f"feed {name}: dtype {a.dtype} is not supported")
                keep.append(a)
                _fill_tensor(arr[i], name, st.tf_code, a.shape, a.ctypes.data if a.size else None)
        return arr, keep

    def resolve(self, feeds: Dict[str, object], fetches: Sequence[str], rows: bool = False) -> List[Tuple[ScalarType, Tuple[int, ...]]]:
        farr, keep = self._feed_array(feeds)
        oarr = (Tensor * max(1, len(fetches)))()
        for i, name in enumerate(fetches):
            _fill_tensor(oarr[i], name, 0, (), None)
        fn = lib().tfs_session_resolve_rows if rows else lib().tfs_session_resolve
        _check(fn(self._h, len(feeds), farr, len(fetches), oarr))
        return [(scalar_type_from_tf(oarr[i].dtype), tuple(oarr[i].dims[d] for d in range(oarr[i].rank)))
                for i in range(len(fetches))]

    def run(self, feeds: Dict[str, object], fetches: Sequence[str], rows: bool = False) -> List[np.ndarray]:
        """Host in, host out: feeds are numpy arrays (or DeviceArray), results are new numpy arrays.
        rows=True maps a per-row graph over the leading dimension of the feeds (tfs_session_run_rows)."""
        metas = self.resolve(feeds, fetches, rows)
        farr, keep = self._feed_array(feeds)
        outs = [np.empty(shape, dtype=st.np_dtype) for st, shape in metas]
        oarr = (Tensor * max(1, len(fetches)))()
        for i, name in enumerate(fetches):
            st, shape = metas[i]
            _fill_tensor(oarr[i], name, st.tf_code, shape, outs[i].ctypes.data if outs[i].size else None)
        fn = lib().tfs_session_run_rows if rows else lib().tfs_session_run
        _check(fn(self._h, len(feeds), farr, len(fetches), oarr))
        return outs

    def run_torch(self, feeds: Dict[str, object], fetches: Sequence[str], device: int):
        """Device in, device out: feeds are torch CUDA tensors (numpy arrays are accepted too and uploaded by the
        call), results are new torch tensors on `device`.  torch only owns the memory; the work is the same
        tfs_session_run.  Inputs produced on torch's current stream are ordered before the run, and the results are
        complete when this returns."""
        import torch
        torch.cuda.current_stream(device).synchronize()
        wrapped = {k: (DeviceArray.from_torch(v.contiguous()) if hasattr(v, "is_cuda") else v) for k, v in feeds.items()}
        metas = self.resolve(wrapped, fetches)
        tdt = {"float64": torch.float64, "float32": torch.float32, "int32": torch.int32, "int64": torch.int64}
        outs = [torch.empty(tuple(shape), dtype=tdt[st.np_dtype.name], device=f"cuda:{device}") for st, shape in metas]
        self.run_into(wrapped, {name: DeviceArray.from_torch(o) for name, o in zip(fetches, outs)})
        self.sync()
        return outs

    def run_into(self, feeds: Dict[str, object], fetches: Dict[str, object]) -> None:
        """Run with caller-provided output buffers (numpy arrays or DeviceArray). Device-only calls
        return as soon as the work is enqueued; use sync()."""
        farr, keep = self._feed_array(feeds)
        oarr, keep2 = self._feed_array(fetches)
        _check(lib().tfs_session_run(self._h, len(feeds), farr, len(fetches), oarr))

    def aggregate(self, keys, feeds: Dict[str, object], fetches: Sequence[str], max_groups: int):
        """Keyed reduction of one block. Returns (keys[int64, n_groups], [per-fetch arrays with n_groups rows])."""
        if isinstance(keys, DeviceArray):
            n_rows, kptr, kkeep = keys.shape[0], keys.ptr, keys
        else:
            kkeep = np.ascontiguousarray(keys, dtype=np.int64)
            n_rows, kptr = int(kkeep.shape[0]), (kkeep.ctypes.data if kkeep.size else None)
        # cell shapes come from running the shape resolution on a 1-row block
        farr, keep = self._feed_array(feeds)
        metas = self.resolve(feeds, fetches)
        outs = [np.empty((max_groups,) + tuple(shape), dtype=st.np_dtype) for st, shape in metas]
        oarr = (Tensor * max(1, len(fetches)))()
        for i, name in enumerate(fetches):
            st, shape = metas[i]
            _fill_tensor(oarr[i], name, st.tf_code, outs[i].shape, outs[i].ctypes.data if outs[i].size else None)
        out_keys = np.empty((max_groups,), dtype=np.int64)
        n_groups = C.c_int64(0)
        _check(lib().tfs_session_aggregate(self._h, n_rows, C.c_void_p(kptr), len(feeds), farr, len(fetches), oarr,
                                           max_groups, C.c_void_p(out_keys.ctypes.data if max_groups else None),
                                           C.byref(n_groups)))
        g = int(n_groups.value)
        return out_keys[:g].copy(), [o[:g].copy() for o in outs]


    def aggregate_into(self, keys: DeviceArray, feeds: Dict[str, object], fetches: Dict[str, object],
                       out_keys: DeviceArray) -> int:
        """Keyed reduction with caller-provided buffers (device or host). fetches[name] has max_groups rows.
        Returns the number of groups."""
        farr, keep = self._feed_array(feeds)
        oarr, keep2 = self._feed_array(fetches)
        n_groups = C.c_int64(0)
        _check(lib().tfs_session_aggregate(self._h, keys.shape[0], C.c_void_p(keys.ptr), len(feeds), farr, len(fetches),
                                           oarr, out_keys.shape[0], C.c_void_p(out_keys.ptr), C.byref(n_groups)))
        return int(n_groups.value)


    # ---- exchange steps over a communicator (one partition per GPU) ----------------------------------------
    def reduce_combine(self, comm: "Comm", partials: Dict[str, DeviceArray], merged: Dict[str, DeviceArray],
                       overlap: bool = False) -> None:
        """The exchange step of reduce_blocks (tfs_session_reduce_combine): `merged` receives the rows reduced over
        every rank's partition.  overlap=True leaves the exchange running on the communicator's stream."""
        parr, keep = self._feed_array(partials)
        marr, keep2 = self._feed_array({k: merged[k] for k in partials})
        _check(lib().tfs_session_reduce_combine(self._h, comm.handle, len(partials), parr, marr, 1 if overlap else 0))

    def combine_wait(self, comm: "Comm") -> None:
        _check(lib().tfs_session_combine_wait(self._h, comm.handle))

    def aggregate_combine(self, comm: "Comm", n_groups: int, keys: DeviceArray, partials: Dict[str, DeviceArray],
                          merged: Dict[str, object], out_keys: DeviceArray) -> int:
        """The exchange step of aggregate (tfs_session_aggregate_combine).  partials[name] holds n_groups rows (a
        larger buffer is fine: only the first n_groups rows are sent).  Returns the number of groups this rank owns."""
        trimmed = {k: DeviceArray(v.ptr, v.dtype, (n_groups,) + tuple(v.shape[1:]), owner=v.owner) for k, v in partials.items()}
        parr, keep = self._feed_array(trimmed)
        marr, keep2 = self._feed_array({k: merged[k] for k in partials})
        n_owned = C.c_int64(0)
        _check(lib().tfs_session_aggregate_combine(self._h, comm.handle, n_groups, C.c_void_p(keys.ptr), len(partials), parr, marr,
                                                   out_keys.shape[0], C.c_void_p(out_keys.ptr), C.byref(n_owned)))
        return int(n_owned.value)


class PinnedBuffer:
    """Page-locked host memory from tfs_host_alloc, viewed as a numpy array."""

    def __init__(self, shape: Sequence[int], dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(int(s) for s in shape)
        nbytes = int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize
        self._p = C.c_void_p()
        _check(lib().tfs_host_alloc(max(nbytes, 1), C.byref(self._p)))
        buf = (C.c_char * max(nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape, dtype=np.int64))).reshape(self.shape)

    def free(self) -> None:
        if self._p:
            self.array = None
            _check(lib().tfs_host_free(self._p))
            self._p = C.c_void_p()
