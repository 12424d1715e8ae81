#!/usr/bin/env python
"""bench.py -- throughput of the TensorFrames block path on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

The default run measures every BASELINE.json performance config at its stated size, one after the other, and
prints ONE JSON line.  The top-level keys are the headline workload, BASELINE configs[1]:

  c2         reduce_blocks, y = reduce_sum(y_input, [0]) and z = reduce_min(z_input, [0]) over two analysed
             DoubleType[?,2] columns of 1e8 rows per GPU                                  (K2, reduce_axis0_tma_kernel)

and `configs` holds one sub-record each, with the same keys (value, ms_per_step, roofline, e2e, cpu_baseline, parity):

  c3         map_blocks, cast_f32(max(x*1.5 + 0.25, 0)) on a 1e9-row float64 column       (K1, chain_kernel)
  reduce_1b  reduce_blocks, reduce_sum + reduce_min on the same 1e9-row float64 column   (K2)
  c4         aggregate, groupBy int64 key (1e6 distinct), reduce_sum FloatType[?,64], 5e8 rows   (K3)
  c5         map_blocks, MatMul FloatType[?,256] x 256x256 + bias, 2^24 rows              (K4, tcgen05)

A "step" is one pass of the block through the graph (N > 1: one partition per GPU, weak scaling, plus the path's
exchange: the combine of the partial rows for reduce_blocks, the key-owner exchange + merge for aggregate).

  value      rows/s, whole job, inputs resident in HBM, timed with CUDA events on the launch stream, max over ranks
  e2e        the same metric through the C ABI with HOST (pinned) column buffers: H2D of the block and D2H of the
             result are inside the timed region
  roofline   the dominant kernel against the measured HBM peak (tensor peak for c5)
  cpu_baseline / --impl reference   oracle/cpu_ref.c, the restatement of the reference's per-block algorithm, on the
             host cores (the real Spark + TF-CPU stack is not installable here); the cpu_baseline leg also checks the
             GPU result of a prefix block against the oracle on the same generated values

Inputs come from one counter-based generator: value(seed, index) = f(splitmix64(seed ^ index)), written
twice below -- numpy for the host legs, torch integer ops for the device -- so that the CPU arm and the GPU arm see the
same numbers (tests/test_bench_generator.py pins the two against each other).

--workload c2|c3|reduce_1b|c4|c5 runs one workload alone (profiling; its record is the line).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

ROWS = {"c2": 100_000_000, "c3": 1_000_000_000, "reduce_1b": 1_000_000_000, "c4": 500_000_000, "c5": 1 << 24}
KEYS_C4 = 1_000_000
ALG_BYTES_PER_ROW = {"c2": 32, "c3": 12, "reduce_1b": 16, "c4": 264, "c5": 2048}
FLOP_PER_ROW_C5 = 131072.0
FALLBACK_HBM_GBS = 6650.0                   # /opt/skills/guides/B200_PROFILING.md
ALL = ["c2", "c3", "reduce_1b", "c5", "c4"]     # run order: the largest block last
METRIC = {
    "c2": "rows/s for reduceBlocks (reduce_sum + reduce_min over DoubleType[?,2])",
    "c3": "rows/s for mapBlocks (fused x*a+b, max(.,0), cast f32 on a float64 column)",
    "reduce_1b": "rows/s for reduceBlocks (reduce_sum + reduce_min over a float64 column)",
    "c4": "rows/s for aggregate (groupBy int64 key, reduce_sum over FloatType[?,64])",
    "c5": "rows/s for mapBlocks (MatMul FloatType[?,256] x 256x256 + bias)",
}
WORKLOAD = {
    "c2": "BASELINE configs[1]: reduce_blocks reduce_sum+reduce_min, DoubleType[?,2], ",
    "c3": "BASELINE configs[2]: map_blocks fused chain cast_f32(max(x*1.5+0.25,0)), float64 column, ",
    "reduce_1b": "north_star target: reduce_blocks reduce_sum+reduce_min on a float64 column, ",
    "c4": "BASELINE configs[3]: aggregate groupBy int64 key (1e6 distinct), reduce_sum FloatType[?,64], ",
    "c5": "BASELINE configs[4]: map_blocks matmul FloatType[?,256] x 256x256 + bias, ",
}
DTYPE = {"c2": "f64", "c3": "f64", "reduce_1b": "f64", "c4": "f32", "c5": "f32 (3xTF32 on tcgen05, fp32 accumulate)"}


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def hbm_peak():
    p = peaks()
    if "hbm_gbs" in p:
        return float(p["hbm_gbs"]), "measured"
    return FALLBACK_HBM_GBS, "fallback"


def ncu_traffic(workload, rows):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/traffic.json), scaled by
    rows when this run's block differs from the captured one.  Returns (bytes or None, source): a capture whose
    summary file is no longer in profiles/ is refused rather than quoted."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f).get(workload)
        if not e:
            return None, None
        if not os.path.exists(os.path.join(ROOT, "profiles", e["source"])):
            return None, f"{e['source']} (missing: not quoted)"
        return e["bytes"] * rows / e["rows"], f"profiles/{e['source']} ({e['kernel']}, captured at {e['rows']} rows)"
    except Exception:
        return None, None


def ncu_extra(workload, key):
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload, {}).get(key)
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------
# The synthetic-column generator, host and device versions of the same arithmetic.
# ---------------------------------------------------------------------------------------------
GAMMA, M1, M2 = 0x9E3779B97F4A7C15, 0xBF58476D1CE4E5B9, 0x94D049BB133111EB
MASK64 = (1 << 64) - 1


def _mix_py(x):
    x = (x + GAMMA) & MASK64
    x = ((x ^ (x >> 30)) * M1) & MASK64
    x = ((x ^ (x >> 27)) * M2) & MASK64
    return x ^ (x >> 31)


def seed_const(seed):
    return _mix_py(seed * 0x632BE59BD9B4E019 & MASK64)


def mix_np(x):
    """splitmix64 of a uint64 array."""
    with np.errstate(over="ignore"):
        x = x + np.uint64(GAMMA)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(M1)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(M2)
        return x ^ (x >> np.uint64(31))


def _i64(c):
    return c - (1 << 64) if c >= (1 << 63) else c


def mix_torch(x):
    """splitmix64 of an int64 tensor holding the bits of a uint64: wrapping multiplies, logical shifts by masking."""
    def lsr(v, k):
        return (v >> k) & ((1 << (64 - k)) - 1)
    x = x + _i64(GAMMA)
    x = (x ^ lsr(x, 30)) * _i64(M1)
    x = (x ^ lsr(x, 27)) * _i64(M2)
    return x ^ lsr(x, 31)


class HostGen:
    """numpy side.  idx: uint64 array of element indices."""

    @staticmethod
    def bits(seed, idx):
        return mix_np(np.uint64(seed_const(seed)) ^ idx)

    @staticmethod
    def u01_f64(seed, idx):
        return (HostGen.bits(seed, idx) >> np.uint64(11)).astype(np.float64) * 2.0 ** -53

    @staticmethod
    def u01_f32(seed, idx):
        return (HostGen.bits(seed, idx) >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)


class DevGen:
    """torch side.  idx: int64 tensor (non-negative)."""

    @staticmethod
    def bits(seed, idx):
        return mix_torch(idx ^ _i64(seed_const(seed)))

    @staticmethod
    def u01_f64(torch, seed, idx):
        h = DevGen.bits(seed, idx)
        return ((h >> 11) & ((1 << 53) - 1)).to(torch.float64) * 2.0 ** -53

    @staticmethod
    def u01_f32(torch, seed, idx):
        h = DevGen.bits(seed, idx)
        return ((h >> 40) & ((1 << 24) - 1)).to(torch.float32) * 2.0 ** -24


def host_columns(wl, row0, rows):
    """The block [row0, row0 + rows) of a workload's columns as numpy arrays."""
    g = np.arange(row0, row0 + rows, dtype=np.uint64)
    if wl == "c2":
        u = HostGen.u01_f64(2, g * np.uint64(2))
        y = np.stack([u, -u], axis=1)
        u = HostGen.u01_f64(2, g * np.uint64(2) + np.uint64(1))
        return {"y_input": y, "z_input": np.stack([u, -u], axis=1)}
    if wl in ("c3", "reduce_1b"):
        return {"x": HostGen.u01_f64(3, g) * 2.0 - 1.0}
    if wl == "c4":
        r = (HostGen.bits(4, g) >> np.uint64(11)) % np.uint64(KEYS_C4)
        keys = mix_np(np.uint64(0x4B455953) ^ r).view(np.int64)
        e = (g[:, None] * np.uint64(64) + np.arange(64, dtype=np.uint64)[None, :])
        return {"keys": keys, "v_input": HostGen.u01_f32(44, e)}
    e = (g[:, None] * np.uint64(256) + np.arange(256, dtype=np.uint64)[None, :])
    return {"x": HostGen.u01_f32(5, e)}


def device_columns(torch, wl, row0, rows, device, chunk=1 << 22):
    """The same block generated on the device, chunk by chunk into preallocated tensors."""
    def rows_of(a, b):
        return torch.arange(row0 + a, row0 + b, dtype=torch.int64, device=device)
    out = {}
    if wl == "c2":
        out = {"y_input": torch.empty((rows, 2), dtype=torch.float64, device=device),
               "z_input": torch.empty((rows, 2), dtype=torch.float64, device=device)}
    elif wl in ("c3", "reduce_1b"):
        out = {"x": torch.empty(rows, dtype=torch.float64, device=device)}
    elif wl == "c4":
        out = {"keys": torch.empty(rows, dtype=torch.int64, device=device),
               "v_input": torch.empty((rows, 64), dtype=torch.float32, device=device)}
    else:
        out = {"x": torch.empty((rows, 256), dtype=torch.float32, device=device)}
    if wl in ("c4", "c5"):
        chunk = 1 << 20
    for a in range(0, rows, chunk):
        b = min(rows, a + chunk)
        g = rows_of(a, b)
        if wl == "c2":
            u = DevGen.u01_f64(torch, 2, g * 2)
            out["y_input"][a:b, 0] = u
            out["y_input"][a:b, 1] = -u
            u = DevGen.u01_f64(torch, 2, g * 2 + 1)
            out["z_input"][a:b, 0] = u
            out["z_input"][a:b, 1] = -u
        elif wl in ("c3", "reduce_1b"):
            out["x"][a:b] = DevGen.u01_f64(torch, 3, g) * 2.0 - 1.0
        elif wl == "c4":
            r = ((DevGen.bits(4, g) >> 11) & ((1 << 53) - 1)) % KEYS_C4
            out["keys"][a:b] = mix_torch(r ^ 0x4B455953)
            e = g[:, None] * 64 + torch.arange(64, dtype=torch.int64, device=device)[None, :]
            out["v_input"][a:b] = DevGen.u01_f32(torch, 44, e)
        else:
            e = g[:, None] * 256 + torch.arange(256, dtype=torch.int64, device=device)[None, :]
            out["x"][a:b] = DevGen.u01_f32(torch, 5, e)
    return out


def c5_constants():
    rng = np.random.default_rng(5)
    return (rng.uniform(-1 / 16, 1 / 16, (256, 256)).astype(np.float32), rng.uniform(-1, 1, 256).astype(np.float32))


def build_graph(wl):
    import tensorframes_b200 as tfs
    from tensorframes_b200 import dsl
    with dsl.withGraph():
        if wl == "c2":
            y = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
            z = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("z_input")
            return dsl.build_graph(dsl.reduce_sum(y, [0]).named("y"), dsl.reduce_min(z, [0]).named("z"))
        if wl == "reduce_1b":
            s = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("s_input")
            m = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("m_input")
            return dsl.build_graph(dsl.reduce_sum(s, [0]).named("s"), dsl.reduce_min(m, [0]).named("m"))
        if wl == "c5":
            w, b = c5_constants()
            x = dsl.placeholder(tfs.FloatType, dsl.Unknown, 256).named("x")
            return dsl.build_graph((dsl.matmul(x, dsl.constant(w)) + dsl.constant(b)).named("y"))
        if wl == "c4":
            v = dsl.placeholder(tfs.FloatType, dsl.Unknown, 64).named("v_input")
            return dsl.build_graph(dsl.reduce_sum(v, [0]).named("v"))
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        return dsl.build_graph(dsl.cast(dsl.maximum(x * 1.5 + 0.25, 0.0), tfs.FloatType).named("z"))


class ClockSampler:
    """SM clock + throttle reasons during the timed region (NVML)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------------------------------------
class Ctx:
    """What every workload needs: torch, the rank layout, the launch stream, the communicator."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from tensorframes_b200 import _native
        self.torch, self.dist, self.native, self.args = torch, dist, _native, args
        self.world = env_int("WORLD_SIZE", 1)
        self.rank = env_int("RANK", 0)
        self.local = env_int("LOCAL_RANK", 0)
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run --nproc-per-node N")
        torch.cuda.set_device(self.local)
        self.device = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.device)
        # a real (non-default) stream: the C ABI treats a NULL stream handle as "use the session's own"
        self.stream = torch.cuda.Stream(device=self.device)
        torch.cuda.set_stream(self.stream)
        self.DA = _native.DeviceArray.from_torch
        self.comm = None
        if self.world > 1:
            # the library's own communicator (NCCL behind the C ABI): rank 0 draws the id, torch.distributed carries it
            ident = torch.zeros(128, dtype=torch.uint8, device=self.device)
            if self.rank == 0:
                ident.copy_(torch.frombuffer(bytearray(_native.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(ident, 0)
            self.comm = _native.Comm(self.local, self.world, self.rank, bytes(ident.cpu().numpy().tobytes()))

    def session(self, wl):
        g = self.native.Graph(build_graph(wl))
        s = self.native.Session(g, self.local)
        s.set_stream(self.stream.cuda_stream)
        return g, s

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step, after_last=None):
        """W warm-up steps, then exactly K steps between barrier + synchronize on both sides; device time of the
        K steps (events on the launch stream), max over ranks; host time spent enqueueing one step."""
        torch, args = self.torch, self.args
        for _ in range(args.warmup):
            step(False)
        if after_last:
            after_last()
        self.barrier()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(self.local) as clocks:
            self.barrier()
            start.record(self.stream)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                step(True)
            host_ms = (time.perf_counter() - t0) * 1e3 / args.steps
            if after_last:
                after_last()            # the last block's exchange belongs to the timed region
            end.record(self.stream)
            self.barrier()
        return self.max_over_ranks(start.elapsed_time(end)), clocks.summary(), host_ms

    def e2e_timed(self, step, k):
        """k end-to-end steps (host buffers in, host results out, synchronous), wall clock, max over ranks."""
        step()
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            step()
        self.barrier()
        return self.max_over_ranks(time.perf_counter() - t0)

    def pinned(self, shape, dtype, like=None):
        pb = self.native.PinnedBuffer(tuple(shape), dtype)
        if like is not None:
            self.torch.from_numpy(pb.array).copy_(like)      # fill the pinned block once, outside the timing
        return pb


def record(ctx, wl, rows, elapsed_ms, clocks, launches, roofline, e2e, cpu, parity, extra_config=None, **more):
    args = ctx.args
    step_s = elapsed_ms / 1e3 / args.steps
    cfg = {"workload": WORKLOAD[wl] + f"{rows} rows per GPU", "rows_per_gpu": rows, "partitions": ctx.world,
           "l2": "inputs larger than L2 (no flush needed)"}
    cfg.update(extra_config or {})
    rec = {"metric": METRIC[wl], "value": ctx.world * rows / step_s, "unit": "rows/s", "n_gpus": ctx.world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": DTYPE[wl], "data": "synthetic (counter-based generator, SURVEY 8(d))",
           "config": cfg, "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
           "cpu_baseline": cpu, "parity": parity}
    rec.update(more)
    return rec


def hbm_roofline(wl, rows, bytes_per_launch, avg_launch_s, kernel):
    peak, src = hbm_peak()
    achieved = bytes_per_launch / avg_launch_s / 1e9
    traffic, tsrc = ncu_traffic("c2" if wl == "reduce_1b" else wl, rows if wl != "reduce_1b" else rows // 2)
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "traffic_source": tsrc, "peak_source": src, "kernel": kernel,
            "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_launch_s * 1e3,
            "frac_of_nominal_8000": achieved / 8000.0}


def cpu_sample(fn, rows_avail, probe_rows, budget_s=10.0):
    """Times fn(rows) on a bounded prefix: a probe sizes the sample so that 3 repeats take about budget_s."""
    from oracle import cpu_baseline as cb
    probe = min(rows_avail, probe_rows)
    t, _ = cb.timed(fn, probe)
    sample = int(min(rows_avail, max(probe, budget_s / max(t, 1e-6) * probe / 3)))
    t, res = cb.timed(fn, sample, repeat=3)
    return sample, t, res


# ---------------------------------------------------------------------------------------------
# reduce_blocks: c2 (two f64[?,2] columns) and reduce_1b (sum + min over one 1e9-row f64 column)
# ---------------------------------------------------------------------------------------------
def run_reduce(ctx, wl, rows, cols=None):
    torch, dist, DA, args = ctx.torch, ctx.dist, ctx.DA, ctx.args
    world, device, stream = ctx.world, ctx.device, ctx.stream
    graph, sess = ctx.session(wl)
    if wl == "c2":
        feeds_t = cols or device_columns(torch, "c2", ctx.rank * rows, rows, device)
        names, cell = ["y", "z"], (2,)
        sum_name, min_name = "y", "z"
    else:
        x = (cols or device_columns(torch, "c3", ctx.rank * rows, rows, device))["x"]
        feeds_t = {"s_input": x, "m_input": x}
        names, cell = ["s", "m"], ()
        sum_name, min_name = "s", "m"
    torch.cuda.synchronize()
    feeds = {k: DA(v) for k, v in feeds_t.items()}
    outs_t = {n: torch.empty(cell, dtype=torch.float64, device=device) for n in names}
    merged_t = {n: torch.empty(cell, dtype=torch.float64, device=device) for n in names}
    outs = {k: DA(v) for k, v in outs_t.items()}
    merged = {k: DA(v) for k, v in merged_t.items()}
    kernel_events = []

    def step(timed):
        if timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        sess.run_into(feeds, outs)
        if timed:
            e1.record(stream)
            kernel_events.append((e0, e1))
        if world > 1:
            # the path's exchange step, behind the C ABI: the partial rows are combined across ranks on the
            # communicator's side stream, so it overlaps the next block's reduction kernels
            sess.reduce_combine(ctx.comm, outs, merged, overlap=True)

    def drain():
        if world > 1:
            sess.combine_wait(ctx.comm)  # launch stream waits for the side stream's last exchange

    launches0 = sess.launch_count
    elapsed_ms, clocks, host_ms = ctx.timed(step, drain)
    launches = (sess.launch_count - launches0) * args.steps // (args.steps + args.warmup)
    kernel_ms = sum(a.elapsed_time(b) for a, b in kernel_events)

    # parity on the device against torch's own reductions over the whole block (and over every rank's block)
    res = merged_t if world > 1 else outs_t
    ys = feeds_t["y_input" if wl == "c2" else "s_input"]
    zs = feeds_t["z_input" if wl == "c2" else "m_input"]
    ref_sum = torch.zeros(cell, dtype=torch.float64, device=device)
    ref_abs = torch.zeros(cell, dtype=torch.float64, device=device)
    for a in range(0, rows, 1 << 26):
        ref_sum += ys[a:a + (1 << 26)].sum(dim=0)
        ref_abs += ys[a:a + (1 << 26)].abs().sum(dim=0)
    ref_min = zs.min(dim=0).values
    if world > 1:
        dist.all_reduce(ref_sum)
        dist.all_reduce(ref_abs)
        dist.all_reduce(ref_min, op=dist.ReduceOp.MIN)
    assert torch.all((res[sum_name] - ref_sum).abs() <= 1e-12 * ref_abs), (res[sum_name], ref_sum)
    assert torch.equal(res[min_name], ref_min)
    parity = {"full_block_vs_torch": "sum within 1e-12 of sum|x|, min bit-exact" + (", merged over all ranks" if world > 1 else "")}

    # ---- end to end: host (pinned) column buffers through the same C-ABI call -----------------
    # one host buffer fed to two placeholders (reduce_1b: sum and min of the same column) is staged once by the library
    h2d = sum(v.numel() * v.element_size() for v in {id(v): v for v in feeds_t.values()}.values())
    d2h = sum(v.numel() * v.element_size() for v in outs_t.values())
    e2e, pinned, host_in = None, [], {}
    if not args.no_e2e:
        try:
            seen = {}
            for k, v in feeds_t.items():
                if id(v) not in seen:
                    pb = ctx.pinned(v.shape, np.float64, v)
                    pinned.append(pb)
                    seen[id(v)] = pb.array
                host_in[k] = seen[id(v)]
            host_out = {}
            for k, v in outs_t.items():
                pb = ctx.pinned(v.shape, np.float64)
                pinned.append(pb)
                host_out[k] = pb.array
            torch.cuda.synchronize()
            k_e2e = max(1, min(args.steps, 5 if wl == "c2" else 3))

            def e2e_step():
                sess.run_into(host_in, host_out)      # H2D + kernels + D2H + sync inside
                if world > 1:
                    for n in names:
                        outs_t[n].copy_(torch.from_numpy(host_out[n]))
                    sess.reduce_combine(ctx.comm, outs, merged, overlap=False)
                    for n in names:
                        torch.from_numpy(host_out[n]).copy_(merged_t[n])
                    torch.cuda.synchronize()

            dt = ctx.e2e_timed(e2e_step, k_e2e)
            e2e = {"value": world * rows * k_e2e / dt, "unit": "rows/s", "steps": k_e2e,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "h2d_gbs_per_gpu": h2d * k_e2e / dt / 1e9}
            if world == 1:
                assert np.array_equal(host_out[min_name], res[min_name].cpu().numpy())
        except Exception as exc:   # e.g. not enough pinnable host memory on this box
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200],
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}

    # ---- CPU baseline beside it (rank 0, N = 1 only) + oracle check of a prefix block ----------------------
    cpu = None
    if ctx.rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb, tf_ops
        cores = cb.cores()
        have_host = len(host_in) == len(feeds_t)
        hy = host_in[list(feeds_t)[0]] if have_host else None
        hz = host_in[list(feeds_t)[1]] if have_host else None
        avail = rows if have_host else min(rows, 20_000_000)
        if not have_host:
            hc = host_columns("c2" if wl == "c2" else "c3", 0, avail)
            hy, hz = (hc["y_input"], hc["z_input"]) if wl == "c2" else (hc["x"], hc["x"])
        y2 = hy if wl == "c2" else hy.reshape(-1, 1)
        z2 = hz if wl == "c2" else hz.reshape(-1, 1)
        sample, t, _ = cpu_sample(lambda r: cb.reduce_blocks_sum_min(y2[:r], z2[:r], cores, cores), avail, 2_000_000)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads, best of 3; "
                         "columnar unboxed input, so an upper bound on the JVM reference"}
        # oracle on a prefix block regenerated on the host: same generator, GPU result through the same session
        n_chk = min(rows, 1 << 20)
        hc = host_columns("c2" if wl == "c2" else "c3", 0, n_chk)
        o_feeds = hc if wl == "c2" else {"s_input": hc["x"], "m_input": hc["x"]}
        want = tf_ops.evaluate(build_graph(wl), o_feeds, names)
        pre_feeds = {k: DA(v[:n_chk]) for k, v in feeds_t.items()}
        sess.run_into(pre_feeds, outs)
        sess.sync()
        got = [outs_t[n].cpu().numpy() for n in names]
        scale = np.abs(o_feeds[list(o_feeds)[0]]).sum(axis=0)
        assert np.all(np.abs(got[0] - want[0]) <= 1e-12 * scale), (got[0], want[0])
        assert np.array_equal(got[1], want[1])
        parity["oracle_prefix_rows"] = n_chk
        parity["oracle"] = "sum within 1e-12 of sum|x| of oracle.tf_ops, min bit-exact, on host-regenerated values"
    host_in = {}
    for pb in pinned:
        pb.free()
    bytes_per_launch = rows * ALG_BYTES_PER_ROW[wl] / 2
    roof = hbm_roofline(wl, rows, bytes_per_launch, kernel_ms / 1e3 / (2 * args.steps), "reduce_axis0_tma_kernel")
    ex = ("grouped ncclAllReduce of the cell-sized partial rows behind the C ABI (tfs_session_reduce_combine), on the "
          "communicator's side stream: overlaps the next block's kernels") if world > 1 else "none"
    rec = record(ctx, wl, rows, elapsed_ms, clocks, launches, roof, e2e, cpu, parity, {"exchange": ex},
                 host_enqueue_ms_per_step=host_ms)
    sess.close()
    graph.close()
    return rec


# ---------------------------------------------------------------------------------------------
def run_c3(ctx, rows, cols=None):
    torch, DA, args = ctx.torch, ctx.DA, ctx.args
    world, device, stream = ctx.world, ctx.device, ctx.stream
    graph, sess = ctx.session("c3")
    x = (cols or device_columns(torch, "c3", ctx.rank * rows, rows, device))["x"]
    z = torch.empty(rows, dtype=torch.float32, device=device)
    torch.cuda.synchronize()
    feeds, outs = {"x": DA(x)}, {"z": DA(z)}

    def step(timed):
        sess.run_into(feeds, outs)

    launches0 = sess.launch_count
    elapsed_ms, clocks, host_ms = ctx.timed(step)
    launches = (sess.launch_count - launches0) * args.steps // (args.steps + args.warmup)
    # parity over the WHOLE block against torch on the device (bit-exact), chunked to bound the temporaries
    for a in range(0, rows, 1 << 26):
        xa = x[a:a + (1 << 26)]
        assert torch.equal(z[a:a + (1 << 26)], torch.clamp_min(xa * 1.5 + 0.25, 0.0).to(torch.float32))
    parity = {"full_block_vs_torch": "bit-exact over all rows"}
    h2d, d2h = rows * 8, rows * 4
    e2e, pinned, hx = None, [], None
    if not args.no_e2e:
        try:
            px = ctx.pinned((rows,), np.float64, x)
            pinned.append(px)
            pz = ctx.pinned((rows,), np.float32)
            pinned.append(pz)
            hx = px.array
            k_e2e = max(1, min(args.steps, 3))
            dt = ctx.e2e_timed(lambda: sess.run_into({"x": px.array}, {"z": pz.array}), k_e2e)
            e2e = {"value": world * rows * k_e2e / dt, "unit": "rows/s", "steps": k_e2e, "h2d_bytes_per_step": int(h2d),
                   "d2h_bytes_per_step": int(d2h), "h2d_gbs_per_gpu": h2d * k_e2e / dt / 1e9}
            assert np.array_equal(pz.array[:1 << 20], z[:1 << 20].cpu().numpy())
        except Exception as exc:
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200], "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}
    cpu = None
    if ctx.rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb, tf_ops
        cores = cb.cores()
        avail = rows if hx is not None else min(rows, 50_000_000)
        hxx = hx if hx is not None else host_columns("c3", 0, avail)["x"]
        sample, t, _ = cpu_sample(lambda r: cb.map_blocks_chain(hxx[:r], 1.5, 0.25, cores, cores), avail, 4_000_000)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads, best of 3; "
                         "columnar unboxed input, so an upper bound on the JVM reference"}
        n_chk = min(rows, 1 << 20)
        hc = host_columns("c3", 0, n_chk)
        (want,) = tf_ops.evaluate(build_graph("c3"), hc, ["z"])
        assert np.array_equal(z[:n_chk].cpu().numpy(), want)
        parity["oracle_prefix_rows"] = n_chk
        parity["oracle"] = "bit-exact against oracle.tf_ops on host-regenerated values"
    hx = None
    for pb in pinned:
        pb.free()
    roof = hbm_roofline("c3", rows, rows * 12.0, elapsed_ms / 1e3 / args.steps, "chain_kernel")
    rec = record(ctx, "c3", rows, elapsed_ms, clocks, launches, roof, e2e, cpu, parity, {"exchange": "none"},
                 host_enqueue_ms_per_step=host_ms)
    sess.close()
    graph.close()
    return rec


# ---------------------------------------------------------------------------------------------
def run_c4(ctx, rows):
    """BASELINE config 4.  One block per GPU; with N > 1 the step also moves every partial group to the rank owning its
    key and merges there (tfs_session_aggregate_combine), so the result is one row per distinct key, sharded by owner."""
    torch, dist, DA, args = ctx.torch, ctx.dist, ctx.DA, ctx.args
    world, device = ctx.world, ctx.device
    graph, sess = ctx.session("c4")
    tried = []
    while True:       # 5e8 rows need 132 GB + ~8 GB of scratch: halve on an allocation failure and say so
        try:
            t = device_columns(torch, "c4", ctx.rank * rows, rows, device)
            keys, v = t["keys"], t["v_input"]
            out_keys = torch.empty(KEYS_C4, dtype=torch.int64, device=device)
            out = torch.empty((KEYS_C4, 64), dtype=torch.float32, device=device)
            torch.cuda.synchronize()
            torch.cuda.empty_cache()      # the generator's temporaries must not sit in torch's cache next to the library's scratch
            n_groups = sess.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(out)}, DA(out_keys))
            break
        except (RuntimeError, ctx.native.NativeError) as exc:
            if "memory" not in str(exc).lower() or rows <= 50_000_000:
                raise
            tried.append({"rows": rows, "error": str(exc)[:160]})
            t = keys = v = out = out_keys = None
            torch.cuda.empty_cache()
            rows //= 2
    if world > 1:      # every rank must run the same block size
        r = torch.tensor([rows], dtype=torch.int64, device=device)
        dist.all_reduce(r, op=dist.ReduceOp.MIN)
        if int(r.item()) != rows:
            rows = int(r.item())
            keys, v = keys[:rows], v[:rows]
    out_keys2 = torch.empty(KEYS_C4, dtype=torch.int64, device=device) if world > 1 else None
    out2 = torch.empty((KEYS_C4, 64), dtype=torch.float32, device=device) if world > 1 else None
    owned = [0]

    def step(timed):
        g = sess.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(out)}, DA(out_keys))
        if world > 1:
            owned[0] = sess.aggregate_combine(ctx.comm, g, DA(out_keys), {"v": DA(out)}, {"v": DA(out2)}, DA(out_keys2))
        return g

    launches0 = sess.launch_count
    elapsed_ms, clocks, host_ms = ctx.timed(step)
    launches = (sess.launch_count - launches0) * args.steps // (args.steps + args.warmup)
    # ---- parity at the full block size: keys exactly once each; sampled groups and the checksum of checksums in f64
    ok = out_keys[:n_groups]
    assert n_groups <= KEYS_C4 and bool(torch.all(ok[1:] > ok[:-1]))               # each key once, ascending
    universe = torch.sort(mix_torch(torch.arange(KEYS_C4, dtype=torch.int64, device=device) ^ 0x4B455953)).values
    assert torch.equal(universe[torch.searchsorted(universe, ok).clamp_max(KEYS_C4 - 1)], ok)     # no key that is not in the data's key table
    sample_keys = ok[torch.linspace(0, n_groups - 1, 64, device=device).long()]
    ref = torch.zeros((64, 64), dtype=torch.float64, device=device)
    col = torch.zeros(64, dtype=torch.float64, device=device)
    for a in range(0, rows, 1 << 24):
        kc, vc = keys[a:a + (1 << 24)], v[a:a + (1 << 24)]
        col += vc.sum(0, dtype=torch.float64)
        assert torch.equal(ok[torch.searchsorted(ok, kc).clamp_max(n_groups - 1)], kc)             # every row's key has its group, bit-exact
        pos = torch.searchsorted(sample_keys, kc).clamp_max(63)
        hit = sample_keys[pos] == kc
        ref.index_add_(0, pos[hit], vc[hit].double())
    got = out[:n_groups][torch.searchsorted(ok, sample_keys)].double()
    assert torch.all((got - ref).abs() <= 1e-5 * ref.abs().clamp_min(1e-30)), "sampled group sums"
    tot = out[:n_groups].sum(0, dtype=torch.float64)
    assert torch.all((tot - col).abs() <= 1e-5 * col), "group sums must add up to the column sums"
    parity = {"full_block_vs_torch": f"{n_groups} keys bit-exact and ascending; 64 sampled groups and the column totals within 1e-5 (f64 reference)"}
    if world > 1:
        tot_owned = torch.tensor([owned[0]], dtype=torch.int64, device=device)
        dist.all_reduce(tot_owned)
        all_keys = torch.tensor([n_groups], dtype=torch.int64, device=device)
        dist.all_reduce(all_keys, op=dist.ReduceOp.MAX)
        assert int(tot_owned.item()) >= int(all_keys.item()) and int(tot_owned.item()) <= KEYS_C4
        tot2 = out2[:owned[0]].sum(0, dtype=torch.float64)
        dist.all_reduce(tot2)
        col_all = col.clone()
        dist.all_reduce(col_all)
        assert torch.all((tot2 - col_all).abs() <= 1e-5 * col_all), "merged group sums must add up to all ranks' column sums"
        parity["merged"] = f"{int(tot_owned.item())} owned groups over all ranks; merged totals within 1e-5 of all ranks' column sums"
    # ---- end to end: host (pinned) key and payload columns through the same C-ABI call; bounded block
    rows_e2e = min(rows, 50_000_000)
    e2e, pinned, host = None, [], {}
    if not args.no_e2e:
        try:
            from tensorframes_b200.shape import LongType
            pk = ctx.pinned((rows_e2e,), np.int64, keys[:rows_e2e]); pinned.append(pk)
            pv = ctx.pinned((rows_e2e, 64), np.float32, v[:rows_e2e]); pinned.append(pv)
            po = ctx.pinned((KEYS_C4, 64), np.float32); pinned.append(po)
            pok = ctx.pinned((KEYS_C4,), np.int64); pinned.append(pok)
            host = {"keys": pk.array, "v_input": pv.array}
            NDA = ctx.native.DeviceArray
            hk = NDA(pk.array.ctypes.data, LongType, pk.array.shape, owner=pk.array)
            hok = NDA(pok.array.ctypes.data, LongType, pok.array.shape, owner=pok.array)
            g_last = [0]

            def e2e_step():
                if world == 1:
                    g_last[0] = sess.aggregate_into(hk, {"v_input": pv.array}, {"v": po.array}, hok)   # H2D, passes, D2H, sync
                    return
                g = sess.aggregate_into(hk, {"v_input": pv.array}, {"v": DA(out)}, DA(out_keys))
                g2 = sess.aggregate_combine(ctx.comm, g, DA(out_keys), {"v": DA(out)}, {"v": DA(out2)}, DA(out_keys2))
                torch.from_numpy(po.array)[:g2].copy_(out2[:g2])
                torch.from_numpy(pok.array)[:g2].copy_(out_keys2[:g2])
                torch.cuda.synchronize()
                g_last[0] = g2

            k_e2e = max(1, min(args.steps, 3))
            dt = ctx.e2e_timed(e2e_step, k_e2e)
            e2e = {"value": world * rows_e2e * k_e2e / dt, "unit": "rows/s", "steps": k_e2e, "rows_per_gpu": rows_e2e,
                   "h2d_bytes_per_step": int(rows_e2e * 264), "d2h_bytes_per_step": int(g_last[0] * 264),
                   "h2d_gbs_per_gpu": rows_e2e * 264 * k_e2e / dt / 1e9}
        except Exception as exc:
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200], "h2d_bytes_per_step": int(rows_e2e * 264), "d2h_bytes_per_step": 0}
    cpu = None
    if ctx.rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb
        cores = cb.cores()
        avail = rows_e2e if host else min(rows, 4_000_000)
        hc = host if host else host_columns("c4", 0, avail)
        hk_np, hv_np = hc["keys"], hc["v_input"]
        sample, t, _ = cpu_sample(lambda r: cb.aggregate_sum(hk_np[:r], hv_np[:r], cores, cores, KEYS_C4), avail, 2_000_000)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads (per-partition hash "
                         "aggregation + merge), best of 3; columnar unboxed input, so an upper bound on the JVM reference"}
        # oracle on a prefix block regenerated on the host: keys bit-exact, sums within 1e-5 of an f64 group-by
        n_chk = min(rows, 1 << 19)
        hcc = host_columns("c4", 0, n_chk)
        assert np.array_equal(hcc["keys"], keys[:n_chk].cpu().numpy()) and np.array_equal(hcc["v_input"][:4096], v[:4096].cpu().numpy())
        uk, inv = np.unique(hcc["keys"], return_inverse=True)
        want = np.zeros((len(uk), 64))
        np.add.at(want, inv, hcc["v_input"].astype(np.float64))
        ok_t = torch.empty(len(uk), dtype=torch.int64, device=device)
        o_t = torch.empty((len(uk), 64), dtype=torch.float32, device=device)
        g = sess.aggregate_into(DA(keys[:n_chk]), {"v_input": DA(v[:n_chk])}, {"v": DA(o_t)}, DA(ok_t))
        assert g == len(uk) and np.array_equal(ok_t.cpu().numpy(), uk)
        assert np.all(np.abs(o_t.cpu().numpy().astype(np.float64) - want) <= 1e-5 * np.maximum(want, 1e-30))
        parity["oracle_prefix_rows"] = n_chk
        parity["oracle"] = "keys bit-exact, sums within 1e-5 of an f64 group-by on host-regenerated values"
    host = {}
    for pb in pinned:
        pb.free()
    peak, peak_src = hbm_peak()
    step_s = elapsed_ms / 1e3 / args.steps
    achieved = rows * 264 / step_s / 1e9
    traffic, tsrc = ncu_traffic("c4", rows)
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
            "traffic_source": tsrc, "peak_source": peak_src,
            "kernel": "whole step (insert, table scan, row partition, gather-reduce; agg_gather_reduce_kernel dominant)",
            "algorithmic_bytes_per_launch": rows * 264.0, "avg_launch_ms": step_s * 1e3, "frac_of_nominal_8000": achieved / 8000.0}
    ex = "none" if world == 1 else "key-owner exchange of partial groups + merge behind the C ABI (tfs_session_aggregate_combine)"
    cfg = {"exchange": ex, "distinct_keys": int(n_groups)}
    if tried:
        cfg["sizes_that_failed"] = tried
    rec = record(ctx, "c4", rows, elapsed_ms, clocks, launches, roof, e2e, cpu, parity, cfg, host_enqueue_ms_per_step=host_ms)
    sess.close()
    graph.close()
    return rec


# ---------------------------------------------------------------------------------------------
def run_c5(ctx, rows):
    torch, DA, args = ctx.torch, ctx.DA, ctx.args
    world, device = ctx.world, ctx.device
    graph, sess = ctx.session("c5")
    x = device_columns(torch, "c5", ctx.rank * rows, rows, device)["x"]
    y = torch.empty((rows, 256), dtype=torch.float32, device=device)
    torch.cuda.synchronize()

    def step(timed):
        sess.run_into({"x": DA(x)}, {"y": DA(y)})

    launches0 = sess.launch_count
    elapsed_ms, clocks, host_ms = ctx.timed(step)
    launches = (sess.launch_count - launches0) * args.steps // (args.steps + args.warmup)
    # parity against float64 on slices at both ends and the middle (normwise 1e-5)
    w_np, b_np = c5_constants()
    w, b = torch.from_numpy(w_np).to(device), torch.from_numpy(b_np).to(device)
    for lo in (0, rows // 2, max(0, rows - 65536)):
        xs = x[lo:lo + 65536].double()
        exact = xs @ w.double() + b.double()
        scale = xs.abs() @ w.double().abs() + b.double().abs()
        assert torch.all((y[lo:lo + 65536].double() - exact).abs() <= 1e-5 * scale)
    parity = {"full_block_vs_torch": "three 65536-row slices within 1e-5 (normwise) of an f64 product"}
    rows_e2e = min(rows, 1 << 22)
    e2e, pinned, hx = None, [], None
    if not args.no_e2e:
        try:
            px = ctx.pinned((rows_e2e, 256), np.float32, x[:rows_e2e]); pinned.append(px)
            py = ctx.pinned((rows_e2e, 256), np.float32); pinned.append(py)
            hx = px.array
            k_e2e = max(1, min(args.steps, 3))
            dt = ctx.e2e_timed(lambda: sess.run_into({"x": px.array}, {"y": py.array}), k_e2e)
            assert np.array_equal(py.array[:4096], y[:4096].cpu().numpy())
            e2e = {"value": world * rows_e2e * k_e2e / dt, "unit": "rows/s", "steps": k_e2e, "rows_per_gpu": rows_e2e,
                   "h2d_bytes_per_step": int(rows_e2e * 1024), "d2h_bytes_per_step": int(rows_e2e * 1024),
                   "h2d_gbs_per_gpu": rows_e2e * 1024 * k_e2e / dt / 1e9}
        except Exception as exc:
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200],
                   "h2d_bytes_per_step": int(rows_e2e * 1024), "d2h_bytes_per_step": int(rows_e2e * 1024)}
    cpu = None
    if ctx.rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb, tf_ops
        cores = cb.cores()
        avail = rows_e2e if hx is not None else min(rows, 262_144)
        xs_np = hx if hx is not None else host_columns("c5", 0, avail)["x"]
        sample, t, _ = cpu_sample(lambda r: cb.matmul_bias(xs_np[:r], w_np, b_np, cores, cores), avail, 65536)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads, best of 3; "
                         "plain C loops (the reference's TF-CPU build calls Eigen's GEMM, which would be faster than this port)"}
        n_chk = 4096
        hc = host_columns("c5", 0, n_chk)
        (want,) = tf_ops.evaluate(build_graph("c5"), hc, ["y"])
        scale = np.abs(hc["x"].astype(np.float64)) @ np.abs(w_np.astype(np.float64)) + np.abs(b_np)
        assert np.all(np.abs(y[:n_chk].cpu().numpy().astype(np.float64) - want.astype(np.float64)) <= 2e-5 * scale)
        parity["oracle_prefix_rows"] = n_chk
        parity["oracle"] = "within 1e-5 (normwise, each side) of oracle.tf_ops' MatMul on host-regenerated values"
    hx = None
    for pb in pinned:
        pb.free()
    pk = peaks()
    bf16 = float(pk.get("bf16_tflops_sustained", 1400.0))
    step_s = elapsed_ms / 1e3 / args.steps
    tflops = rows * FLOP_PER_ROW_C5 / step_s / 1e12
    hbm, _ = hbm_peak()
    traffic, tsrc = ncu_traffic("c5", rows)
    roof = {"bound": "tensor", "achieved": tflops, "peak": bf16, "unit": "TFLOP/s", "frac": tflops / bf16,
            "traffic": traffic, "traffic_source": tsrc,
            "peak_source": "measured bf16 sustained; tf32 runs at half of it and each fp32 product costs 3 tf32 MMAs, "
                           "so the attainable ceiling is ~peak/6",
            "kernel": "matmul_tc_pair_kernel", "mma_tflops_issued": 3 * tflops, "frac_of_attainable_peak_over_6": tflops / (bf16 / 6),
            "avg_launch_ms": step_s * 1e3, "hbm_gbs": rows * 2048.0 / step_s / 1e9, "hbm_frac": rows * 2048.0 / step_s / 1e9 / hbm,
            "ncu_tensor_pipe_active_pct_of_elapsed": ncu_extra("c5", "tensor_pipe_active_pct_of_elapsed")}
    rec = record(ctx, "c5", rows, elapsed_ms, clocks, launches, roof, e2e, cpu, parity, {"exchange": "none"},
                 host_enqueue_ms_per_step=host_ms)
    sess.close()
    graph.close()
    return rec


def run_ours(args):
    ctx = Ctx(args)
    torch = ctx.torch
    todo = ALL if args.workload == "all" else [args.workload]
    recs = {}
    shared_x = None
    for wl in todo:
        rows = args.rows or ROWS[wl]
        try:
            if wl == "c2":
                recs[wl] = run_reduce(ctx, "c2", rows)
            elif wl == "c3":
                # the 1e9-row column is generated once and shared with reduce_1b
                if "reduce_1b" in todo and (args.rows or ROWS["reduce_1b"]) == rows:
                    shared_x = device_columns(torch, "c3", ctx.rank * rows, rows, ctx.device)
                recs[wl] = run_c3(ctx, rows, shared_x)
            elif wl == "reduce_1b":
                recs[wl] = run_reduce(ctx, "reduce_1b", rows, shared_x)
                shared_x = None
            elif wl == "c4":
                recs[wl] = run_c4(ctx, rows)
            else:
                recs[wl] = run_c5(ctx, rows)
        except Exception as exc:      # one workload failing must not take the headline line with it
            if wl == todo[0]:
                raise
            import traceback
            recs[wl] = {"metric": METRIC[wl], "value": None, "error": (str(exc) or repr(exc))[:300],
                        "where": traceback.format_exc().strip().splitlines()[-3:]}
        if wl != "c3":
            shared_x = None
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
    if ctx.rank == 0:
        line = dict(recs[todo[0]])
        if len(todo) > 1:
            line["configs"] = {wl: recs[wl] for wl in todo[1:]}
        print(json.dumps(line))
    if ctx.comm is not None:
        ctx.comm.close()
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
REF_ROWS = {"c2": 100_000_000, "c3": 100_000_000, "reduce_1b": 100_000_000, "c4": 4_000_000, "c5": 262_144}


def reference_record(args, wl, steps, warmup):
    """oracle/cpu_ref.c on all host threads, one bounded block per step (same generator as the GPU arm)."""
    from oracle import cpu_baseline as cb
    cores = cb.cores()
    rows_full = args.rows or ROWS[wl]
    sample = min(rows_full, REF_ROWS[wl])
    cols = host_columns("c3" if wl == "reduce_1b" else wl, 0, sample)
    if wl == "c2":
        fn = lambda: cb.reduce_blocks_sum_min(cols["y_input"], cols["z_input"], cores, cores)
    elif wl == "reduce_1b":
        x2 = cols["x"].reshape(-1, 1)
        fn = lambda: cb.reduce_blocks_sum_min(x2, x2, cores, cores)
    elif wl == "c3":
        fn = lambda: cb.map_blocks_chain(cols["x"], 1.5, 0.25, cores, cores)
    elif wl == "c4":
        fn = lambda: cb.aggregate_sum(cols["keys"], cols["v_input"], cores, cores, KEYS_C4)
    else:
        w, b = c5_constants()
        fn = lambda: cb.matmul_bias(cols["x"], w, b, cores, cores)
    for _ in range(warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = time.perf_counter() - t0
    value = sample * steps / dt
    same = sample == rows_full
    desc = (f"{sample} rows per step ({'the whole block' if same else 'bounded sample of the ' + str(rows_full) + '-row block'}), "
            f"{cores} partitions on {cores} threads; oracle/cpu_ref.c restates the reference's per-block algorithm on "
            "columnar input (Spark + TF-Java are not installable here)")
    return {
        "impl": "reference", "metric": METRIC[wl], "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": DTYPE[wl], "data": "synthetic (counter-based generator, SURVEY 8(d))",
        "config": {"workload": WORKLOAD[wl] + f"{rows_full} rows per GPU", "rows_per_gpu": rows_full, "partitions": args.gpus,
                   "l2": "inputs larger than L2 (no flush needed)", "exchange": "none"},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def run_reference(args):
    """The reference arm: the restatement of the reference's own CPU algorithm (oracle/cpu_ref.c), all host threads,
    same workloads / metric / unit.  Rank 0 only."""
    if env_int("RANK", 0) != 0:
        return
    todo = ALL if args.workload == "all" else [args.workload]
    line = reference_record(args, todo[0], args.steps, args.warmup)
    if len(todo) > 1:
        line["configs"] = {wl: reference_record(args, wl, max(1, min(args.steps, 5)), min(args.warmup, 1)) for wl in todo[1:]}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all"] + ALL)
    ap.add_argument("--rows", type=int, default=0, help="rows per GPU (default: the BASELINE config's)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    import __graft_entry__ as entry
    if not os.path.exists(entry.LIB_PATH) or not os.path.exists(entry.ORACLE_LIB):
        if env_int("LOCAL_RANK", 0) == 0:
            entry.build()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
