#!/usr/bin/env python
"""bench.py -- throughput of the TensorFrames block path on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (BASELINE.json configs[1], name in config.workload): reduce_blocks with
y = reduce_sum(y_input, [0]) and z = reduce_min(z_input, [0]) over two analysed
DoubleType[?,2] columns of 100M synthetic rows per GPU.  A "step" is one pass of that block
through the reducer graph (N > 1: one partition per GPU, weak scaling, plus the exchange of the
cell-sized partial rows and the merge run).

  value      rows/s, whole job, inputs resident in HBM, timed with CUDA events on the launch stream
  e2e        the same metric through the C ABI with HOST (pinned) column buffers: H2D of the block
             and D2H of the result row are inside the timed region
  roofline   the dominant kernel (reduce_axis0_tma_kernel) against the measured HBM peak
  cpu_baseline / --impl reference   oracle/cpu_ref.c, the restatement of the reference's
             per-block algorithm, on the host cores (the real Spark + TF-CPU stack is not installable here)

--workload c3 | c4 | c5 measure the other BASELINE configs the same way: map_blocks on the fused chain (chain_kernel),
the keyed aggregate (insert / sort / gather passes) and the tcgen05 MatMul (matmul_tc_pair_kernel; roofline.bound =
"tensor").  With N > 1 the reduce exchange of a block runs on a second stream and overlaps the next block's kernels.
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

ROWS_C2 = 100_000_000
ROWS_C3 = 1_000_000_000
ROWS_C4 = 250_000_000          # BASELINE config 4 names 5e8 rows (132 GB); see config.workload
KEYS_C4 = 1_000_000
ROWS_C5 = 1 << 24               # BASELINE config 5: FloatType[?,256] x 256x256 + bias, 2^24 rows per GPU
ALG_BYTES_PER_ROW = {"c2": 32, "c3": 12, "c4": 264, "c5": 2048}
FALLBACK_HBM_GBS = 6650.0                   # /opt/skills/guides/B200_PROFILING.md


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback"


def ncu_traffic(workload, rows=None):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/traffic.json),
    scaled by rows when this run's block differs from the captured one; None when there is no capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f).get(workload)
        if not e:
            return None
        if rows is None or rows == e["rows"]:
            return e["bytes"]
        return e["bytes"] * rows / e["rows"]
    except Exception:
        return None


def ncu_extra(workload, key):
    """Another figure of the committed ncu capture of the workload's dominant kernel (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload, {}).get(key)
    except Exception:
        return None


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
def build_graph(workload):
    import tensorframes_b200 as tfs
    from tensorframes_b200 import dsl
    with dsl.withGraph():
        if workload == "c2":
            y = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
            z = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("z_input")
            return dsl.build_graph(dsl.reduce_sum(y, [0]).named("y"), dsl.reduce_min(z, [0]).named("z"))
        if workload == "c5":
            rng = np.random.default_rng(5)
            w = rng.uniform(-1 / 16, 1 / 16, (256, 256)).astype(np.float32)
            b = rng.uniform(-1, 1, 256).astype(np.float32)
            x = dsl.placeholder(tfs.FloatType, dsl.Unknown, 256).named("x")
            return dsl.build_graph((dsl.matmul(x, dsl.constant(w)) + dsl.constant(b)).named("y"))
        if workload == "c4":
            v = dsl.placeholder(tfs.FloatType, dsl.Unknown, 64).named("v_input")
            return dsl.build_graph(dsl.reduce_sum(v, [0]).named("v"))
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown).named("x")
        return dsl.build_graph(dsl.cast(dsl.maximum(x * 1.5 + 0.25, 0.0), tfs.FloatType).named("z"))


def make_inputs(torch, workload, rows, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    if workload == "c2":
        u = torch.rand(rows, generator=g, device=device, dtype=torch.float64)
        y = torch.stack([u, -u], dim=1).contiguous()
        u = torch.rand(rows, generator=g, device=device, dtype=torch.float64)
        z = torch.stack([u, -u], dim=1).contiguous()
        return {"y_input": y, "z_input": z}
    if workload == "c4":
        # keys: a table of 1e6 distinct random 64-bit values (not dense ids, so the hash is really exercised)
        # (the key universe is shared by all ranks, as the partitions of one DataFrame would share it)
        gk = torch.Generator(device=device)
        gk.manual_seed(4)
        table = torch.randint(-2**62, 2**62, (KEYS_C4,), generator=gk, device=device, dtype=torch.int64)
        table = torch.unique(table)
        idx = torch.randint(0, table.numel(), (rows,), generator=g, device=device, dtype=torch.int64)
        keys = table[idx]
        del idx
        v = torch.rand((rows, 64), generator=g, device=device, dtype=torch.float32)
        return {"keys": keys, "v_input": v}
    x = torch.rand(rows, generator=g, device=device, dtype=torch.float64) * 2.0 - 1.0
    return {"x": x}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from tensorframes_b200 import _native
    from tensorframes_b200.distributed import combine_reduce_blocks

    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run --nproc-per-node N")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    wl = args.workload
    rows = args.rows or {"c2": ROWS_C2, "c3": ROWS_C3, "c4": ROWS_C4, "c5": ROWS_C5}[wl]
    DA = _native.DeviceArray.from_torch
    if wl == "c4":
        return run_c4(args, torch, dist, _native, world, rank, local, device, rows)
    if wl == "c5":
        return run_c5(args, torch, dist, _native, world, rank, local, device, rows)

    graph_bytes = build_graph(wl)
    graph = _native.Graph(graph_bytes)
    sess = _native.Session(graph, local)
    # a real (non-default) stream: the C ABI treats a NULL stream handle as "use the session's own"
    stream = torch.cuda.Stream(device=device)
    torch.cuda.set_stream(stream)
    sess.set_stream(stream.cuda_stream)
    feeds_t = make_inputs(torch, wl, rows, seed=(2 if wl == "c2" else 3) + 1000 * rank, device=device)
    feeds = {k: DA(v) for k, v in feeds_t.items()}
    if wl == "c2":
        outs_t = {"y": torch.empty(2, dtype=torch.float64, device=device),
                  "z": torch.empty(2, dtype=torch.float64, device=device)}
        merged_t = {k: torch.empty_like(v) for k, v in outs_t.items()}
    else:
        outs_t = {"z": torch.empty(rows, dtype=torch.float32, device=device)}
        merged_t = {}
    outs = {k: DA(v) for k, v in outs_t.items()}
    names = list(outs_t)
    # N > 1: the exchange of a block (all_gather of the partial rows + merge run) goes on a second stream with its own
    # session, so it overlaps the next block's reduction kernels; the partial rows are kept in a small ring of buffers for that
    overlap = wl == "c2" and world > 1
    if overlap:
        side = torch.cuda.Stream(device=device)
        merge_sess = _native.Session(graph, local)
        merge_sess.set_stream(side.cuda_stream)
        depth = max(2, env_int("TFS_BENCH_EXCHANGE_DEPTH", 4))      # partial-row buffers in flight
        part_t = [outs_t] + [{k: torch.empty_like(v) for k, v in outs_t.items()} for _ in range(depth - 1)]
        part = [{k: DA(v) for k, v in p.items()} for p in part_t]
        reduced_ev = [torch.cuda.Event() for _ in range(depth)]
        exchanged_ev = [torch.cuda.Event() for _ in range(depth)]
        step_no = [0]

    def merge(feeds_stacked):
        (merge_sess if overlap else sess).run_into({k: DA(v) for k, v in feeds_stacked.items()}, {k: DA(v) for k, v in merged_t.items()})
        return [merged_t[n] for n in names]

    kernel_events = []

    def step(timed):
        if timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        if overlap:
            b = step_no[0] % depth
            step_no[0] += 1
            stream.wait_event(exchanged_ev[b])          # the exchange that last read this buffer set (`depth` blocks ago)
            sess.run_into(feeds, part[b])
        else:
            sess.run_into(feeds, outs)
        if timed:
            e1.record(stream)
            kernel_events.append((e0, e1))
        if overlap:
            reduced_ev[b].record(stream)
            with torch.cuda.stream(side):
                side.wait_event(reduced_ev[b])
                combine_reduce_blocks([part_t[b][n] for n in names], names, merge)
                exchanged_ev[b].record(side)

    def barrier():
        if overlap:
            stream.wait_stream(side)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(False)
    barrier()
    launches0 = sess.launch_count
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        start.record(stream)
        t_host0 = time.perf_counter()
        for _ in range(args.steps):
            step(True)
        host_enqueue_ms = (time.perf_counter() - t_host0) * 1e3 / args.steps   # CPU time to enqueue one step
        if overlap:
            stream.wait_stream(side)       # the last block's exchange belongs to the timed region
        end.record(stream)
        barrier()
    elapsed_ms = start.elapsed_time(end)
    launches = sess.launch_count - launches0
    big_launches = args.steps * (2 if wl == "c2" else 1)
    kernel_ms = sum(a.elapsed_time(b) for a, b in kernel_events)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())

    # parity spot check on the device (the tests hold the real parity checks)
    if wl == "c2":
        res = merged_t if world > 1 else outs_t
        y = feeds_t["y_input"]
        ref_sum, ref_abs, ref_min = y.sum(dim=0), y.abs().sum(dim=0), feeds_t["z_input"].min(dim=0).values
        if world > 1:       # the merged row must be the reduction over every rank's partition
            dist.all_reduce(ref_sum)
            dist.all_reduce(ref_abs)
            dist.all_reduce(ref_min, op=dist.ReduceOp.MIN)
        assert torch.all((res["y"] - ref_sum).abs() <= 1e-12 * ref_abs), (res["y"], ref_sum)
        assert torch.equal(res["z"], ref_min)
    else:
        n_chk = min(rows, 1 << 24)
        x = feeds_t["x"][:n_chk]
        assert torch.equal(outs_t["z"][:n_chk], torch.clamp_min(x * 1.5 + 0.25, 0.0).to(torch.float32))

    # ---- end to end: host (pinned) column buffers through the same C-ABI call -----------------
    e2e = None
    host_in = {}
    pinned = []
    h2d = sum(v.numel() * v.element_size() for v in feeds_t.values())
    d2h = sum(v.numel() * v.element_size() for v in outs_t.values())
    if not args.no_e2e:
        try:
            for k, v in feeds_t.items():
                pb = _native.PinnedBuffer(tuple(v.shape), np.float64)
                pinned.append(pb)
                torch.from_numpy(pb.array).copy_(v)       # fill the pinned block once, outside the timing
                host_in[k] = pb.array
            host_out = {}
            for k, v in outs_t.items():
                pb = _native.PinnedBuffer(tuple(v.shape), np.float64 if v.dtype == torch.float64 else np.float32)
                pinned.append(pb)
                host_out[k] = pb.array
            torch.cuda.synchronize()
            k_e2e = max(1, min(args.steps, 5))

            def e2e_step():
                sess.run_into(host_in, host_out)      # H2D + kernels + D2H + sync inside
                if wl == "c2" and world > 1:
                    parts = [torch.from_numpy(host_out[n]).to(device) for n in names]
                    if overlap:      # the merge session lives on the side stream: order it after the copies above
                        side.wait_stream(torch.cuda.current_stream())
                        with torch.cuda.stream(side):
                            combine_reduce_blocks(parts, names, merge)
                    else:
                        combine_reduce_blocks(parts, names, merge)
                    torch.cuda.synchronize()

            e2e_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                e2e_step()
            barrier()
            dt = time.perf_counter() - t0
            tt = torch.tensor([dt], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e2e = {"value": world * rows * k_e2e / float(tt.item()), "unit": "rows/s", "steps": k_e2e,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}
            if wl == "c2" and world == 1:
                assert np.array_equal(host_out["z"], outs_t["z"].cpu().numpy())
        except Exception as exc:   # e.g. not enough pinnable host memory on this box
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200],
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}

    # ---- CPU baseline beside it (rank 0, N = 1 only): the oracle's restatement of the reference ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        have_host = len(host_in) == len(feeds_t)
        cpu = cpu_baseline_leg(wl, host_in if have_host else {k: v.cpu().numpy() for k, v in feeds_t.items()}, rows)
    host_in = {}
    for pb in pinned:
        pb.free()

    if rank == 0:
        peak, peak_src = hbm_peak()
        alg = ALG_BYTES_PER_ROW[wl]
        bytes_per_launch = rows * alg / (2 if wl == "c2" else 1)
        avg_launch_s = kernel_ms / 1e3 / big_launches
        achieved = bytes_per_launch / avg_launch_s / 1e9
        line = {
            "metric": "rows/s for reduceBlocks (reduce_sum + reduce_min over DoubleType[?,2])" if wl == "c2"
                      else "rows/s for mapBlocks (fused x*a+b, max(.,0), cast f32 on a float64 column)",
            "value": world * rows * args.steps / (elapsed_ms / 1e3),
            "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": elapsed_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": ("BASELINE configs[1]: reduce_blocks reduce_sum+reduce_min, DoubleType[?,2], "
                                    if wl == "c2" else
                                    "BASELINE configs[2]: map_blocks fused chain cast_f32(max(x*1.5+0.25,0)), float64 column, ")
                                   + f"{rows} rows per GPU", "rows_per_gpu": rows, "partitions": world,
                       "l2": "inputs larger than L2 (no flush needed)",
                       "exchange": ("all_gather of partial rows + merge run, on a second stream: overlaps the next block's kernels"
                                    if (wl == "c2" and world > 1) else "none")},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic(wl, rows), "peak_source": peak_src,
                         "kernel": "reduce_axis0_tma_kernel" if wl == "c2" else "chain_kernel",
                         "algorithmic_bytes_per_launch": bytes_per_launch,
                         "avg_launch_ms": avg_launch_s * 1e3, "frac_of_nominal_8000": achieved / 8000.0},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
            "cpu_baseline": cpu,
            "host_enqueue_ms_per_step": host_enqueue_ms,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_c4(args, torch, dist, _native, world, rank, local, device, rows):
    """BASELINE config 4: groupBy(int64 key).aggregate(reduce_sum(v_input, [0])), FloatType[?,64].
    One block per GPU; with N > 1 the step also moves every partial group to the rank owning its key
    (all_to_all) and merges there, so the result is one row per distinct key, sharded by owner."""
    DA = _native.DeviceArray.from_torch
    graph = _native.Graph(build_graph("c4"))
    sess = _native.Session(graph, local)
    stream = torch.cuda.Stream(device=device)
    torch.cuda.set_stream(stream)
    sess.set_stream(stream.cuda_stream)
    t = make_inputs(torch, "c4", rows, seed=4 + 1000 * rank, device=device)
    keys, v = t["keys"], t["v_input"]
    out_keys = torch.empty(KEYS_C4, dtype=torch.int64, device=device)
    out = torch.empty((KEYS_C4, 64), dtype=torch.float32, device=device)

    out_keys2 = torch.empty(KEYS_C4, dtype=torch.int64, device=device)
    out2 = torch.empty((KEYS_C4, 64), dtype=torch.float32, device=device)
    owned = [0]

    def step():
        g = sess.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(out)}, DA(out_keys))
        if world > 1:
            # exchange: partial groups travel to the rank that owns their key, which merges them
            from tensorframes_b200.distributed import exchange_groups
            rk, (rv,) = exchange_groups(out_keys[:g], [out[:g]])
            owned[0] = sess.aggregate_into(DA(rk), {"v_input": DA(rv)}, {"v": DA(out2)}, DA(out_keys2))
        return g

    for _ in range(args.warmup):
        n_groups = step()
    torch.cuda.synchronize()
    launches0 = sess.launch_count
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        start.record(stream)
        for _ in range(args.steps):
            n_groups = step()
        end.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    elapsed_ms = start.elapsed_time(end)
    tt = torch.tensor([elapsed_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    elapsed_ms = float(tt.item())
    launches = sess.launch_count - launches0
    # parity spot check on the device: keys exactly once each, sums within the f32 tolerance
    ref_keys, inv = torch.unique(keys, return_inverse=True)
    assert n_groups == ref_keys.numel()
    order = torch.argsort(out_keys[:n_groups])
    assert torch.equal(out_keys[:n_groups][order], ref_keys)
    ref = torch.zeros((n_groups, 64), dtype=torch.float64, device=device)
    chk = min(rows, 20_000_000)
    ref.index_add_(0, inv[:chk], v[:chk].double())
    if chk == rows:
        assert torch.all((out[:n_groups][order].double() - ref).abs() <= 1e-5 * ref.abs().clamp_min(1e-30))
    # ---- end to end: host (pinned) key and payload columns through the same C-ABI call ----------------
    # bounded block (the full 66 GB block would have to be pinned and staged next to the resident one)
    e2e = None
    rows_e2e = min(rows, 50_000_000)
    pinned = []
    host = {}
    if not args.no_e2e:
        try:
            from tensorframes_b200.shape import LongType
            pk = _native.PinnedBuffer((rows_e2e,), np.int64); pinned.append(pk)
            pv = _native.PinnedBuffer((rows_e2e, 64), np.float32); pinned.append(pv)
            po = _native.PinnedBuffer((KEYS_C4, 64), np.float32); pinned.append(po)
            pok = _native.PinnedBuffer((KEYS_C4,), np.int64); pinned.append(pok)
            torch.from_numpy(pk.array).copy_(keys[:rows_e2e])
            torch.from_numpy(pv.array).copy_(v[:rows_e2e])
            host = {"keys": pk.array, "v_input": pv.array}
            hk = _native.DeviceArray(pk.array.ctypes.data, LongType, pk.array.shape, owner=pk.array)
            hok = _native.DeviceArray(pok.array.ctypes.data, LongType, pok.array.shape, owner=pok.array)

            def e2e_step():
                if world == 1:
                    return sess.aggregate_into(hk, {"v_input": pv.array}, {"v": po.array}, hok)   # H2D, passes, D2H, sync
                from tensorframes_b200.distributed import exchange_groups
                g = sess.aggregate_into(hk, {"v_input": pv.array}, {"v": DA(out)}, DA(out_keys))
                rk, (rv,) = exchange_groups(out_keys[:g], [out[:g]])
                g2 = sess.aggregate_into(DA(rk), {"v_input": DA(rv)}, {"v": DA(out2)}, DA(out_keys2))
                torch.from_numpy(po.array)[:g2].copy_(out2[:g2])
                torch.from_numpy(pok.array)[:g2].copy_(out_keys2[:g2])
                torch.cuda.synchronize()
                return g2

            k_e2e = max(1, min(args.steps, 3))
            g_e2e = e2e_step()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                g_e2e = e2e_step()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            dt = time.perf_counter() - t0
            td = torch.tensor([dt], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(td, op=dist.ReduceOp.MAX)
            e2e = {"value": world * rows_e2e * k_e2e / float(td.item()), "unit": "rows/s", "steps": k_e2e,
                   "rows_per_gpu": rows_e2e, "h2d_bytes_per_step": int(rows_e2e * 264),
                   "d2h_bytes_per_step": int(g_e2e * 264)}
        except Exception as exc:
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200],
                   "h2d_bytes_per_step": int(rows_e2e * 264), "d2h_bytes_per_step": 0}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb
        cores = cb.cores()
        hk_np = host["keys"] if host else keys[:rows_e2e].cpu().numpy()
        hv_np = host["v_input"] if host else v[:rows_e2e].cpu().numpy()
        probe = min(rows_e2e, 2_000_000)
        t, _ = cb.timed(cb.aggregate_sum, hk_np[:probe], hv_np[:probe], cores, cores, KEYS_C4)
        sample = int(min(rows_e2e, max(probe, 12.0 / max(t, 1e-6) * probe / 3)))
        t, _ = cb.timed(cb.aggregate_sum, hk_np[:sample], hv_np[:sample], cores, cores, KEYS_C4, repeat=3)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads (per-partition hash "
                         "aggregation + merge), best of 3; columnar unboxed input, so an upper bound on the JVM reference"}
    host = {}
    for pb in pinned:
        pb.free()
    if rank == 0:
        peak, peak_src = hbm_peak()
        step_s = elapsed_ms / 1e3 / args.steps
        achieved = rows * 264 / step_s / 1e9
        print(json.dumps({
            "metric": "rows/s for aggregate (groupBy int64 key, reduce_sum over FloatType[?,64])",
            "value": world * rows / step_s, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"BASELINE configs[3]: aggregate groupBy int64 key ({n_groups} distinct), reduce_sum "
                                   f"FloatType[?,64], {rows} rows per GPU (config names 5e8; reduced to bound memory)",
                       "rows_per_gpu": rows, "l2": "inputs larger than L2 (no flush needed)", "exchange": "none" if world == 1 else "key-owner all_to_all of partial groups + merge aggregate"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic("c4", rows), "peak_source": peak_src, "kernel": "whole step (insert, table scan, pair sort, gather-reduce; agg_gather_reduce_kernel dominant)",
                         "algorithmic_bytes_per_launch": rows * 264.0, "frac_of_nominal_8000": achieved / 8000.0},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks.summary(), "cpu_baseline": cpu}))
    if world > 1:
        dist.destroy_process_group()


def run_c5(args, torch, dist, _native, world, rank, local, device, rows):
    """BASELINE config 5: map_blocks with MatMul(x[rows,256], W[256,256]) + bias on the tcgen05 path."""
    DA = _native.DeviceArray.from_torch
    graph = _native.Graph(build_graph("c5"))
    sess = _native.Session(graph, local)
    stream = torch.cuda.Stream(device=device)
    torch.cuda.set_stream(stream)
    sess.set_stream(stream.cuda_stream)
    g = torch.Generator(device=device)
    g.manual_seed(5 + 1000 * rank)
    x = torch.rand((rows, 256), generator=g, device=device, dtype=torch.float32)
    y = torch.empty((rows, 256), dtype=torch.float32, device=device)

    def step():
        sess.run_into({"x": DA(x)}, {"y": DA(y)})

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    launches0 = sess.launch_count
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        start.record(stream)
        for _ in range(args.steps):
            step()
        end.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    elapsed_ms = start.elapsed_time(end)
    tt = torch.tensor([elapsed_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    elapsed_ms = float(tt.item())
    launches = sess.launch_count - launches0
    # parity spot check against float64 on a slice (normwise 1e-5)
    rng = np.random.default_rng(5)
    w = torch.from_numpy(rng.uniform(-1 / 16, 1 / 16, (256, 256)).astype(np.float32)).to(device)
    b = torch.from_numpy(rng.uniform(-1, 1, 256).astype(np.float32)).to(device)
    xs = x[:65536].double()
    exact = xs @ w.double() + b.double()
    scale = xs.abs() @ w.double().abs() + b.double().abs()
    assert torch.all((y[:65536].double() - exact).abs() <= 1e-5 * scale)
    # ---- end to end: pinned host block in, pinned host block out (bounded block; PCIe-bound) ---------
    e2e = None
    rows_e2e = min(rows, 1 << 22)
    pinned = []
    hx = None
    if not args.no_e2e:
        try:
            px = _native.PinnedBuffer((rows_e2e, 256), np.float32); pinned.append(px)
            py = _native.PinnedBuffer((rows_e2e, 256), np.float32); pinned.append(py)
            torch.from_numpy(px.array).copy_(x[:rows_e2e])
            hx = px.array
            k_e2e = max(1, min(args.steps, 3))
            sess.run_into({"x": px.array}, {"y": py.array})
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                sess.run_into({"x": px.array}, {"y": py.array})      # H2D + kernel + D2H + sync inside
            if world > 1:
                dist.barrier()
            dt = time.perf_counter() - t0
            td = torch.tensor([dt], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(td, op=dist.ReduceOp.MAX)
            assert np.array_equal(py.array[:4096], y[:4096].cpu().numpy())
            e2e = {"value": world * rows_e2e * k_e2e / float(td.item()), "unit": "rows/s", "steps": k_e2e,
                   "rows_per_gpu": rows_e2e, "h2d_bytes_per_step": int(rows_e2e * 1024), "d2h_bytes_per_step": int(rows_e2e * 1024)}
        except Exception as exc:
            e2e = {"value": None, "unit": "rows/s", "error": str(exc)[:200],
                   "h2d_bytes_per_step": int(rows_e2e * 1024), "d2h_bytes_per_step": int(rows_e2e * 1024)}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_baseline as cb
        cores = cb.cores()
        xs_np = hx if hx is not None else x[:rows_e2e].cpu().numpy()
        w_np, b_np = w.cpu().numpy(), b.cpu().numpy()
        probe = min(rows_e2e, 65536)
        t, _ = cb.timed(cb.matmul_bias, xs_np[:probe], w_np, b_np, cores, cores)
        sample = int(min(rows_e2e, max(probe, 12.0 / max(t, 1e-6) * probe / 3)))
        t, _ = cb.timed(cb.matmul_bias, xs_np[:sample], w_np, b_np, cores, cores, repeat=3)
        cpu = {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads, best of 3; "
                         "plain C loops (the reference's TF-CPU build calls Eigen's GEMM, which would be faster than this port)"}
    hx = None
    for pb in pinned:
        pb.free()
    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        bf16 = float(peaks.get("bf16_tflops_sustained", 1400.0))
        step_s = elapsed_ms / 1e3 / args.steps
        tflops = rows * 131072.0 / step_s / 1e12
        hbm, _ = hbm_peak()
        print(json.dumps({
            "metric": "rows/s for mapBlocks (MatMul FloatType[?,256] x 256x256 + bias)",
            "value": world * rows / step_s, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (3xTF32 on tcgen05, fp32 accumulate)", "data": "synthetic",
            "config": {"workload": f"BASELINE configs[4]: map_blocks matmul FloatType[?,256] x 256x256 + bias, {rows} rows per GPU",
                       "rows_per_gpu": rows, "l2": "inputs larger than L2 (no flush needed)", "exchange": "none"},
            "roofline": {"bound": "tensor", "achieved": tflops, "peak": bf16, "unit": "TFLOP/s", "frac": tflops / bf16,
                         "traffic": ncu_traffic("c5", rows), "peak_source": "measured bf16 sustained; tf32 runs at half of it and "
                         "each fp32 product costs 3 tf32 MMAs, so the attainable ceiling is ~peak/6",
                         "kernel": "matmul_tc_pair_kernel", "mma_tflops_issued": 3 * tflops,
                         "hbm_gbs": rows * 2048.0 / step_s / 1e9, "hbm_frac": rows * 2048.0 / step_s / 1e9 / hbm,
                         "ncu_tensor_pipe_active_pct_of_elapsed": ncu_extra("c5", "tensor_pipe_active_pct_of_elapsed")},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks.summary(), "cpu_baseline": cpu}))
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline_leg(wl, host_feeds, rows, budget_s=12.0):
    """oracle/cpu_ref.c on a bounded sample: the whole block if it finishes within the budget,
    else a prefix sized from a calibration run."""
    from oracle import cpu_baseline as cb
    cores = cb.cores()
    if wl == "c2":
        y, z = host_feeds["y_input"], host_feeds["z_input"]
        probe = min(rows, 2_000_000)
        t, _ = cb.timed(cb.reduce_blocks_sum_min, y[:probe], z[:probe], cores, cores)
        sample = int(min(rows, max(probe, budget_s / max(t, 1e-6) * probe / 3)))
        t, (s, m) = cb.timed(cb.reduce_blocks_sum_min, y[:sample], z[:sample], cores, cores, repeat=3)
    else:
        x = host_feeds["x"]
        probe = min(rows, 4_000_000)
        t, _ = cb.timed(cb.map_blocks_chain, x[:probe], 1.5, 0.25, cores, cores)
        sample = int(min(rows, max(probe, budget_s / max(t, 1e-6) * probe / 3)))
        t, _ = cb.timed(cb.map_blocks_chain, x[:sample], 1.5, 0.25, cores, cores, repeat=3)
    return {"value": sample / t, "unit": "rows/s", "cores": cores, "kind": "port",
            "sample": f"first {sample} rows of the same block, {cores} partitions on {cores} threads, best of 3; "
                      "columnar unboxed input, so an upper bound on the JVM reference"}


def run_reference(args):
    """The reference arm: the restatement of the reference's own CPU algorithm (oracle/cpu_ref.c),
    all host threads, same workload / metric / unit. Rank 0 only."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    wl = args.workload
    cores = cb.cores()
    rows_full = args.rows or {"c2": ROWS_C2, "c3": ROWS_C3, "c4": ROWS_C4, "c5": ROWS_C5}[wl]
    rng = np.random.default_rng({"c2": 2, "c3": 3, "c4": 4, "c5": 5}[wl])
    # bounded sample: sized so that warmup + steps stay within a few minutes
    sample = min(rows_full, {"c2": 20_000_000, "c3": 50_000_000, "c4": 4_000_000, "c5": 262_144}[wl])
    if wl == "c2":
        u = rng.random(sample)
        y = np.stack([u, -u], axis=1)
        u = rng.random(sample)
        z = np.stack([u, -u], axis=1)
        fn = lambda: cb.reduce_blocks_sum_min(y, z, cores, cores)
    elif wl == "c3":
        x = rng.random(sample) * 2.0 - 1.0
        fn = lambda: cb.map_blocks_chain(x, 1.5, 0.25, cores, cores)
    elif wl == "c4":
        table = np.unique(rng.integers(-2**62, 2**62, KEYS_C4, dtype=np.int64))
        keys = table[rng.integers(0, table.size, sample)]
        v = rng.random((sample, 64), dtype=np.float32)
        fn = lambda: cb.aggregate_sum(keys, v, cores, cores, KEYS_C4)
    else:
        w = rng.uniform(-1 / 16, 1 / 16, (256, 256)).astype(np.float32)
        b = rng.uniform(-1, 1, 256).astype(np.float32)
        x = rng.random((sample, 256), dtype=np.float32)
        fn = lambda: cb.matmul_bias(x, w, b, cores, cores)
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    desc = (f"{sample} rows per step (bounded sample of the {rows_full}-row block), {cores} partitions on "
            f"{cores} threads; oracle/cpu_ref.c restates the reference's per-block algorithm on columnar input "
            "(Spark + TF-Java are not installable here)")
    line = {
        "impl": "reference",
        "metric": {"c2": "rows/s for reduceBlocks (reduce_sum + reduce_min over DoubleType[?,2])",
                   "c3": "rows/s for mapBlocks (fused x*a+b, max(.,0), cast f32 on a float64 column)",
                   "c4": "rows/s for aggregate (groupBy int64 key, reduce_sum over FloatType[?,64])",
                   "c5": "rows/s for mapBlocks (MatMul FloatType[?,256] x 256x256 + bias)"}[wl],
        "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64" if wl in ("c2", "c3") else "f32", "data": "synthetic",
        "config": {"workload": {"c2": "BASELINE configs[1]: reduce_blocks reduce_sum+reduce_min, DoubleType[?,2], ",
                                "c3": "BASELINE configs[2]: map_blocks fused chain, float64 column, ",
                                "c4": "BASELINE configs[3]: aggregate groupBy int64 key, reduce_sum FloatType[?,64], ",
                                "c5": "BASELINE configs[4]: map_blocks matmul FloatType[?,256] x 256x256 + bias, "}[wl]
                               + f"{rows_full} rows per GPU", "rows_per_gpu": rows_full},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
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
