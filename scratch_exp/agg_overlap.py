"""Experiment: throughput of the keyed aggregate with 1, 2, 3 blocks in flight (one session + stream + host thread each)."""
import sys, threading, time
sys.path.insert(0, ".")
import torch
import bench
from tensorframes_b200 import _native

rows = 250_000_000 // 2     # two resident blocks of half the bench size keep memory in bounds
device = torch.device("cuda", 0)
DA = _native.DeviceArray.from_torch
graph = _native.Graph(bench.build_graph("c4"))
blocks = [bench.make_inputs(torch, "c4", rows, seed=4 + i, device=device) for i in range(2)]

def worker(idx, n_steps, barrier, times):
    sess = _native.Session(graph, 0)
    stream = torch.cuda.Stream(device=device)
    sess.set_stream(stream.cuda_stream)
    t = blocks[idx % len(blocks)]
    out = torch.empty((bench.KEYS_C4, 64), dtype=torch.float32, device=device)
    out_keys = torch.empty(bench.KEYS_C4, dtype=torch.int64, device=device)
    sess.aggregate_into(DA(t["keys"]), {"v_input": DA(t["v_input"])}, {"v": DA(out)}, DA(out_keys))
    barrier.wait()
    t0 = time.perf_counter()
    for _ in range(n_steps):
        sess.aggregate_into(DA(t["keys"]), {"v_input": DA(t["v_input"])}, {"v": DA(out)}, DA(out_keys))
    times[idx] = time.perf_counter() - t0

for n_threads in (1, 2, 3, 4):
    n_steps = 6
    barrier = threading.Barrier(n_threads)
    times = [0.0] * n_threads
    th = [threading.Thread(target=worker, args=(i, n_steps, barrier, times)) for i in range(n_threads)]
    for x in th: x.start()
    for x in th: x.join()
    total = n_threads * n_steps * rows
    print(f"{n_threads} blocks in flight: {max(times) / n_steps * 1e3 / n_threads:.2f} ms per {rows}-row block, {total / max(times):.3e} rows/s")
