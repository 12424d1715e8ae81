"""Do the phases of the keyed aggregate overlap when two blocks are reduced at once?  Two sessions (own streams, own
workspaces) reduce the same 2.5e8-row block one after the other, then from two host threads at the same time.
TFS_AGG_GATHER_PAD_KB was a temporary switch (unused dynamic shared memory on the gather launch, capping its CTAs per SM
so that the other session's insert / partition CTAs find room); the library no longer reads it."""
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

rows = int(os.environ.get("ROWS", 250_000_000))
torch.cuda.set_device(0)
cols = bench.device_columns(torch, "c4", 0, rows, "cuda")
keys, v = cols["keys"], cols["v_input"]
DA = _native.DeviceArray.from_torch
with dsl.withGraph():
    p = dsl.placeholder(tfs.FloatType, dsl.Unknown, 64).named("v_input")
    g = dsl.build_graph(dsl.reduce_sum(p, [0]).named("v"))
G = 1_000_000
sessions, outs = [], []
graph = _native.Graph(g)
for i in range(2):
    s = _native.Session(graph, 0)
    st = torch.cuda.Stream()
    s.set_stream(st.cuda_stream)
    sessions.append((s, st))
    outs.append((torch.empty((G, 64), dtype=torch.float32, device="cuda"), torch.empty(G, dtype=torch.int64, device="cuda")))


def one(i):
    s, _ = sessions[i]
    o, ok = outs[i]
    return s.aggregate_into(DA(keys), {"v_input": DA(v)}, {"v": DA(o)}, DA(ok))


for i in range(2):
    for _ in range(2):
        one(i)
torch.cuda.synchronize()
reps = 5
t0 = time.perf_counter()
for _ in range(reps):
    one(0)
    one(1)
torch.cuda.synchronize()
seq = (time.perf_counter() - t0) / reps


def loop(i):
    for _ in range(reps):
        one(i)


t0 = time.perf_counter()
th = [threading.Thread(target=loop, args=(i,)) for i in range(2)]
for t in th:
    t.start()
for t in th:
    t.join()
torch.cuda.synchronize()
conc = (time.perf_counter() - t0) / reps
assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
print("rows %d  pad_kb %s  two blocks one after the other %.2f ms, at the same time %.2f ms" % (rows, os.environ.get("TFS_AGG_GATHER_PAD_KB", "0"), seq * 1e3, conc * 1e3))
