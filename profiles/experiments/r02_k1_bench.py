"""K1 micro-benchmark: achieved HBM GB/s (bytes read + written / CUDA-event time) of fused element-wise groups on
device-resident columns, chain kernel vs interpreter (TFS_ELT_IMPL=interp).  usage: python r02_k1_bench.py [rows]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000_000
dev = torch.device("cuda:0")
x = torch.rand(n, dtype=torch.float64, device=dev) * 2 - 1
y = torch.rand(n, dtype=torch.float64, device=dev) + 1
DA = _native.DeviceArray.from_torch
D, F = tfs.DoubleType, tfs.FloatType


def graphs():
    def ph():
        return dsl.placeholder(D, dsl.Unknown).named("x"), dsl.placeholder(D, dsl.Unknown).named("y")
    out = []
    with dsl.withGraph():
        a, b = ph()
        out.append(("c3 chain", dsl.build_graph(dsl.cast(dsl.maximum(a * 1.5 + 0.25, 0.0), F).named("z")), ["x"], {"z": torch.float32}))
    with dsl.withGraph():
        a, b = ph()
        d = a - b
        out.append(("(x-y)^2/(y^2+1)", dsl.build_graph(dsl.div(d * d, b * b + 1.0).named("z")), ["x", "y"], {"z": torch.float64}))
    with dsl.withGraph():
        a, b = ph()
        t = a
        for k in range(6):
            t = (t + b) * 0.5
        out.append(("12-slot chain", dsl.build_graph(t.named("z")), ["x", "y"], {"z": torch.float64}))
    with dsl.withGraph():
        a, b = ph()
        out.append(("mid-chain cast", dsl.build_graph((dsl.cast(a * 2.0, F) * dsl.constant(__import__("numpy").float32(3.0))).named("z")), ["x"], {"z": torch.float32}))
    with dsl.withGraph():
        a, b = ph()
        out.append(("two fetches", dsl.build_graph((a * 2.0).named("p"), (a + 3.0).named("q")), ["x"], {"p": torch.float64, "q": torch.float64}))
    with dsl.withGraph():
        a, b = ph()
        s_ = a + b
        d_ = a - b
        out.append(("(x+y)*(x-y)+x*y", dsl.build_graph((s_ * d_ + a * b).named("z")), ["x", "y"], {"z": torch.float64}))
    return out


res = {}
for name, g, ins, outs in graphs():
    feeds = {k: DA({"x": x, "y": y}[k]) for k in ins}
    ot = {k: torch.empty(n, dtype=dt, device=dev) for k, dt in outs.items()}
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        st = torch.cuda.Stream()
        s.set_stream(st.cuda_stream)
        torch.cuda.synchronize()
        for _ in range(3):
            s.run_into(feeds, {k: DA(v) for k, v in ot.items()})
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(10):
            s.run_into(feeds, {k: DA(v) for k, v in ot.items()})
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
    byt = len(ins) * 8 * n + sum(v.element_size() * n for v in ot.values())
    res[name] = round(byt / ms / 1e6, 1)
print(json.dumps({"impl": os.environ.get("TFS_ELT_IMPL", "default"), "rows": n, "GBps": res}))
