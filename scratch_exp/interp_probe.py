"""Experiment: achieved HBM bandwidth of fused element-wise groups that need the interpreter (not linear chains)."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

DA = _native.DeviceArray.from_torch
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)

def run(name, build, feeds, outs, bytes_moved):
    with dsl.withGraph():
        g = build()
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        s.set_stream(stream.cuda_stream)
        f = {k: DA(v) for k, v in feeds.items()}
        o = {k: DA(v) for k, v in outs.items()}
        for _ in range(3):
            s.run_into(f, o)
        torch.cuda.synchronize()
        l0 = s.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(10):
            s.run_into(f, o)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{name:58s} {ms:7.3f} ms  {bytes_moved / ms / 1e6:7.0f} GB/s  ({(s.launch_count - l0) // 10} launches)")

n = 400_000_000
x = torch.rand(n, dtype=torch.float64, device=dev); y = torch.rand(n, dtype=torch.float64, device=dev) + 1
o1 = torch.empty(n, dtype=torch.float64, device=dev); o2 = torch.empty(n, dtype=torch.float64, device=dev)
P = lambda nm, dt=tfs.DoubleType, *sh: dsl.placeholder(dt, dsl.Unknown, *sh).named(nm)
run("chain  z = x*1.5+0.25 (f64 -> f64)", lambda: dsl.build_graph((P("x") * 1.5 + 0.25).named("z")), {"x": x}, {"z": o1}, 16 * n)
run("chain  z = x + y", lambda: dsl.build_graph((P("x") + P("y")).named("z")), {"x": x, "y": y}, {"z": o1}, 24 * n)
def dag():
    a, b = P("x"), P("y")
    return dsl.build_graph(((a + b) * (a - b)).named("z"))
run("interp z = (x+y)*(x-y)", dag, {"x": x, "y": y}, {"z": o1}, 24 * n)
def two_out():
    a = P("x")
    return dsl.build_graph((a * 2.0).named("y2"), (a + 3.0).named("z"))
run("interp y = x*2, z = x+3 (two fetches)", two_out, {"x": x}, {"y2": o1, "z": o2}, 24 * n)
def xx():
    a = P("x")
    return dsl.build_graph((a + a).named("z"))
run("?      z = x + x", xx, {"x": x}, {"z": o1}, 16 * n)
del x, y, o1, o2
rows, c = 8_000_000, 64
xf = torch.rand((rows, c), dtype=torch.float32, device=dev); of = torch.empty_like(xf)
w = np.random.default_rng(0).uniform(-1, 1, c).astype(np.float32); b = np.random.default_rng(1).uniform(-1, 1, c).astype(np.float32)
run("interp relu(x*w+b), w,b row vectors, f32[n,64]", lambda: dsl.build_graph(dsl.relu(P("x", tfs.FloatType, c) * dsl.constant(w) + dsl.constant(b)).named("z")), {"x": xf}, {"z": of}, 8 * rows * c)
