"""Byte-level mutations of valid GraphDefs: the library reads graph bytes a caller hands it, so a damaged graph has
to end in a status code, never in a crash.  Shared by a CPU test (import only) and a GPU test (import, plan, run)."""
import numpy as np

import tensorframes_b200 as tfs
from tensorframes_b200 import dsl


def seed_graphs():
    import test_kmeans_ops as tk
    _, centers = tk.data(n=16)
    graphs = [tk.step1_graph(centers, 6)[0], tk.step2_graph(centers, 6)[0]]
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, dsl.Unknown, 4).named("x")
        y = dsl.relu(dsl.matmul(p, dsl.constant(np.ones((4, 3), np.float32))) + dsl.constant(np.ones(3, np.float32)))
        graphs.append(dsl.build_graph(y.named("y"), dsl.reduce_sum(p, [0]).named("s"), dsl.cast(p, tfs.LongType).named("c")))
    return graphs


def mutate(rng, graph_bytes, structural):
    g = bytearray(graph_bytes)
    kind = int(rng.integers(0, 5 if structural else 3))
    if kind == 0:
        for _ in range(int(rng.integers(1, 4))):
            g[int(rng.integers(0, len(g)))] ^= 1 << int(rng.integers(0, 8))
    elif kind == 1:
        g[int(rng.integers(0, len(g)))] = int(rng.integers(0, 256))
    elif kind == 2:
        i = int(rng.integers(0, len(g)))
        g[i] = (g[i] + int(rng.integers(-2, 3))) & 255
    elif kind == 3:
        g = g[:int(rng.integers(0, len(g)))]
    else:
        i = int(rng.integers(0, len(g)))
        g[i:i] = bytes(rng.integers(0, 256, int(rng.integers(1, 9)), dtype=np.uint8))
    return bytes(g)
