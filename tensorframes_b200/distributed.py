"""One process per GPU: partitions map one-to-one onto ranks.

mapBlocks shards with no communication (reference: rdd.mapPartitions,
impl/DebugRowOps.scala:377-391).  reduceBlocks has one exchange step: the reference
pulls every partition's result row to the driver and merges them by running the same
graph on 2-row blocks (RDD.reduce with reducePairBlock, impl/DebugRowOps.scala:524-525,
741-750).  Here every rank all-gathers the cell-sized partial rows and runs the graph
once more on the stacked [world, cell...] block, which is the same merge for the
associative reducers the X / X_input contract allows -- and it works for any reducer
graph, not only Sum/Min/Max.  torch.distributed is the plumbing (NCCL on GPUs, gloo in
the CPU tests); the merge itself runs through the same C-ABI session as everything else.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence

import numpy as np


def shard_rows(n_rows: int, rank: int, world: int):
    """Contiguous row range of `rank` (Spark partition <-> GPU, one to one)."""
    return (n_rows * rank) // world, (n_rows * (rank + 1)) // world


def gather_partials(partials: Sequence, dist=None) -> List:
    """All-gather each rank's partial result rows.  `partials` are torch tensors (on the rank's
    device for NCCL, on the CPU for gloo); returns one [world, cell...] tensor per fetch."""
    import torch
    import torch.distributed as td
    dist = dist or td
    world = dist.get_world_size()
    flats = [torch.empty(world * p.numel(), dtype=p.dtype, device=p.device) for p in partials]
    srcs = [p.contiguous().reshape(-1) for p in partials]     # flat form works on NCCL and gloo alike
    coalesce = getattr(dist, "_coalescing_manager", None)
    if coalesce is not None and len(partials) > 1 and partials[0].is_cuda:
        # one NCCL group = one launch for all fetches: the rows are a few bytes, so the step is pure launch latency
        with coalesce(device=partials[0].device):
            for flat, src in zip(flats, srcs):
                dist.all_gather_into_tensor(flat, src)
    else:
        for flat, src in zip(flats, srcs):
            dist.all_gather_into_tensor(flat, src)
    return [flat.reshape((world,) + tuple(p.shape)) for flat, p in zip(flats, partials)]


def combine_reduce_blocks(partials: Sequence, names: Sequence[str], run_graph: Callable[[Dict[str, object]], List]):
    """The exchange step of reduceBlocks.  `run_graph(feeds)` evaluates the reducer graph on a
    block (feeds keyed `X_input`) and returns the fetches in the order of `names`."""
    stacked = gather_partials(partials)
    return run_graph({n + "_input": s for n, s in zip(names, stacked)})


def owner_of_key(keys: np.ndarray, world: int) -> np.ndarray:
    """Which rank owns a group key in aggregate's exchange step (a fixed 64-bit mix, so every
    rank and every run agrees)."""
    x = keys.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    # the high half: the low bits of the same mix pick the hash-table slot of the merge on the owner, and a rank whose keys
    # all agree in them modulo a power-of-two world would use one slot in `world` as home slot
    return ((x >> np.uint64(32)) % np.uint64(world)).astype(np.int64)


def owner_of_key_torch(keys, world: int):
    """owner_of_key on a torch int64 tensor (any device), bit-identical to the numpy version:
    the 64-bit mix is done in wrapping int64 arithmetic with logical shifts emulated by masks."""
    import torch

    def lsr(x, k):   # logical shift right on int64
        return (x >> k) & ((1 << (64 - k)) - 1)

    def mul(x, c):   # wrapping multiply by an unsigned 64-bit constant
        return x * torch.tensor(c - (1 << 64) if c >= (1 << 63) else c, dtype=torch.int64, device=x.device)

    x = keys.clone()
    x = x ^ lsr(x, 30)
    x = mul(x, 0xBF58476D1CE4E5B9)
    x = x ^ lsr(x, 27)
    x = mul(x, 0x94D049BB133111EB)
    x = x ^ lsr(x, 31)
    return lsr(x, 32) % world      # (the high half: non-negative in int64)


def exchange_groups(keys, values: Sequence, dist=None):
    """The exchange step of aggregate: every rank holds partial groups (keys [G], values[i] [G, cell...]);
    afterwards every rank holds all partial rows of the keys it owns (owner = high half of mix(key), mod world), ready
    for one more keyed reduction.  The reference does this with Spark's hash shuffle between the UDAF's
    partial and final aggregation (impl/DebugRowOps.scala:573-576).  Returns (keys, values) of the
    received rows; row order is by source rank."""
    import torch
    import torch.distributed as td
    dist = dist or td
    world = dist.get_world_size()
    owner = owner_of_key_torch(keys, world)
    order = torch.argsort(owner, stable=True)
    send_counts = torch.bincount(owner, minlength=world).to(torch.int64)
    recv_counts = torch.empty_like(send_counts)
    dist.all_to_all_single(recv_counts, send_counts)
    send_splits = send_counts.tolist()
    recv_splits = recv_counts.tolist()
    n_recv = int(sum(recv_splits))

    def a2a(t):
        t = t[order].contiguous()
        out = torch.empty((n_recv,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_to_all_single(out, t, output_split_sizes=recv_splits, input_split_sizes=send_splits)
        return out

    return a2a(keys), [a2a(v) for v in values]
