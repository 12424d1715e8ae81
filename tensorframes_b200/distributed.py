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
    out = []
    for p in partials:
        flat = torch.empty(world * p.numel(), dtype=p.dtype, device=p.device)
        dist.all_gather_into_tensor(flat, p.contiguous().reshape(-1))   # flat form works on NCCL and gloo alike
        out.append(flat.reshape((world,) + tuple(p.shape)))
    return out


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
    return (x % np.uint64(world)).astype(np.int64)
