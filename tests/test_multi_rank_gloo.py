"""The N > 1 host logic on CPU: two gloo ranks, one partition each.  The device work is replaced by
the oracle here (no GPU in this container); what is under test is the sharding, the exchange of the
partial rows and the merge run -- the same code bench.py --gpus N and a multi-GPU job go through."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import tensorframes_b200 as tfs
        from oracle import tf_ops
        from tensorframes_b200 import dsl
        from tensorframes_b200.distributed import combine_reduce_blocks, owner_of_key, shard_rows

        rng = np.random.default_rng(5)               # every rank draws the same full block ...
        y = rng.uniform(0, 1, (10_001, 2))
        z = rng.uniform(-1, 1, (10_001, 2))
        lo, hi = shard_rows(y.shape[0], rank, world)  # ... and keeps its own partition
        with dsl.withGraph():
            yi = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
            zi = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("z_input")
            g = dsl.build_graph(dsl.reduce_sum(yi, [0]).named("y"), dsl.reduce_min(zi, [0]).named("z"))
        names = ["y", "z"]

        def run_graph(feeds):                         # stands in for session.run on the rank's GPU
            return tf_ops.evaluate(g, {k: np.asarray(v) for k, v in feeds.items()}, names)

        partial = run_graph({"y_input": y[lo:hi], "z_input": z[lo:hi]})
        merged = combine_reduce_blocks([torch.from_numpy(p) for p in partial], names, run_graph)
        whole = run_graph({"y_input": y, "z_input": z})
        assert np.all(np.abs(merged[0] - whole[0]) <= 1e-12 * np.abs(y).sum(axis=0))
        assert np.array_equal(merged[1], whole[1])
        # every rank ends with the same merged row (the reference returns it to the driver)
        t = torch.from_numpy(np.concatenate([merged[0], merged[1]]))
        ref = t.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(t, ref)
        # shards tile the block exactly, and key ownership is a partition of the key space
        spans = [shard_rows(10_001, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == 10_001 and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        keys = rng.integers(-2**62, 2**62, 1000)
        own = owner_of_key(keys, world)
        assert own.min() >= 0 and own.max() < world and np.array_equal(own, owner_of_key(keys.copy(), world))
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_reduce_blocks_exchange_two_ranks(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]


def _agg_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tensorframes_b200.distributed import exchange_groups, owner_of_key, owner_of_key_torch, shard_rows

        rng = np.random.default_rng(11)
        n, n_keys = 6000, 97
        table = rng.integers(-2**63, 2**63 - 1, n_keys, dtype=np.int64)
        keys = table[rng.integers(0, n_keys, n)]
        v = rng.integers(-50, 50, (n, 3)).astype(np.int64)         # integers: the merged sums must be exact
        assert np.array_equal(owner_of_key_torch(torch.from_numpy(table), world).numpy(), owner_of_key(table, world))
        lo, hi = shard_rows(n, rank, world)

        def local_aggregate(k, x):      # stands in for tfs_session_aggregate on the rank's GPU
            uk, inv = np.unique(k, return_inverse=True)
            out = np.zeros((len(uk), x.shape[1]), dtype=x.dtype)
            np.add.at(out, inv, x)
            return uk, out

        pk, pv = local_aggregate(keys[lo:hi], v[lo:hi])                                 # partial groups of this partition
        rk, (rv,) = exchange_groups(torch.from_numpy(pk), [torch.from_numpy(pv)])      # key-owner exchange
        fk, fv = local_aggregate(rk.numpy(), rv.numpy())                                # final merge on the owner
        wk, wv = local_aggregate(keys, v)
        mine = owner_of_key(wk, world) == rank
        assert np.array_equal(fk, wk[mine]) and np.array_equal(fv, wv[mine])           # every key once, on its owner
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_aggregate_exchange_two_ranks(tmp_path):
    world = 2
    mp.spawn(_agg_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]
