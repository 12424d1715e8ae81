"""Worker of tests/test_multi_gpu_comm.py: one rank per GPU under torchrun.  Exercises the exchange steps through
the C ABI (tfs_comm_create, tfs_session_reduce_combine, tfs_session_aggregate_combine) and checks them against numpy
on the data of every rank (which each rank regenerates from the seeds)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import tensorframes_b200 as tfs  # noqa: E402
from tensorframes_b200 import _native, dsl  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    ident = [_native.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ident, 0)
    comm = _native.Comm(local, world, rank, ident[0])
    assert comm.world == world and comm.rank == rank
    DA = _native.DeviceArray.from_torch
    dev = torch.device("cuda", local)

    def block(r, n, cell, seed):
        return np.random.default_rng(1000 * seed + r).uniform(-1, 1, (n,) + cell)

    # ---- reduce_blocks exchange: bare Sum / Min (all-reduce path), in order and overlapped -----------------
    with dsl.withGraph():
        y = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 3).named("y_input")
        z = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 3).named("z_input")
        g = dsl.build_graph(dsl.reduce_sum(y, [0]).named("y"), dsl.reduce_min(z, [0]).named("z"))
    n = 50_000 + 17 * rank
    mine = block(rank, n, (3,), 1)
    everyone = np.concatenate([block(r, 50_000 + 17 * r, (3,), 1) for r in range(world)])
    with _native.Graph(g) as gr, _native.Session(gr, local) as s:
        d_in = torch.from_numpy(mine).to(dev)
        outs = {k: torch.empty(3, dtype=torch.float64, device=dev) for k in ("y", "z")}
        merged = {k: torch.empty(3, dtype=torch.float64, device=dev) for k in ("y", "z")}
        torch.cuda.synchronize()
        for overlap in (False, True, True, True, True, True, True):       # more overlapped rounds than ring slots
            s.run_into({"y_input": DA(d_in), "z_input": DA(d_in)}, {k: DA(v) for k, v in outs.items()})
            s.reduce_combine(comm, {k: DA(v) for k, v in outs.items()}, {k: DA(v) for k, v in merged.items()}, overlap=overlap)
            if overlap:
                for v in outs.values():
                    v.zero_()            # partials may be overwritten at once: the exchange works on its own copy
                s.combine_wait(comm)
            s.sync()
            assert np.all(np.abs(merged["y"].cpu().numpy() - everyone.sum(axis=0)) <= 1e-12 * np.abs(everyone).sum(axis=0))
            assert np.array_equal(merged["z"].cpu().numpy(), everyone.min(axis=0))
            for v in merged.values():
                v.zero_()
            torch.cuda.synchronize()

    # ---- reduce_blocks exchange: a reducer graph that is not a bare reduction (all-gather + merge run) ------
    with dsl.withGraph():
        x = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2, 2).named("x_input")
        g = dsl.build_graph(dsl.neg(dsl.reduce_max(dsl.neg(x), [0])).named("x"))      # = min, written the long way
    with _native.Graph(g) as gr, _native.Session(gr, local) as s:
        mine = block(rank, 999, (2, 2), 2)
        everyone = np.concatenate([block(r, 999, (2, 2), 2) for r in range(world)])
        part = torch.empty((2, 2), dtype=torch.float64, device=dev)
        merged = torch.empty((2, 2), dtype=torch.float64, device=dev)
        s.run_into({"x_input": mine}, {"x": DA(part)})
        s.reduce_combine(comm, {"x": DA(part)}, {"x": DA(merged)}, overlap=True)    # falls back to stream order
        s.sync()
        assert np.array_equal(merged.cpu().numpy(), everyone.min(axis=0))

    # ---- aggregate exchange: groups travel to the owner of their key and are merged there ---------------
    with dsl.withGraph():
        v = dsl.placeholder(tfs.FloatType, dsl.Unknown, 8).named("v_input")
        c = dsl.placeholder(tfs.LongType, dsl.Unknown).named("c_input")
        g = dsl.build_graph(dsl.reduce_sum(v, [0]).named("v"), dsl.reduce_max(c, [0]).named("c"))

    def agg_block(r):
        rng = np.random.default_rng(77 + r)
        table = np.random.default_rng(5).integers(-2**62, 2**62, 5000)           # one key universe for all ranks
        k = table[rng.integers(0, 5000, 200_000)].astype(np.int64)
        return k, rng.uniform(0, 1, (200_000, 8)).astype(np.float32), rng.integers(-10**9, 10**9, 200_000).astype(np.int64)

    k, vv, cc = agg_block(rank)
    with _native.Graph(g) as gr, _native.Session(gr, local) as s:
        cap = 8192
        ok = torch.empty(cap, dtype=torch.int64, device=dev)
        ov = torch.empty((cap, 8), dtype=torch.float32, device=dev)
        oc = torch.empty(cap, dtype=torch.int64, device=dev)
        ng = s.aggregate_into(DA(torch.from_numpy(k).to(dev)), {"v_input": vv, "c_input": cc}, {"v": DA(ov), "c": DA(oc)}, DA(ok))
        mk = torch.empty(cap, dtype=torch.int64, device=dev)
        mv = torch.empty((cap, 8), dtype=torch.float32, device=dev)
        mc = torch.empty(cap, dtype=torch.int64, device=dev)
        for _ in range(2):
            owned = s.aggregate_combine(comm, ng, DA(ok), {"v": DA(ov), "c": DA(oc)}, {"v": DA(mv), "c": DA(mc)}, DA(mk))
        got_k, got_v, got_c = mk[:owned].cpu().numpy(), mv[:owned].cpu().numpy(), mc[:owned].cpu().numpy()
    from tensorframes_b200.distributed import owner_of_key
    blocks = [agg_block(r) for r in range(world)]
    all_k = np.concatenate([b[0] for b in blocks])
    all_v = np.concatenate([b[1] for b in blocks])
    all_c = np.concatenate([b[2] for b in blocks])
    uniq, inv = np.unique(all_k, return_inverse=True)
    mine_mask = owner_of_key(uniq, world) == rank
    assert np.array_equal(got_k, uniq[mine_mask])                                 # exactly the keys this rank owns, ascending
    want_v = np.zeros((len(uniq), 8))
    np.add.at(want_v, inv, all_v.astype(np.float64))
    want_c = np.full(len(uniq), -2**63, dtype=np.int64)
    np.maximum.at(want_c, inv, all_c)
    assert np.all(np.abs(got_v - want_v[mine_mask]) <= 1e-5 * want_v[mine_mask])
    assert np.array_equal(got_c, want_c[mine_mask])
    counts = [None] * world
    dist.all_gather_object(counts, int(owned))
    assert sum(counts) == len(uniq)                                                # every key has exactly one owner
    comm.close()
    dist.barrier()
    if rank == 0:
        print("comm worker ok", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
