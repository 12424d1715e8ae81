"""The exchange steps of reduceBlocks and aggregate through the C ABI, on real GPUs: one rank per GPU under torchrun
(tests/_comm_worker.py), and several devices driven by the threads of one process.  Needs at least two GPUs."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import tensorframes_b200 as tfs
from tensorframes_b200 import _native, dsl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(_native.device_count() < 2, reason="needs two GPUs")]


@pytest.mark.timeout(600)
def test_exchange_steps_one_rank_per_gpu():
    world = min(_native.device_count(), 4)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                          "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "tests", "_comm_worker.py")],
                         capture_output=True, text=True, timeout=540)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert f"comm worker ok {world}" in out.stdout


@pytest.mark.timeout(300)
def test_devices_of_one_process_one_thread_each():
    """tfs_comm_create_all: one thread creates the communicators, then every device is driven by its own thread (the
    way local[N] task threads of one JVM would)."""
    torch = pytest.importorskip("torch")
    n_dev = min(_native.device_count(), 4)
    comms = _native.Comm.create_all(list(range(n_dev)))
    with dsl.withGraph():
        y = dsl.placeholder(tfs.DoubleType, dsl.Unknown, 2).named("y_input")
        g = dsl.build_graph(dsl.reduce_sum(y, [0]).named("y"))
    blocks = [np.random.default_rng(d).uniform(0, 1, (30_000 + d, 2)) for d in range(n_dev)]
    results, errors = [None] * n_dev, []

    def work(d):
        try:
            with _native.Graph(g) as gr, _native.Session(gr, d) as s:
                part = torch.empty(2, dtype=torch.float64, device=f"cuda:{d}")
                merged = torch.empty(2, dtype=torch.float64, device=f"cuda:{d}")
                DA = _native.DeviceArray.from_torch
                s.run_into({"y_input": blocks[d]}, {"y": DA(part)})
                s.reduce_combine(comms[d], {"y": DA(part)}, {"y": DA(merged)})
                s.sync()
                results[d] = merged.cpu().numpy()
        except Exception as exc:
            errors.append((d, repr(exc)))

    threads = [threading.Thread(target=work, args=(d,)) for d in range(n_dev)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    total = np.concatenate(blocks)
    for r in results:
        assert np.all(np.abs(r - total.sum(axis=0)) <= 1e-12 * np.abs(total).sum(axis=0))
        assert np.array_equal(r, results[0])              # every rank holds the same merged row, bit for bit
    for c in comms:
        c.close()


@pytest.mark.timeout(300)
def test_operators_over_a_device_list_match_one_device():
    """B200Ops(devices=[...]): partition i runs on devices[i % P]; reduce_blocks and aggregate combine the per-device
    results through the communicators.  Same answers as one device (sums within tolerance, the rest bit for bit)."""
    import _golden
    from tensorframes_b200.frame import TensorFrame
    from tensorframes_b200.ops import B200Ops
    rng = np.random.default_rng(2024)
    n = 40_000
    cols = {"k": rng.integers(0, 500, n).astype(np.int64), "x": rng.uniform(-1, 1, (n, 4))}
    rows = [(int(cols["k"][i]), f"t{i % 7}", cols["x"][i].tolist()) for i in range(n)]
    df = tfs.analyze(TensorFrame.from_rows(rows, ["k", "tag", "x"], num_partitions=5))
    one, many = B200Ops(0), B200Ops(devices=list(range(min(_native.device_count(), 4))))
    with dsl.withGraph():
        z = (tfs.block(df, "x") * 2.0 + 1.0).named("z")
        a, b = one.mapBlocks(df, z), many.mapBlocks(df, z)
    assert a.columns == b.columns and len(a.partitions) == len(b.partitions)
    for pa, pb in zip(a.partitions, b.partitions):
        assert np.array_equal(pa["z"], pb["z"])
    with dsl.withGraph():
        xi = tfs.block(df, "x", tf_name="x_input")
        s_ = dsl.reduce_sum(xi, [0]).named("x")
        r1, r2 = one.reduceBlocks(df.select("x"), s_), many.reduceBlocks(df.select("x"), s_)
        g1 = one.aggregate(df.select("k", "x").groupBy("k"), s_)
        g2 = many.aggregate(df.select("k", "x").groupBy("k"), s_)
    with dsl.withGraph():
        xi = tfs.block(df, "x", tf_name="x_input")
        m_ = dsl.reduce_min(xi, [0]).named("x")
        assert one.reduceBlocks(df.select("x"), m_) == many.reduceBlocks(df.select("x"), m_)
        t1 = one.aggregate(df.select("k", "tag", "x").groupBy("tag", "k"), m_)
        t2 = many.aggregate(df.select("k", "tag", "x").groupBy("tag", "k"), m_)
    assert np.allclose(np.asarray(r1.x), np.asarray(r2.x), rtol=0, atol=1e-9)
    assert len(g2.partitions) > 1                       # sharded by key owner
    rows1 = sorted(g1.collect(), key=lambda r: r.k)
    rows2 = sorted(g2.collect(), key=lambda r: r.k)
    assert [r.k for r in rows1] == [r.k for r in rows2] == sorted(set(cols["k"].tolist()))
    assert np.allclose(np.asarray([r.x for r in rows1]), np.asarray([r.x for r in rows2]), rtol=0, atol=1e-9)
    assert _golden.normalize(t1.collect()) == _golden.normalize(t2.collect())
