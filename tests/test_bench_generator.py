"""bench.py's synthetic-column generator is written twice -- numpy for the host legs (CPU baseline, reference arm,
oracle checks) and torch integer ops for the device -- and both must produce the same bits, or the CPU arm and the
GPU arm would not be reducing the same values.  torch runs on the CPU here; the integer arithmetic is the same."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def test_mix_matches_python_integers():
    xs = [0, 1, 2**63, 2**64 - 1, 0x123456789ABCDEF]
    got = bench.mix_np(np.asarray(xs, dtype=np.uint64))
    assert [int(g) for g in got] == [bench._mix_py(x) for x in xs]


@pytest.mark.parametrize("wl", ["c2", "c3", "c4", "c5"])
@pytest.mark.parametrize("row0", [0, 999_999_937])
def test_host_and_device_generators_agree(wl, row0):
    torch = pytest.importorskip("torch")
    rows = 3001 if wl in ("c4", "c5") else 70_001
    h = bench.host_columns(wl, row0, rows)
    d = bench.device_columns(torch, wl, row0, rows, torch.device("cpu"), chunk=4096)
    assert sorted(h) == sorted(d)
    for k in h:
        assert h[k].dtype == d[k].numpy().dtype and h[k].tobytes() == d[k].numpy().tobytes(), k


def test_columns_have_the_stated_shape_and_range():
    c2 = bench.host_columns("c2", 0, 1000)
    assert np.array_equal(c2["y_input"][:, 1], -c2["y_input"][:, 0]) and c2["y_input"][:, 0].min() >= 0 and c2["y_input"].max() < 1
    assert not np.array_equal(c2["y_input"], c2["z_input"])
    x = bench.host_columns("c3", 0, 100_000)["x"]
    assert -1 <= x.min() < -0.99 and 0.99 < x.max() < 1 and abs(x.mean()) < 0.01
    c4 = bench.host_columns("c4", 0, 50_000)
    assert c4["v_input"].shape == (50_000, 64) and c4["v_input"].dtype == np.float32 and 0 <= c4["v_input"].min() and c4["v_input"].max() < 1
    # keys come from a table of KEYS_C4 distinct 64-bit values: sparse, not dense ids
    table = bench.mix_np(np.uint64(0x4B455953) ^ np.arange(bench.KEYS_C4, dtype=np.uint64)).view(np.int64)
    assert len(np.unique(table)) == bench.KEYS_C4 and np.isin(c4["keys"], table).all()
    assert len(np.unique(c4["keys"])) > 45_000          # 50000 draws from 1e6 keys: almost all distinct
    # a later block continues the same columns
    assert np.array_equal(bench.host_columns("c3", 500, 100)["x"], bench.host_columns("c3", 0, 600)["x"][500:])
