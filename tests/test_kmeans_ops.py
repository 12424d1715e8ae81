"""The reference's one application: the k-means and geometric-mean snippets
(src/main/python/tensorframes_snippets/kmeans_demo.py, geom_mean.py), rebuilt with the DSL node for node
(Shape -> StridedSlice -> Pack -> Tile, Square, Sum/Min over axis 1, MatMul(transpose_b), ExpandDims, ArgMin,
UnsortedSegmentSum, Inv, OnesLike) and run through map_blocks / aggregate / reduce_blocks.

Not-GPU part: the oracle's evaluation of those graphs against plain numpy formulas, and the static shapes the
C ABI reports for the fetches (what TF's shape inference gives the reference).  GPU part: the executors against
the oracle.  The snippets have no expected values in the reference (they print), so parity for these ops is
unpinned by the reference and pinned to the published TF op definitions through numpy."""
import numpy as np
import pytest

import tensorframes_b200 as tfs
from oracle import tf_ops
from tensorframes_b200 import _native, dsl
from tensorframes_b200.frame import TensorFrame
from tensorframes_b200.shape import Shape, Unknown


def distances_graph(points, centers):
    """tf_compute_distances (kmeans_demo.py:13-43)."""
    num_centroids = centers.shape[0]
    num_points = dsl.shape(points)[0]
    c = dsl.constant(centers)
    squares = dsl.reduce_sum(dsl.square(points), [1])
    center_squares = dsl.reduce_sum(dsl.square(c), [1])
    prods = dsl.matmul(points, c, transpose_b=True)
    t1 = dsl.tile(dsl.expand_dims(center_squares, 0), dsl.stack([num_points, 1]))
    t2 = dsl.tile(dsl.expand_dims(squares, 1), dsl.stack([1, num_centroids]))
    return t1 + t2 - 2.0 * prods


def step1_graph(centers, m):
    """run_one_step's map graph (kmeans_demo.py:68-80): nearest centroid, a column of ones, the distance."""
    with dsl.withGraph():
        points = dsl.placeholder(tfs.DoubleType, Unknown, m).named("features")
        num_points = dsl.stack([dsl.shape(points)[0]]).named("num_points")
        distances = distances_graph(points, centers)
        indexes = dsl.argmin(distances, 1).named("indexes")
        min_distances = dsl.reduce_min(distances, [1]).named("min_distances")
        counts = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), num_points).named("count")
        fetches = [indexes, counts, min_distances]
        return dsl.build_graph(*fetches), fetches


def step2_graph(centers, m):
    """run_one_step2's map graph (kmeans_demo.py:124-146): the per-block pre-aggregation."""
    k = centers.shape[0]
    with dsl.withGraph():
        points = dsl.placeholder(tfs.DoubleType, Unknown, m).named("features")
        distances = distances_graph(points, centers)
        indexes = dsl.argmin(distances, 1).named("indexes")
        min_distances = dsl.reduce_min(distances, [1]).named("min_distances")
        num_points = dsl.stack([dsl.shape(points)[0]]).named("num_points")
        counts = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), num_points).named("count")
        block_points = dsl.unsorted_segment_sum(points, indexes, k).named("block_points")
        block_counts = dsl.unsorted_segment_sum(counts, indexes, k).named("block_counts")
        block_distances = dsl.reduce_sum(min_distances).named("block_distances")
        fetches = [dsl.expand_dims(block_points, 0).named("agg_points"), dsl.expand_dims(block_counts, 0).named("agg_counts"),
                   dsl.expand_dims(block_distances, 0).named("agg_distances")]
        return dsl.build_graph(*fetches), fetches


def numpy_step(x, centers):
    d = (centers ** 2).sum(1)[None, :] + (x ** 2).sum(1)[:, None] - 2.0 * (x @ centers.T)
    idx = d.argmin(1)
    return d, idx, d.min(1)


def data(n=4000, m=6, k=5, seed=11):
    rng = np.random.default_rng(seed)
    true = rng.uniform(-4, 4, (k, m))
    x = true[rng.integers(0, k, n)] + rng.normal(0, 0.5, (n, m))
    centers = x[rng.choice(n, k, replace=False)].copy()
    return x, centers


# ---------------------------------------------------------------------------------------------------------------
def test_oracle_evaluates_the_kmeans_graphs_like_numpy():
    x, centers = data(n=257)
    g, fetches = step1_graph(centers, x.shape[1])
    idx, cnt, md = tf_ops.evaluate(g, {"features": x}, [f.name for f in fetches])
    d, ref_idx, ref_md = numpy_step(x, centers)
    assert idx.dtype == np.int64 and np.array_equal(idx, ref_idx)
    assert cnt.dtype == np.int32 and np.array_equal(cnt, np.ones(len(x), dtype=np.int32))
    np.testing.assert_allclose(md, ref_md, rtol=1e-12, atol=1e-12)
    g2, fetches2 = step2_graph(centers, x.shape[1])
    pts, cts, dist = tf_ops.evaluate(g2, {"features": x}, [f.name for f in fetches2])
    k = centers.shape[0]
    assert pts.shape == (1, k, x.shape[1]) and cts.shape == (1, k) and dist.shape == (1,)
    for j in range(k):
        np.testing.assert_allclose(pts[0, j], x[ref_idx == j].sum(0), rtol=1e-12, atol=1e-12)
        assert cts[0, j] == (ref_idx == j).sum()
    np.testing.assert_allclose(dist[0], ref_md.sum(), rtol=1e-12)


def test_static_shapes_of_the_kmeans_fetches():
    x, centers = data(n=16)
    k, m = centers.shape
    g, fetches = step1_graph(centers, m)
    assert [f.shape for f in fetches] == [Shape(Unknown), Shape(Unknown), Shape(Unknown)]     # the DSL's own inference
    with _native.Graph(g) as gr:
        for f, dt in zip(fetches, (tfs.LongType, tfs.IntegerType, tfs.DoubleType)):
            st, sh, is_ph, _ = gr.node_info(f.name)
            assert st is dt and sh == Shape(Unknown) and not is_ph
        ops = dict(gr.nodes())
        assert {"Shape", "StridedSlice", "Pack", "Tile", "ExpandDims", "ArgMin", "MatMul", "Square"} <= set(ops.values())
    g2, fetches2 = step2_graph(centers, m)
    assert [f.shape for f in fetches2] == [Shape(1, k, m), Shape(1, k), Shape(1)]
    with _native.Graph(g2) as gr:
        assert gr.node_info("agg_points")[1] == Shape(1, k, m)
        assert gr.node_info("agg_counts")[:2] == (tfs.IntegerType, Shape(1, k))
        assert gr.node_info("agg_distances")[1] == Shape(1)


def test_unsupported_shape_plumbing_is_refused_by_name():
    with dsl.withGraph():
        p = dsl.placeholder(tfs.IntegerType, Unknown).named("x")
        bad = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), p).named("t")     # multiples from DATA, not shapes
        g = dsl.build_graph(bad)
    with _native.Graph(g) as gr:
        assert gr.node_info("t")[1] == Shape(Unknown)


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_kmeans_step_through_map_blocks_and_aggregate():
    """run_one_step: map_blocks([indexes, count, min_distances]) then groupBy(indexes).aggregate(sums)."""
    x, centers = data()
    k, m = centers.shape
    df = tfs.analyze(TensorFrame.from_columns({"features": x}, num_partitions=3))     # kmeans_demo.py:220 does the same
    g, fetches = step1_graph(centers, m)
    with dsl.withGraph():
        points = dsl.placeholder(tfs.DoubleType, Unknown, m).named("features")
        distances = distances_graph(points, centers)
        indexes = dsl.argmin(distances, 1).named("indexes")
        min_distances = dsl.reduce_min(distances, [1]).named("min_distances")
        counts = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), dsl.stack([dsl.shape(points)[0]])).named("count")
        df2 = tfs.map_blocks([indexes, counts, min_distances], df)
    d, ref_idx, ref_md = numpy_step(x, centers)
    got_idx = np.concatenate([p["indexes"] for p in df2.partitions])
    got_md = np.concatenate([p["min_distances"] for p in df2.partitions])
    got_cnt = np.concatenate([p["count"] for p in df2.partitions])
    assert got_idx.dtype == np.int64 and got_cnt.dtype == np.int32 and np.all(got_cnt == 1)
    srt = np.sort(d, axis=1)
    clear = (srt[:, 1] - srt[:, 0]) > 1e-9 * np.abs(srt[:, 0]).clip(1e-30)     # rows without a near-tie
    assert clear.mean() > 0.99 and np.array_equal(got_idx[clear], ref_idx[clear])
    np.testing.assert_allclose(got_md, ref_md, rtol=1e-11, atol=1e-11)
    # the oracle on the same graph (block = one partition)
    part = df.partitions[1]["features"]
    o_idx, o_cnt, o_md = tf_ops.evaluate(g, {"features": part}, [f.name for f in fetches])
    assert np.array_equal(df2.partitions[1]["count"], o_cnt)
    np.testing.assert_allclose(df2.partitions[1]["min_distances"], o_md, rtol=1e-11, atol=1e-11)
    # the aggregation half of the step
    with dsl.withGraph():
        x_in = tfs.block(df2, "features", tf_name="features_input")
        c_in = tfs.block(df2, "count", tf_name="count_input")
        md_in = tfs.block(df2, "min_distances", tf_name="min_distances_input")
        df3 = tfs.aggregate([dsl.reduce_sum(x_in, [0]).named("features"), dsl.reduce_sum(c_in, [0]).named("count"),
                             dsl.reduce_sum(md_in, [0]).named("min_distances")], df2.groupBy("indexes"))
    rows = df3.collect()
    assert sorted(r.indexes for r in rows) == sorted(set(got_idx.tolist()))
    new_centers = {r.indexes: np.asarray(r.features) / r["count"] for r in rows}
    for j, c in new_centers.items():
        np.testing.assert_allclose(c, x[got_idx == j].mean(0), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(sum(r.min_distances for r in rows), got_md.sum(), rtol=1e-10)


@pytest.mark.gpu
def test_kmeans_preaggregated_step_through_trimmed_map_and_reduce():
    """run_one_step2: map_blocks(trim=True) returns one pre-aggregated row per block, reduce_blocks adds them."""
    x, centers = data(n=5000, m=4, k=7, seed=12)
    k, m = centers.shape
    df = tfs.analyze(TensorFrame.from_columns({"features": x}, num_partitions=4))
    g2, fetches2 = step2_graph(centers, m)
    with dsl.withGraph():
        points = dsl.placeholder(tfs.DoubleType, Unknown, m).named("features")
        distances = distances_graph(points, centers)
        indexes = dsl.argmin(distances, 1).named("indexes")
        min_distances = dsl.reduce_min(distances, [1]).named("min_distances")
        counts = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), dsl.stack([dsl.shape(points)[0]])).named("count")
        agg_points = dsl.expand_dims(dsl.unsorted_segment_sum(points, indexes, k), 0).named("agg_points")
        agg_counts = dsl.expand_dims(dsl.unsorted_segment_sum(counts, indexes, k), 0).named("agg_counts")
        agg_distances = dsl.expand_dims(dsl.reduce_sum(min_distances), 0).named("agg_distances")
        df2 = tfs.map_blocks([agg_points, agg_counts, agg_distances], df, trim=True)
    assert df2.count() == 4 and df2.columns == ["agg_counts", "agg_distances", "agg_points"]
    # each block against the oracle on the same graph
    for part_in, part_out in zip(df.partitions, df2.partitions):
        o_pts, o_cts, o_dist = tf_ops.evaluate(g2, {"features": part_in["features"]}, [f.name for f in fetches2])
        assert np.array_equal(part_out["agg_counts"], o_cts)
        np.testing.assert_allclose(part_out["agg_points"], o_pts, rtol=1e-11, atol=1e-11)
        np.testing.assert_allclose(part_out["agg_distances"], o_dist, rtol=1e-11)
    with dsl.withGraph():
        x_in = dsl.placeholder(tfs.DoubleType, Unknown, k, m).named("agg_points_input")
        c_in = dsl.placeholder(tfs.IntegerType, Unknown, k).named("agg_counts_input")
        d_in = dsl.placeholder(tfs.DoubleType, Unknown).named("agg_distances_input")
        x_, count_, total = tfs.reduce_blocks([dsl.reduce_sum(x_in, [0]).named("agg_points"),
                                               dsl.reduce_sum(c_in, [0]).named("agg_counts"),
                                               dsl.reduce_sum(d_in, [0]).named("agg_distances")], df2)
    d, ref_idx, ref_md = numpy_step(x, centers)
    x_, count_ = np.asarray(x_), np.asarray(count_)
    assert count_.sum() == len(x)
    for j in range(k):
        if np.all(np.sort(d, 1)[:, 1] - np.sort(d, 1)[:, 0] > 1e-9):
            assert count_[j] == (ref_idx == j).sum()
    new_centers = (x_.T / (count_ + 1e-7)).T
    ref_centers = np.stack([x[ref_idx == j].sum(0) / ((ref_idx == j).sum() + 1e-7) for j in range(k)])
    np.testing.assert_allclose(new_centers, ref_centers, rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(total, ref_md.sum(), rtol=1e-10)


@pytest.mark.gpu
def test_geometric_mean_snippet():
    """geom_mean.py:25-46 (it computes a harmonic mean): inv + ones_like, keyed sums, count / invs."""
    rng = np.random.default_rng(13)
    n = 1000
    x = rng.uniform(1, 5, n)
    key = rng.integers(0, 4, n)
    df = TensorFrame.from_columns({"key": key, "x": x}, num_partitions=2)
    with dsl.withGraph():
        xb = tfs.block(df, "x")
        invs = dsl.inv(dsl.cast(xb, tfs.DoubleType)).named("invs")
        df2 = tfs.map_blocks([invs, dsl.ones_like(invs).named("count")], df)
    with dsl.withGraph():
        x_in = tfs.block(df2, "invs", tf_name="invs_input")
        c_in = tfs.block(df2, "count", tf_name="count_input")
        df3 = tfs.aggregate([dsl.reduce_sum(x_in, [0]).named("invs"), dsl.reduce_sum(c_in, [0]).named("count")],
                            df2.select("key", "invs", "count").groupBy("key"))
    with dsl.withGraph():
        invs = tfs.block(df3, "invs")
        count = tfs.block(df3, "count")
        df4 = tfs.map_blocks((dsl.cast(count, tfs.DoubleType) / invs).named("harmonic_mean"), df3).select("key", "harmonic_mean")
    got = {r.key: r.harmonic_mean for r in df4.collect()}
    for kk in range(4):
        sel = x[key == kk]
        np.testing.assert_allclose(got[kk], len(sel) / (1.0 / sel).sum(), rtol=1e-12)


@pytest.mark.gpu
def test_shape_ops_one_by_one_against_the_oracle():
    rng = np.random.default_rng(14)
    a = rng.uniform(-1, 1, (37, 5)).astype(np.float32)
    ids32 = rng.integers(0, 6, 37).astype(np.int32)
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, Unknown, 5).named("a")
        s = dsl.placeholder(tfs.IntegerType, Unknown).named("ids")
        fetches = [
            dsl.tile(p, [2, 3]).named("tiled"),
            dsl.tile(dsl.expand_dims(p, 1), dsl.stack([1, 4, 1])).named("tiled3"),
            dsl.argmax(p, 1).named("amax1"), dsl.argmin(p, 0).named("amin0"), dsl.argmax(p, -1).named("amax_neg"),
            dsl.shape(p).named("shp"), dsl.stack([dsl.shape(p)[1], dsl.shape(p)[0], 7]).named("packed"),
            dsl.unsorted_segment_sum(p, s, 6).named("seg"), dsl.unsorted_segment_sum(s, s, 9).named("seg_int"),
            dsl.zeros_like(p).named("z"), dsl.ones_like(s).named("o"),
            dsl.expand_dims(p, -1).named("ex"),
        ]
        g = dsl.build_graph(*fetches)
    names = [f.name for f in fetches]
    want = tf_ops.evaluate(g, {"a": a, "ids": ids32}, names)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as sess:
        got = sess.run({"a": a, "ids": ids32}, names)
    for nm, w, gt in zip(names, want, got):
        assert gt.dtype == w.dtype and gt.shape == w.shape, nm
        if nm == "seg":
            np.testing.assert_allclose(gt, w, rtol=1e-5, atol=1e-6, err_msg=nm)
        else:
            assert gt.tobytes() == w.tobytes(), nm


@pytest.mark.gpu
def test_segment_ids_out_of_range_raise():
    with dsl.withGraph():
        p = dsl.placeholder(tfs.DoubleType, Unknown, 2).named("a")
        s = dsl.placeholder(tfs.LongType, Unknown).named("ids")
        g = dsl.build_graph(dsl.unsorted_segment_sum(p, s, 3).named("seg"))
    a = np.ones((5, 2))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as sess:
        with pytest.raises(_native.NativeError, match="out of range"):
            sess.run({"a": a, "ids": np.asarray([0, 1, 2, 3, 0], dtype=np.int64)}, ["seg"])     # 4 distinct ids > 3 segments
        with pytest.raises(_native.NativeError, match="out of range"):
            sess.run({"a": a, "ids": np.asarray([0, 1, -1, 1, 0], dtype=np.int64)}, ["seg"])    # in count, out of range
        (ok,) = sess.run({"a": a, "ids": np.asarray([2, 2, 0, 2, 0], dtype=np.int64)}, ["seg"])
        assert ok.tolist() == [[2.0, 2.0], [0.0, 0.0], [3.0, 3.0]]


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_many_distinct_bad_segment_ids_raise():
    """200000 distinct ids against 3 segments: far more keys than the (1024-slot) table of the keyed pass holds.
    The run must fail with the out-of-range error instead of probing a full table for ever."""
    with dsl.withGraph():
        p = dsl.placeholder(tfs.FloatType, Unknown, 2).named("a")
        s = dsl.placeholder(tfs.LongType, Unknown).named("ids")
        g = dsl.build_graph(dsl.unsorted_segment_sum(p, s, 3).named("seg"))
    n = 200_000
    a = np.ones((n, 2), dtype=np.float32)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as sess:
        with pytest.raises(_native.NativeError, match="out of range"):
            sess.run({"a": a, "ids": np.arange(n, dtype=np.int64) + 5}, ["seg"])
        (ok,) = sess.run({"a": a, "ids": (np.arange(n) % 3).astype(np.int64)}, ["seg"])
        assert ok[:, 0].tolist() == [float(len(range(k, n, 3))) for k in range(3)]


@pytest.mark.gpu
def test_float64_matmul_with_transposes():
    rng = np.random.default_rng(15)
    a = rng.uniform(-1, 1, (300, 17))
    b = rng.uniform(-1, 1, (9, 17))
    with dsl.withGraph():
        pa = dsl.placeholder(tfs.DoubleType, Unknown, 17).named("a")
        pb = dsl.placeholder(tfs.DoubleType, 9, 17).named("b")
        g = dsl.build_graph(dsl.matmul(pa, pb, transpose_b=True).named("ab"), dsl.matmul(pb, pa, transpose_b=True).named("ba"))
    with _native.Graph(g) as gr, _native.Session(gr, 0) as sess:
        ab, ba = sess.run({"a": a, "b": b}, ["ab", "ba"])
    scale = np.abs(a) @ np.abs(b).T
    assert np.all(np.abs(ab - a @ b.T) <= 1e-12 * scale) and np.all(np.abs(ba - b @ a.T) <= 1e-12 * scale.T)


@pytest.mark.gpu
def test_kmeans_graphs_on_an_empty_block():
    """Zero rows: every step degenerates to an empty launch, the segment sums are zero rows of the right shape."""
    _, centers = data(n=16)
    k, m = centers.shape
    x = np.zeros((0, m))
    g, fetches = step1_graph(centers, m)
    g2, fetches2 = step2_graph(centers, m)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        idx, cnt, md = s.run({"features": x}, [f.name for f in fetches])
    assert idx.shape == (0,) and idx.dtype == np.int64 and cnt.shape == (0,) and md.shape == (0,)
    with _native.Graph(g2) as gr, _native.Session(gr, 0) as s:
        pts, cts, dist = s.run({"features": x}, [f.name for f in fetches2])
    assert pts.shape == (1, k, m) and not pts.any() and cts.shape == (1, k) and not cts.any() and dist.tolist() == [0.0]


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(40))
def test_random_shape_op_graphs_match_the_oracle(seed):
    """Tile / ExpandDims / ArgMin / ArgMax / UnsortedSegmentSum / Shape arithmetic on random shapes and dtypes, fed
    into each other and into element-wise ops."""
    rng = np.random.default_rng(13000 + seed)
    dts = [("f4", tfs.FloatType), ("f8", tfs.DoubleType), ("i4", tfs.IntegerType), ("i8", tfs.LongType)]
    name, st = dts[seed % 4]
    npdt = np.dtype("<" + name)
    n = int(rng.choice([1, 6, 257, 5001]))
    c = int(rng.choice([1, 3, 8, 33]))
    x = rng.integers(-5, 6, (n, c)).astype(npdt)
    n_seg = int(rng.choice([1, 4, 70]))
    ids = rng.integers(0, n_seg, n).astype([np.int32, np.int64][seed % 2])
    with dsl.withGraph():
        p = dsl.placeholder(st, dsl.Unknown, c).named("x")
        pid = dsl.placeholder([tfs.IntegerType, tfs.LongType][seed % 2], dsl.Unknown).named("ids")
        rows = dsl.shape(p)[0]
        reps = int(rng.integers(1, 4))
        fetches = [
            dsl.tile(p, [reps, int(rng.integers(1, 3))]).named("t0"),
            dsl.tile(dsl.expand_dims(dsl.reduce_sum(p, [0]), 0), dsl.stack([rows, 1])).named("t1"),        # a row vector per row
            (p - dsl.tile(dsl.expand_dims(dsl.reduce_max(p, [1]), 1), dsl.stack([1, c]))).named("t2"),    # x - rowmax
            dsl.argmin(p, 1).named("a1"), dsl.argmax(p, 0).named("a0"),
            dsl.unsorted_segment_sum(p, pid, n_seg).named("seg"),
            dsl.unsorted_segment_sum(dsl.ones_like(dsl.reduce_sum(p, [1])), pid, n_seg).named("cnt"),
            dsl.expand_dims(dsl.reduce_min(p, [0]), int(rng.integers(0, 2))).named("e"),
            dsl.stack([rows, c, 7]).named("shp"),
        ]
        g = dsl.build_graph(*fetches)
    names = [f.name for f in fetches]
    feeds = {"x": x, "ids": ids}
    want = tf_ops.evaluate(g, feeds, names)
    with _native.Graph(g) as gr, _native.Session(gr, 0) as s:
        got = s.run(feeds, names)
    for nm, w, gt in zip(names, want, got):
        assert gt.dtype == w.dtype and gt.shape == w.shape, (seed, nm, gt.dtype, w.dtype, gt.shape, w.shape)
        assert gt.tobytes() == np.ascontiguousarray(w).tobytes(), (seed, nm)
