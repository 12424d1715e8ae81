"""K-means on a TensorFrame, following the reference's snippet step for step
(src/main/python/tensorframes_snippets/kmeans_demo.py: tf_compute_distances, run_one_step2, kmeanstf).

Each iteration is one trimmed map_blocks (distances to the centroids, nearest centroid, per-block segment sums of
the points -- all on the GPU, one pre-aggregated row per partition) and one reduce_blocks over those rows.

    python examples/kmeans_demo.py [n_points] [n_features] [k] [n_partitions] [--host]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # run from a checkout

import tensorframes_b200 as tfs
from tensorframes_b200 import dsl
from tensorframes_b200.frame import TensorFrame
from tensorframes_b200.shape import Unknown


def compute_distances(points, centers):
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


def run_one_step2(df, centers):
    """run_one_step2 (kmeans_demo.py:101-169): most of the aggregation happens inside the block graph."""
    k, m = centers.shape
    with dsl.withGraph():
        points = dsl.placeholder(tfs.DoubleType, Unknown, m).named("features")
        distances = compute_distances(points, centers)
        indexes = dsl.argmin(distances, 1).named("indexes")
        min_distances = dsl.reduce_min(distances, [1]).named("min_distances")
        counts = dsl.tile(dsl.constant(np.asarray([1], dtype=np.int32)), dsl.stack([dsl.shape(points)[0]])).named("count")
        agg_points = dsl.expand_dims(dsl.unsorted_segment_sum(points, indexes, k), 0).named("agg_points")
        agg_counts = dsl.expand_dims(dsl.unsorted_segment_sum(counts, indexes, k), 0).named("agg_counts")
        agg_distances = dsl.expand_dims(dsl.reduce_sum(min_distances), 0).named("agg_distances")
        df2 = tfs.map_blocks([agg_points, agg_counts, agg_distances], df, trim=True)
    with dsl.withGraph():
        x_in = dsl.placeholder(tfs.DoubleType, Unknown, k, m).named("agg_points_input")
        c_in = dsl.placeholder(tfs.IntegerType, Unknown, k).named("agg_counts_input")
        d_in = dsl.placeholder(tfs.DoubleType, Unknown).named("agg_distances_input")
        x_, count_, total = tfs.reduce_blocks([dsl.reduce_sum(x_in, [0]).named("agg_points"),
                                               dsl.reduce_sum(c_in, [0]).named("agg_counts"),
                                               dsl.reduce_sum(d_in, [0]).named("agg_distances")], df2)
    x_, count_ = np.asarray(x_), np.asarray(count_)
    return (x_.T / (count_ + 1e-7)).T, float(total)


def kmeanstf(df, init_centers, num_iters=5):
    """kmeanstf (kmeans_demo.py:172-196)."""
    c, d, ds = init_centers, np.inf, []
    for i in range(num_iters):
        t0 = time.perf_counter()
        c1, d1 = run_one_step2(df, c)
        print(f"Step = {i}, overall distance = {d1:.6e}  ({(time.perf_counter() - t0) * 1e3:.1f} ms)")
        c = c1
        if d == d1:
            break
        d = d1
        ds.append(d1)
    return c, ds


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    n = int(args[0]) if len(args) > 0 else 2_000_000
    m = int(args[1]) if len(args) > 1 else 16
    k = int(args[2]) if len(args) > 2 else 10
    parts = int(args[3]) if len(args) > 3 else 4
    rng = np.random.default_rng(0)
    true = rng.uniform(-10, 10, (k, m))
    x = true[rng.integers(0, k, n)] + rng.normal(0, 1.0, (n, m))
    df = tfs.analyze(TensorFrame.from_columns({"features": x}, num_partitions=parts))
    if "--host" not in sys.argv:
        df = df.to_device(0)          # the points stay in HBM across iterations; only the [k, m] sums come back
    init = x[rng.choice(n, k, replace=False)].copy()
    centers, ds = kmeanstf(df, init, num_iters=8)
    assert all(b <= a * (1 + 1e-12) for a, b in zip(ds, ds[1:])), "k-means must not increase the total distance"
    # one numpy step from the same start, for comparison
    d = (init ** 2).sum(1)[None, :] + (x ** 2).sum(1)[:, None] - 2.0 * (x @ init.T)
    print(f"numpy first-step distance = {d.min(1).sum():.6e}; first step here = {ds[0]:.6e}")
    print("final centers (first 3):")
    print(np.round(centers[:3], 3))


if __name__ == "__main__":
    main()
