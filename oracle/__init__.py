"""CPU oracle for the TensorFrames block path.  TEST INFRASTRUCTURE ONLY.

Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import or execute it, and
only as the checker or the timed CPU baseline.  The product (tensorframes_b200/)
never imports this package and has no CPU execution path.

What it restates: the reference's per-block algorithm
(/root/reference/src/main/scala/org/tensorframes/impl/DebugRowOps.scala,
TFDataOps.scala, DataOps.scala, datatypes.scala) on top of a numpy evaluator of the
TensorFlow 1.1 op definitions the reference delegates to.

Parity status.  The arithmetic lives in a third-party dependency that is not in
/root/reference: org.tensorflow:tensorflow:1.1.0-rc1 (project/Dependencies.scala:17,
project/Build.scala:54), and neither a JVM, Spark nor TensorFlow exist in this
environment, so the reference itself cannot be run.  The oracle is pinned against
every known-answer vector the reference's own tests and README hold for this path
(tests/golden/reference_vectors.json, made by tests/golden/make_golden.py):
Identity, Add, Const, Sum, the trimmed maps, reduce_blocks, aggregate, analyze, and
the two GraphDef fixtures.  The graphs those cases run on are written by oracle/graph_writer.py
(the reference's node layout restated, no product code); the product's DSL is checked against it.  Ops with no reference test -- Sub, Mul, Div, Neg, Cast,
Minimum, Maximum, Relu, reduce_min (README only), reduce_max, MatMul, and the k-means /
geometric-mean snippet vocabulary (Square, Inv, OnesLike, ZerosLike, Shape, StridedSlice,
Pack, ExpandDims, Tile, ArgMin, ArgMax, UnsortedSegmentSum; the snippets print, they hold
no expected values) -- are PARITY UNPINNED: they follow the published TF op definitions
(IEEE-754 round-to-nearest arithmetic, C++ static_cast casts, truncating integer division,
first-extreme-wins arg reductions, row-order segment sums).
"""
