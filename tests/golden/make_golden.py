"""Writes tests/golden/reference_vectors.json: every known-answer vector the reference's own
tests and README hold for the block path, transcribed with the file:line it comes from, plus
the two GraphDef fixtures of the reference (read from /root/reference, hex-encoded).

Run from the repo root in the build container (where /root/reference exists):
    python tests/golden/make_golden.py
The GPU box has no /root/reference; tests only read the JSON this script commits.

Shorthand in `cite`: T/ = src/test/scala/org/tensorframes/, P/ = src/main/python/tensorframes/.
Graph spec: placeholders [name, dtype, shape], consts [name, dtype, value], ops [name, op, args...].
"""
import json
import os

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors.json")

DTYPES = {
    "int": ("int", int),
    "long": ("long", int),
    "float": ("float", float),
    "double": ("double", float),
}


def u(dt, v):
    """`.u` of CommonOperationsSuite: literals are written as doubles and converted (T/CommonOperationsSuite.scala:24-86)."""
    conv = DTYPES[dt][1]
    if isinstance(v, list):
        return [u(dt, x) for x in v]
    if isinstance(v, str):
        return v
    return conv(v)


cases = []

# ---------------- mapBlocks ----------------
cases += [
    dict(id="map_identity", cite="T/BasicOperationsSuite.scala:19-25", kind="map_blocks",
         columns=[["in", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["in", "double", [-1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[1.0, 1.0], [2.0, 2.0]]),
    dict(id="map_simple_add", cite="T/BasicOperationsSuite.scala:27-34", kind="map_blocks",
         columns=[["a", "double"], ["b", "double"]], rows=[[1.0, 1.1], [2.0, 2.2]],
         graph=dict(placeholders=[["a", "double", [-1]], ["b", "double", [-1]]], ops=[["out", "add", "a", "b"]],
                    fetches=["out"]),
         select=["a", "b", "out"], expected=[[1.0, 1.1, 2.1], [2.0, 2.2, 4.2]]),
    dict(id="map_identity_1dim", cite="T/BasicOperationsSuite.scala:36-43", kind="map_blocks", analyze=True,
         columns=[["in", "double"]], rows=[[[1.0]], [[2.0]]],
         graph=dict(placeholders=[["in", "double", [-1, 1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[[1.0], [1.0]], [[2.0], [2.0]]]),
    dict(id="map_simple_add_1dim", cite="T/BasicOperationsSuite.scala:45-58", kind="map_blocks", analyze=True,
         columns=[["a", "double"], ["b", "double"]], rows=[[[1.0], [1.1]], [[2.0], [2.2]]],
         graph=dict(placeholders=[["a", "double", [-1, 1]], ["b", "double", [-1, 1]]], ops=[["out", "add", "a", "b"]],
                    fetches=["out"]),
         select=["a", "b", "out"], expected=[[[1.0], [1.1], [2.1]], [[2.0], [2.2], [4.2]]]),
    dict(id="map_two_fetches", cite="T/DSLOperationsSuite.scala:29-36", kind="map_blocks",
         columns=[["x", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["x", "double", [-1]]], ops=[["y", "identity", "x"], ["z", "add", "x", "x"]],
                    fetches=["y", "z"]),
         select=["x", "y", "z"], expected=[[1.0, 1.0, 2.0], [2.0, 2.0, 4.0]]),
    dict(id="map_block_extraction_scalar_const", cite="T/DSLOperationsSuite.scala:62-70", kind="map_blocks",
         columns=[["a", "double"]], rows=[[1.0], [2.0]],
         graph=dict(block_placeholders=[["a", "a"]], consts=[["c", "double", 2.0]], ops=[["b", "add", "a", "c"]],
                    fetches=["b"]),
         select=["a", "b"], expected=[[1.0, 3.0], [2.0, 4.0]]),
    dict(id="map_2tensor_1", cite="T/BasicOperationsSuite.scala:212-218", kind="map_blocks", analyze=True,
         columns=[["x", "double"]], rows=[[[[1.0]]]],
         graph=dict(block_placeholders=[["x", "x"]], ops=[["y", "identity", "x"]], fetches=["y"]),
         select=["y"], expected=[[[[1.0]]]]),
    dict(id="map_2tensor_2", cite="T/BasicOperationsSuite.scala:220-226", kind="map_blocks", analyze=True,
         columns=[["x", "double"]], rows=[[[[1.0, 2.0]]]],
         graph=dict(block_placeholders=[["x", "x"]], ops=[["y", "identity", "x"]], fetches=["y"]),
         select=["y"], expected=[[[[1.0, 2.0]]]]),
    dict(id="map_2tensor_3", cite="T/BasicOperationsSuite.scala:228-237", kind="map_blocks", analyze=True,
         columns=[["x", "double"]], rows=[[[[1.0, 2.0], [3.0, 4.0]]]],
         graph=dict(block_placeholders=[["x", "x"]], ops=[["y", "identity", "x"]], fetches=["y"]),
         select=["y"], expected=[[[[1.0, 2.0], [3.0, 4.0]]]]),
    # direct executor call: TF columns come FIRST in the produced rows
    # (performMap is called below mapBlocks' validation in the reference test, with a [1] placeholder on a [?]
    #  column; through the public operator the column is analysed first so that the shapes agree)
    dict(id="perform_map_identity", cite="T/DebugRowOpsSuite.scala:15-25", kind="map_blocks", analyze=True,
         columns=[["x", "double"]], rows=[[1.0]],
         graph=dict(placeholders=[["x", "double", [1]]], ops=[["y", "identity", "x"]], fetches=["y"]),
         select=None, expected=[[1.0, 1.0]]),
    dict(id="perform_map_add", cite="T/DebugRowOpsSuite.scala:27-37", kind="map_blocks", analyze=True,
         columns=[["x", "double"]], rows=[[1.0]],
         graph=dict(placeholders=[["x", "double", [1]]], ops=[["y", "add", "x", "x"]], fetches=["y"]),
         select=None, expected=[[2.0, 1.0]]),
    dict(id="readme_map_blocks_x_plus_3", cite="README.md:56-87; P/core_test.py:39-50", kind="map_blocks",
         columns=[["x", "double"]], rows=[[float(i)] for i in range(10)],
         graph=dict(placeholders=[["x", "double", [-1]]], consts=[["c", "double", 3.0]], ops=[["z", "add", "x", "c"]],
                    fetches=["z"]),
         select=["z", "x"], expected=[[float(i) + 3.0, float(i)] for i in range(10)]),
]

for dt in DTYPES:
    cases += [
        dict(id=f"type_identity_{dt}", cite="T/type_suites.scala:11-18", kind="map_blocks",
             columns=[["in", dt]], rows=u(dt, [[1.0], [20.0]]),
             graph=dict(placeholders=[["in", dt, [-1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
             select=["in", "out"], expected=u(dt, [[1.0, 1.0], [20.0, 20.0]])),
        dict(id=f"type_identity_1dim_{dt}", cite="T/type_suites.scala:20-27", kind="map_blocks", analyze=True,
             columns=[["in", dt]], rows=u(dt, [[[1.0]], [[20.0]]]),
             graph=dict(placeholders=[["in", dt, [-1, 1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
             select=["in", "out"], expected=u(dt, [[[1.0], [1.0]], [[20.0], [20.0]]])),
        dict(id=f"type_simple_add_{dt}", cite="T/type_suites.scala:74-81", kind="map_blocks",
             columns=[["a", dt], ["b", dt]], rows=u(dt, [[10.0, 11.0], [20.0, 22.0]]),
             graph=dict(placeholders=[["a", dt, [-1]], ["b", dt, [-1]]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
             select=["a", "b", "out"], expected=u(dt, [[10.0, 11.0, 21.0], [20.0, 22.0, 42.0]])),
        dict(id=f"type_simple_add_1dim_{dt}", cite="T/type_suites.scala:83-96", kind="map_blocks", analyze=True,
             columns=[["a", dt], ["b", dt]], rows=u(dt, [[[10.0], [11.0]], [[20.0], [22.0]]]),
             graph=dict(placeholders=[["a", dt, [-1, 1]], ["b", dt, [-1, 1]]], ops=[["out", "add", "a", "b"]],
                        fetches=["out"]),
             select=["a", "b", "out"], expected=u(dt, [[[10.0], [11.0], [21.0]], [[20.0], [22.0], [42.0]]])),
        dict(id=f"type_reduce_block_sum_{dt}", cite="T/type_suites.scala:158-164", kind="reduce_blocks",
             columns=[["x", dt]], rows=u(dt, [[1.0], [2.0]]),
             graph=dict(placeholders=[["x_input", dt, [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
             expected=u(dt, [3.0])),
        dict(id=f"type_reduce_block_sum_extra_column_{dt}", cite="T/type_suites.scala:166-175", kind="reduce_blocks",
             columns=[["key2", "string"], ["x", dt]], rows=[["1", u(dt, 10.0)], ["2", u(dt, 11.0)], ["3", u(dt, 20.0)]],
             graph=dict(placeholders=[["x_input", dt, [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
             expected=u(dt, [41.0])),
        dict(id=f"type_aggregate_{dt}", cite="T/type_suites.scala:177-186", kind="aggregate", keys=["key"],
             columns=[["key", "string"], ["x", dt]], rows=[["a", u(dt, 10.0)], ["a", u(dt, 11.0)], ["b", u(dt, 20.0)]],
             graph=dict(placeholders=[["x_input", dt, [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
             select=["key", "x"], expected=[["a", u(dt, 21.0)], ["b", u(dt, 20.0)]]),
    ]

# ---------------- trimmed maps ----------------
cases += [
    dict(id="trim_less_rows", cite="T/TrimmingOperationsSuite.scala:17-23", kind="map_blocks_trimmed",
         columns=[["in", "double"]], rows=[[1.0], [2.0]],
         graph=dict(consts=[["out", "double", [1.0]]], fetches=["out"]),
         select=None, expected_columns=["out"], expected=[[1.0]]),
    dict(id="trim_more_rows", cite="T/TrimmingOperationsSuite.scala:25-31", kind="map_blocks_trimmed",
         columns=[["in", "double"]], rows=[[3.0]],
         graph=dict(consts=[["out", "double", [1.0, 2.0]]], fetches=["out"]),
         select=None, expected_columns=["out"], expected=[[1.0], [2.0]]),
    dict(id="trim_as_many_rows", cite="T/TrimmingOperationsSuite.scala:33-39", kind="map_blocks_trimmed",
         columns=[["in", "double"]], rows=[[3.0], [4.0]],
         graph=dict(consts=[["out", "double", [1.0, 2.0]]], fetches=["out"]),
         select=None, expected_columns=["out"], expected=[[1.0], [2.0]]),
    dict(id="trim_higher_dim", cite="T/TrimmingOperationsSuite.scala:41-47", kind="map_blocks_trimmed", analyze=True,
         columns=[["in", "double"]], rows=[[[1.0]], [[2.0]]],
         graph=dict(consts=[["out", "double", [[1.0]]]], fetches=["out"]),
         select=None, expected_columns=["out"], expected=[[[1.0]]]),
    dict(id="py_trim_const_int", cite="P/core_test.py:105-116", kind="map_blocks_trimmed",
         columns=[["x", "double"]], rows=[[0.0], [1.0], [2.0]],
         graph=dict(consts=[["z", "int", [2]]], fetches=["z"]),
         select=None, expected_columns=["z"], expected=[[2]]),
]

# ---------------- reduceBlocks ----------------
cases += [
    dict(id="reduce_block_sum_double", cite="T/BasicOperationsSuite.scala:170-176", kind="reduce_blocks",
         columns=[["x", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["x_input", "double", [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
         expected=[3.0]),
    dict(id="reduce_block_sum_extra_column", cite="T/BasicOperationsSuite.scala:178-187", kind="reduce_blocks",
         columns=[["key2", "string"], ["x", "double"]], rows=[["1", 1.0], ["2", 1.1], ["3", 2.0]],
         graph=dict(placeholders=[["x_input", "double", [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
         expected=[4.1]),
    dict(id="reduce_block_two_partitions", cite="T/BasicOperationsSuite.scala:189-198", kind="reduce_blocks",
         partitions=2, analyze=True,
         columns=[["x", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["x_input", "double", [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
         expected=[3.0]),
    dict(id="py_reduce_blocks_all_axes", cite="P/core_test.py:93-103", kind="reduce_blocks",
         columns=[["x", "double"]], rows=[[float(i)] for i in range(5)],
         graph=dict(placeholders=[["x_input", "double", [-1]]], ops=[["x", "reduce_sum", "x_input", None]], fetches=["x"]),
         expected=[10.0]),
    dict(id="readme_reduce_blocks_sum_min", cite="README.md:92-124 (documentation, the only stated answer for Min)",
         kind="reduce_blocks", analyze=True,
         columns=[["y", "double"], ["z", "double"]], rows=[[[float(i), float(-i)], [float(i), float(-i)]] for i in range(10)],
         graph=dict(block_placeholders=[["y", "y_input"], ["z", "z_input"]],
                    ops=[["y", "reduce_sum", "y_input", [0]], ["z", "reduce_min", "z_input", [0]]], fetches=["y", "z"]),
         expected=[[45.0, -45.0], [0.0, -9.0]]),
]

# ---------------- aggregate ----------------
cases += [
    dict(id="aggregate_int_key", cite="T/BasicOperationsSuite.scala:200-210", kind="aggregate", keys=["key"],
         columns=[["key", "int"], ["x", "double"]], rows=[[1, 1.0], [1, 1.1], [2, 2.0]],
         graph=dict(placeholders=[["x_input", "double", [-1]]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
         select=["key", "x"], expected=[[1, 2.1], [2, 2.0]]),
    dict(id="py_groupby", cite="P/core_test.py:118-127", kind="aggregate", keys=["key"],
         columns=[["x", "double"], ["key", "string"]], rows=[[float(x), str(x % 2)] for x in range(4)],
         graph=dict(block_placeholders=[["x", "x_input"]], ops=[["x", "reduce_sum", "x_input", [0]]], fetches=["x"]),
         select=["key", "x"], expected=[["0", 2.0], ["1", 4.0]]),
]

# ---------------- mapRows / reduceRows ----------------
cases += [
    dict(id="rows_identity", cite="T/BasicOperationsSuite.scala:78-84", kind="map_rows",
         columns=[["in", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["in", "double", []]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[1.0, 1.0], [2.0, 2.0]]),
    dict(id="rows_simple_add", cite="T/BasicOperationsSuite.scala:86-98", kind="map_rows",
         columns=[["a", "double"], ["b", "double"]], rows=[[1.0, 1.1], [2.0, 2.2]],
         graph=dict(placeholders=[["a", "double", []], ["b", "double", []]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
         select=["a", "b", "out"], expected=[[1.0, 1.1, 2.1], [2.0, 2.2, 4.2]]),
    dict(id="rows_identity_known_dim", cite="T/BasicOperationsSuite.scala:100-112", kind="map_rows", analyze=True,
         columns=[["in", "double"]], rows=[[[1.0]], [[2.0]]],
         graph=dict(placeholders=[["in", "double", [1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[[1.0], [1.0]], [[2.0], [2.0]]]),
    dict(id="rows_identity_unknown_dim", cite="T/BasicOperationsSuite.scala:114-126", kind="map_rows", analyze=True,
         columns=[["in", "double"]], rows=[[[1.0]], [[2.0]]],
         graph=dict(placeholders=[["in", "double", [-1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[[1.0], [1.0]], [[2.0], [2.0]]]),
    dict(id="rows_identity_variable_sizes", cite="T/BasicOperationsSuite.scala:128-140", kind="map_rows", analyze=True,
         columns=[["in", "double"]], rows=[[[1.0]], [[2.0, 2.1]]],
         graph=dict(placeholders=[["in", "double", [-1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
         select=["in", "out"], expected=[[[1.0], [1.0]], [[2.0, 2.1], [2.0, 2.1]]]),
    dict(id="rows_add_1dim", cite="T/BasicOperationsSuite.scala:142-155", kind="map_rows", analyze=True,
         columns=[["a", "double"], ["b", "double"]], rows=[[[1.0], [1.1]], [[2.0], [2.2]]],
         graph=dict(placeholders=[["a", "double", [1]], ["b", "double", [1]]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
         select=["a", "b", "out"], expected=[[[1.0], [1.1], [2.1]], [[2.0], [2.2], [4.2]]]),
    dict(id="rows_add_unknown_rows", cite="T/BasicOperationsSuite.scala:157-172", kind="map_rows", analyze=True,
         columns=[["a", "double"], ["b", "double"]], rows=[[[1.0, 1.0], [1.1, 1.1]], [[2.0], [2.2]]],
         graph=dict(placeholders=[["a", "double", [-1]], ["b", "double", [-1]]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
         select=["a", "b", "out"], expected=[[[1.0, 1.0], [1.1, 1.1], [2.1, 2.1]], [[2.0], [2.2], [4.2]]]),
    dict(id="rows_const_reduce", cite="T/DSLOperationsSuite.scala:12-19", kind="map_rows",
         columns=[["a", "int"]], rows=[[1]],
         graph=dict(consts=[["x", "double", [1.0, 1.0]]], ops=[["out", "reduce_sum", "x", [0]]], fetches=["out"], keep_consts=["x"]),
         select=["a", "out"], expected=[[1, 2.0]]),
    dict(id="rows_constant", cite="T/DSLOperationsSuite.scala:21-26", kind="map_rows",
         columns=[["a", "int"]], rows=[[1]],
         graph=dict(consts=[["x", "double", 1.0]], fetches=["x"]),
         select=["a", "x"], expected=[[1, 1.0]]),
    dict(id="rows_row_extraction", cite="T/DSLOperationsSuite.scala:52-60", kind="map_rows",
         columns=[["a", "double"]], rows=[[1.0], [2.0]],
         graph=dict(row_placeholders=[["a", "a"]], consts=[["c", "double", 2.0]], ops=[["b", "add", "a", "c"]], fetches=["b"]),
         select=["a", "b"], expected=[[1.0, 3.0], [2.0, 4.0]]),
    dict(id="rows_2tensor_constant_output", cite="T/BasicOperationsSuite.scala:239-246", kind="map_rows", analyze=True,
         columns=[["x", "int"]], rows=[[1]],
         graph=dict(consts=[["y", "double", [[1.0]]]], fetches=["y"]),
         select=["y"], expected=[[[[1.0]]]]),
    dict(id="py_map_rows_feed_dict", cite="P/core_test.py:64-75", kind="map_rows", feed_dict={"x": "y"},
         columns=[["y", "double"]], rows=[[float(i)] for i in range(5)],
         graph=dict(placeholders=[["x", "double", []]], consts=[["c", "double", 3.0]], ops=[["z", "add", "x", "c"]], fetches=["z"]),
         select=["z", "y"], expected=[[float(i) + 3.0, float(i)] for i in range(5)]),
    dict(id="reduce_rows_sum_double", cite="T/BasicOperationsSuite.scala:60-67", kind="reduce_rows",
         columns=[["x", "double"]], rows=[[1.0], [2.0]],
         graph=dict(placeholders=[["x_1", "double", []], ["x_2", "double", []]], ops=[["x", "add", "x_1", "x_2"]], fetches=["x"]),
         expected=[3.0]),
    dict(id="reduce_rows_sum_int", cite="T/BasicOperationsSuite.scala:69-76", kind="reduce_rows",
         columns=[["x", "int"]], rows=[[1], [2], [3], [4]],
         graph=dict(placeholders=[["x_1", "int", []], ["x_2", "int", []]], ops=[["x", "add", "x_1", "x_2"]], fetches=["x"]),
         expected=[10]),
    dict(id="py_reduce_rows", cite="P/core_test.py:78-90", kind="reduce_rows",
         columns=[["x", "double"]], rows=[[float(i)] for i in range(5)],
         graph=dict(placeholders=[["x_1", "double", []], ["x_2", "double", []]], ops=[["x", "add", "x_1", "x_2"]], fetches=["x"]),
         expected=[10.0]),
]
for dt in DTYPES:
    cases += [
        dict(id=f"type_rows_identity_known_dim_{dt}", cite="T/type_suites.scala:29-40", kind="map_rows", analyze=True,
             columns=[["in", dt]], rows=u(dt, [[[1.0]], [[20.0]]]),
             graph=dict(placeholders=[["in", dt, [1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
             select=["in", "out"], expected=u(dt, [[[1.0], [1.0]], [[20.0], [20.0]]])),
        dict(id=f"type_rows_variable_sizes_{dt}", cite="T/type_suites.scala:55-66", kind="map_rows", analyze=True,
             columns=[["in", dt]], rows=[[u(dt, [1.0])], [u(dt, [20.0, 21.0])]],
             graph=dict(placeholders=[["in", dt, [-1]]], ops=[["out", "identity", "in"]], fetches=["out"]),
             select=["in", "out"], expected=[[u(dt, [1.0]), u(dt, [1.0])], [u(dt, [20.0, 21.0]), u(dt, [20.0, 21.0])]]),
        dict(id=f"type_rows_add_one_row_{dt}", cite="T/type_suites.scala:114-124", kind="map_rows",
             columns=[["a", dt], ["b", dt]], rows=u(dt, [[10.0, 11.0], [20.0, 22.0]]),
             graph=dict(placeholders=[["a", dt, []], ["b", dt, []]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
             select=["a", "b", "out"], expected=u(dt, [[10.0, 11.0, 21.0], [20.0, 22.0, 42.0]])),
        dict(id=f"type_rows_add_unknown_rows_{dt}", cite="T/type_suites.scala:142-156", kind="map_rows", analyze=True,
             columns=[["a", dt], ["b", dt]], rows=[[u(dt, [10.0, 10.0]), u(dt, [11.0, 11.0])], [u(dt, [20.0]), u(dt, [22.0])]],
             graph=dict(placeholders=[["a", dt, [-1]], ["b", dt, [-1]]], ops=[["out", "add", "a", "b"]], fetches=["out"]),
             select=["a", "b", "out"],
             expected=[[u(dt, [10.0, 10.0]), u(dt, [11.0, 11.0]), u(dt, [21.0, 21.0])], [u(dt, [20.0]), u(dt, [22.0]), u(dt, [42.0])]]),
        dict(id=f"type_reduce_rows_sum_{dt}", cite="T/type_suites.scala:98-105", kind="reduce_rows",
             columns=[["x", dt]], rows=u(dt, [[1.0], [2.0]]),
             graph=dict(placeholders=[["x_1", dt, []], ["x_2", dt, []]], ops=[["x", "add", "x_1", "x_2"]], fetches=["x"]),
             expected=u(dt, [3.0])),
    ]

# ---------------- analyze / schema ----------------
analyze_cases = [
    dict(id="explain_double", cite="T/ExtraOperationsSuite.scala:15-23", analyze=False,
         columns=[["a", "double"]], rows=[[0.0]], expected={"a": ["double", [-1]]}),
    dict(id="explain_int", cite="T/ExtraOperationsSuite.scala:25-33", analyze=False,
         columns=[["a", "int"]], rows=[[0]], expected={"a": ["int", [-1]]}),
    dict(id="explain_arrays", cite="T/ExtraOperationsSuite.scala:35-49", analyze=False,
         columns=[["a", "double"], ["b", "double"], ["c", "double"]], rows=[[0.0, [1.0], [[1.0]]]],
         expected={"a": ["double", [-1]], "b": ["double", [-1, -1]], "c": ["double", [-1, -1, -1]]}),
    dict(id="analyze_simple", cite="T/ExtraOperationsSuite.scala:51-60", analyze=True,
         columns=[["a", "double"]], rows=[[0.0]], expected={"a": ["double", [1]]}),
    dict(id="analyze_uneven_partitions", cite="T/ExtraOperationsSuite.scala:62-71", analyze=True, repartition=3,
         columns=[["a", "double"]], rows=[[0.0]] * 10, expected={"a": ["double", [-1]]}),
    dict(id="analyze_variable_sizes", cite="T/ExtraOperationsSuite.scala:73-84", analyze=True,
         columns=[["a", "double"], ["b", "double"]], rows=[[0.0, [0.0]], [1.0, [1.0, 1.0]]],
         expected={"b": ["double", [2, -1]]}),
    dict(id="analyze_second_order", cite="T/ExtraOperationsSuite.scala:86-98", analyze=True,
         columns=[["a", "double"], ["b", "double"]], rows=[[0.0, [0.0, 0.0]], [1.0, [1.0, 1.0]], [2.0, [2.0, 2.0]]],
         expected={"b": ["double", [3, 2]]}),
]

# ---------------- DSL -> NodeDef (what python TF prints for the same program) ----------------
dsl_cases = [
    dict(id="dsl_constant", cite="T/dsl/BasicSuite.scala:16-19", program=[["constant", None, 3]],
         expected=[dict(name="Const", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[3]))]),
    dict(id="dsl_named_constant", cite="T/dsl/BasicSuite.scala:21-24", program=[["constant", "x", 3]],
         expected=[dict(name="x", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[3]))]),
    dict(id="dsl_two_constants", cite="T/dsl/BasicSuite.scala:26-34", program=[["constant", None, 3], ["constant", None, 3]],
         expected=[dict(name="Const", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[3])),
                   dict(name="Const_1", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[3]))]),
    dict(id="dsl_fill", cite="T/dsl/BasicSuite.scala:36-38", program=[["fill", None, [2], 3]],
         expected=[dict(name="Fill", op="Fill", inputs=["Fill/dims", "Fill/value"], T="DT_INT32"),
                   dict(name="Fill/dims", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[1], values=[2])),
                   dict(name="Fill/value", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[3]))]),
    dict(id="dsl_add", cite="T/dsl/BasicOpsSuite.scala:12-22",
         program=[["constant", "x", 1], ["constant", "y", 2], ["add", "z", "x", "y"]],
         expected=[dict(name="x", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[1])),
                   dict(name="y", op="Const", inputs=[], dtype="DT_INT32", value=dict(dtype="DT_INT32", shape=[], values=[2])),
                   dict(name="z", op="Add", inputs=["x", "y"], T="DT_INT32")]),
]

# ---------------- GraphDef fixtures of the reference ----------------
fixtures = {}
for fname, cite, nodes in [
    ("graph.pb", "src/test/resources/graph.pb (T/TFInitializationSuite.scala:23-27)",
     [dict(name="matrix1", op="Const", dtype="DT_FLOAT", shape=[1, 2], values=[3.0, 3.0]),
      dict(name="x", op="Placeholder", dtype="DT_FLOAT", shape=[2])]),
    ("graph2.pb", "src/test/resources/graph2.pb (T/TFInitializationSuite.scala:23-27)",
     [dict(name="z_1", op="Placeholder", dtype="DT_FLOAT", shape=[2, 2]),
      dict(name="z_2", op="Placeholder", dtype="DT_FLOAT", shape=[2, 2]),
      dict(name="out", op="Add", dtype="DT_FLOAT", shape=[2, 2], inputs=["z_1", "z_2"])]),
]:
    with open(os.path.join(REF, "src/test/resources", fname), "rb") as f:
        fixtures[fname] = dict(cite=cite, hex=f.read().hex(), nodes=nodes)

# DenseTensor endianness (T/DenseTensorSuite.scala:8-18)
dense = [dict(id="dense_double", cite="T/DenseTensorSuite.scala:9-13", value=1.0, dtype="double", field="double_val"),
         dict(id="dense_int", cite="T/DenseTensorSuite.scala:15-18", value=1, dtype="int", field="int_val")]

with open(OUT, "w") as f:
    json.dump(dict(cases=cases, analyze=analyze_cases, dsl=dsl_cases, fixtures=fixtures, dense=dense), f, indent=1)
print(f"{len(cases)} executor cases, {len(analyze_cases)} analyze cases, {len(dsl_cases)} dsl cases -> {OUT}")
