// GraphDef model for the op subset, plus dtype/shape inference.
//
// This is the library-side stand-in for what the reference gets from TF-Java's
// Graph.importGraphDef + Operation.output(i).dataType()/shape()
// (reference: src/main/scala/org/tensorframes/impl/TensorFlowOps.scala:87-150).
// The wire format parsed here is the vendored TF 1.x schema
// (reference: src/main/protobuf/tensorflow/core/framework/{graph,attr_value,
// tensor,tensor_shape,types}.proto).
#pragma once
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"

namespace tfs {

// A fully materialised constant (TensorProto expanded to dense row-major bytes).
struct TensorValue {
  int dtype = 0;
  Shape shape;
  std::vector<uint8_t> bytes;
  int64_t numel() const { return shape.numel(); }
};

struct Attr {
  enum Kind { NONE, S, I, F, B, TYPE, SHAPE, TENSOR, LIST } kind = NONE;
  std::string s;
  int64_t i = 0;
  float f = 0.f;
  bool b = false;
  int type = 0;
  Shape shape;
  TensorValue tensor;
  std::vector<int64_t> list_i;
};

enum class OpClass {
  Placeholder,
  Const,
  Unary,    // Identity, Neg, Relu, Square, Abs, Sqrt, Reciprocal/Inv, Cast
  Binary,   // Add, Sub, Mul, Div, RealDiv, Minimum, Maximum, BiasAdd (broadcasting)
  Reduce,   // Sum, Min, Max with Const reduction_indices
  MatMul,
  Fill,
  // shape plumbing of the k-means / geometric-mean snippets (tensorframes_snippets/kmeans_demo.py:26-43, 76-79, 136-144)
  ShapeOf,       // Shape: int vector known once the input shape is
  StridedSlice,  // on such int vectors only (tf.shape(x)[0])
  Pack,          // stack of int scalars -> int vector (tf.stack([n, 1]))
  ExpandDims,
  Tile,
  ArgMinMax,     // ArgMin / ArgMax along one axis
  SegmentSum,    // UnsortedSegmentSum(data, segment_ids, num_segments)
};

// Element-wise opcodes; shared with the device interpreter (program.h mirrors these values).
enum EltOp : int {
  E_IDENTITY = 0, E_NEG, E_RELU, E_SQUARE, E_ABS, E_SQRT, E_RECIP, E_CAST,
  E_ADD, E_SUB, E_MUL, E_DIV, E_MIN, E_MAX,
  E_ONES_LIKE, E_ZEROS_LIKE,   // unary by arity only: the operand gives the shape, its data is never read
};
enum RedOp : int { R_SUM = 0, R_MIN = 1, R_MAX = 2 };

struct Node {
  std::string name, op;
  std::vector<std::string> raw_inputs;
  std::vector<int> in;             // indices of data inputs (output 0 of each)
  std::map<std::string, Attr> attrs;

  OpClass cls = OpClass::Const;
  int dtype = 0;                   // dtype of output 0
  Shape sshape;                    // statically inferred shape (may hold -1 / unknown rank)
  // parsed parameters
  int elt_op = 0;                  // Unary/Binary
  int red_op = 0;                  // Reduce
  std::vector<int> axes_raw;       // Reduce: reduction_indices as written (may be negative)
  bool keep_dims = false;          // Reduce
  bool transpose_a = false, transpose_b = false;  // MatMul
  TensorValue value;               // Const
  int axis = 0;                    // ExpandDims dim / ArgMinMax dimension (as written, may be negative)
  bool arg_min = true;             // ArgMinMax
  int ss_begin_mask = 0, ss_end_mask = 0, ss_shrink_mask = 0;   // StridedSlice
  // Small integer tensors whose VALUE follows from shapes and constants (Shape, StridedSlice, Pack, int Consts):
  // statically known entries, -1 where the value depends on a dimension not known at import.
  bool has_ival = false;
  std::vector<int64_t> ival;
};

struct Graph {
  std::vector<Node> nodes;
  std::unordered_map<std::string, int> by_name;
  // Looks a node up by name, accepting a trailing ":0"; throws on ":k" with k != 0.
  int find(const std::string& name) const;
};

// Parses + validates + statically analyses a binary GraphDef. Throws tfs::Error.
Graph import_graphdef(const uint8_t* data, size_t len);

// Integer value of an input as far as it is known (nullptr: not an integer-valued node).
using IntVals = std::vector<const std::vector<int64_t>*>;

// Shape of node `n`'s output given the shapes of its inputs (dims may be -1) and, for the ops whose output
// shape depends on an input VALUE (Fill dims, Tile multiples, UnsortedSegmentSum num_segments), those values.
// Used with static shapes at import and with concrete shapes at run time.
Shape infer_shape(const Graph& g, const Node& n, const std::vector<Shape>& in_shapes, const IntVals& in_vals);

// Integer value of node `n` (Shape / StridedSlice / Pack / small int Const) from its inputs' shapes and values;
// false when `n` is not such a node. Entries are -1 where unknown.
bool infer_int_value(const Graph& g, const Node& n, const std::vector<Shape>& in_shapes, const IntVals& in_vals,
                     std::vector<int64_t>& out);

// Normalised, sorted, de-duplicated reduction axes for an input of `rank`.
std::vector<int> normalize_axes(const Node& n, int rank);

}  // namespace tfs
