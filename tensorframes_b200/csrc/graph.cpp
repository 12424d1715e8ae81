// GraphDef import: protobuf wire decoding, node validation, static shape inference.
// See graph.h for the reference call sites this replaces.
#include "graph.h"

#include <algorithm>
#include <set>

namespace tfs {
namespace {

// ---------------------------------------------------------------------------
// Minimal protobuf wire reader (varint / fixed32 / fixed64 / length-delimited).
// ---------------------------------------------------------------------------
struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  Reader(const uint8_t* b, size_t n) : p(b), end(b + n) {}
  bool done() const { return p >= end; }
  uint64_t varint() {
    uint64_t v = 0;
    int shift = 0;
    while (true) {
      TFS_CHECK(p < end, TFS_ERR_INVALID, "GraphDef: truncated varint");
      uint8_t b = *p++;
      v |= uint64_t(b & 0x7f) << shift;
      if (!(b & 0x80)) break;
      shift += 7;
      TFS_CHECK(shift < 70, TFS_ERR_INVALID, "GraphDef: varint too long");
    }
    return v;
  }
  uint32_t fixed32() {
    TFS_CHECK(end - p >= 4, TFS_ERR_INVALID, "GraphDef: truncated fixed32");
    uint32_t v; memcpy(&v, p, 4); p += 4; return v;
  }
  uint64_t fixed64() {
    TFS_CHECK(end - p >= 8, TFS_ERR_INVALID, "GraphDef: truncated fixed64");
    uint64_t v; memcpy(&v, p, 8); p += 8; return v;
  }
  Reader sub() {
    uint64_t n = varint();
    TFS_CHECK(uint64_t(end - p) >= n, TFS_ERR_INVALID, "GraphDef: truncated length-delimited field");
    Reader r(p, size_t(n));
    p += n;
    return r;
  }
  std::string str() {
    Reader r = sub();
    return std::string(reinterpret_cast<const char*>(r.p), size_t(r.end - r.p));
  }
  // returns false at end of message
  bool tag(int& field, int& wt) {
    if (done()) return false;
    uint64_t t = varint();
    field = int(t >> 3);
    wt = int(t & 7);
    return true;
  }
  void skip(int wt) {
    switch (wt) {
      case 0: varint(); break;
      case 1: fixed64(); break;
      case 2: sub(); break;
      case 5: fixed32(); break;
      default: fail(TFS_ERR_INVALID, "GraphDef: unsupported wire type " + std::to_string(wt));
    }
  }
};

Shape parse_shape(Reader r) {  // TensorShapeProto (tensor_shape.proto:12-45)
  Shape s; s.rank = 0;
  bool unknown_rank = false;
  int f, wt;
  while (r.tag(f, wt)) {
    if (f == 2 && wt == 2) {
      Reader d = r.sub();
      int64_t size = 0;
      int f2, wt2;
      while (d.tag(f2, wt2)) {
        if (f2 == 1 && wt2 == 0) size = int64_t(d.varint());
        else d.skip(wt2);
      }
      TFS_CHECK(s.rank < TFS_MAX_RANK, TFS_ERR_UNSUPPORTED, "tensor rank above " + std::to_string(TFS_MAX_RANK));
      TFS_CHECK(size >= -1, TFS_ERR_INVALID, "shape dimension below -1");
      s.d[s.rank++] = size;
    } else if (f == 3 && wt == 0) {
      unknown_rank = r.varint() != 0;
    } else {
      r.skip(wt);
    }
  }
  if (unknown_rank) s.rank = -1;
  return s;
}

template <typename T>
void append_scalar(std::vector<uint8_t>& out, T v) {
  size_t n = out.size();
  out.resize(n + sizeof(T));
  memcpy(out.data() + n, &v, sizeof(T));
}

TensorValue parse_tensor(Reader r) {  // TensorProto (tensor.proto:13-60)
  TensorValue t;
  t.shape.rank = 0;
  std::string content;
  std::vector<float> fv;
  std::vector<double> dv;
  std::vector<int64_t> iv;   // int_val
  std::vector<int64_t> lv;   // int64_val
  int f, wt;
  while (r.tag(f, wt)) {
    switch (f) {
      case 1: TFS_CHECK(wt == 0, TFS_ERR_INVALID, "TensorProto.dtype wire type"); t.dtype = int(r.varint()); break;
      case 2: TFS_CHECK(wt == 2, TFS_ERR_INVALID, "TensorProto.tensor_shape wire type"); t.shape = parse_shape(r.sub()); break;
      case 4: TFS_CHECK(wt == 2, TFS_ERR_INVALID, "TensorProto.tensor_content wire type"); content = r.str(); break;
      case 5:
        if (wt == 2) { Reader s = r.sub(); while (!s.done()) { uint32_t u = s.fixed32(); float x; memcpy(&x, &u, 4); fv.push_back(x); } }
        else if (wt == 5) { uint32_t u = r.fixed32(); float x; memcpy(&x, &u, 4); fv.push_back(x); }
        else r.skip(wt);
        break;
      case 6:
        if (wt == 2) { Reader s = r.sub(); while (!s.done()) { uint64_t u = s.fixed64(); double x; memcpy(&x, &u, 8); dv.push_back(x); } }
        else if (wt == 1) { uint64_t u = r.fixed64(); double x; memcpy(&x, &u, 8); dv.push_back(x); }
        else r.skip(wt);
        break;
      case 7:
        if (wt == 2) { Reader s = r.sub(); while (!s.done()) iv.push_back(int64_t(int32_t(s.varint()))); }
        else if (wt == 0) iv.push_back(int64_t(int32_t(r.varint())));
        else r.skip(wt);
        break;
      case 10:
        if (wt == 2) { Reader s = r.sub(); while (!s.done()) lv.push_back(int64_t(s.varint())); }
        else if (wt == 0) lv.push_back(int64_t(r.varint()));
        else r.skip(wt);
        break;
      default: r.skip(wt);
    }
  }
  TFS_CHECK(dtype_supported(t.dtype), TFS_ERR_UNSUPPORTED,
            std::string("constant of unsupported dtype code ") + std::to_string(t.dtype));
  TFS_CHECK(t.shape.fully_known(), TFS_ERR_INVALID, "constant with unknown shape");
  const int64_t n = t.shape.numel();
  const int es = dtype_size(t.dtype);
  if (!content.empty()) {
    TFS_CHECK(int64_t(content.size()) == n * es, TFS_ERR_INVALID,
              "tensor_content holds " + std::to_string(content.size()) + " bytes, shape " +
                  t.shape.str() + " needs " + std::to_string(n * es));
    t.bytes.assign(content.begin(), content.end());
    return t;
  }
  // Typed value lists: fewer values than elements repeat the last one
  // ("single value fills the shape", tensor.proto:25-28); none means zeros.
  t.bytes.reserve(size_t(n * es));
  auto emit = [&](auto const& vals, auto cast) {
    TFS_CHECK(int64_t(vals.size()) <= n || n == 0, TFS_ERR_INVALID, "constant has more values than its shape holds");
    for (int64_t i = 0; i < n; ++i) {
      if (vals.empty()) cast(0);
      else cast(vals[size_t(std::min<int64_t>(i, int64_t(vals.size()) - 1))]);
    }
  };
  switch (t.dtype) {
    case TFS_FLOAT: emit(fv, [&](auto v) { append_scalar<float>(t.bytes, float(v)); }); break;
    case TFS_DOUBLE: emit(dv, [&](auto v) { append_scalar<double>(t.bytes, double(v)); }); break;
    case TFS_INT32: emit(iv, [&](auto v) { append_scalar<int32_t>(t.bytes, int32_t(v)); }); break;
    case TFS_INT64: emit(lv, [&](auto v) { append_scalar<int64_t>(t.bytes, int64_t(v)); }); break;
  }
  return t;
}

Attr parse_attr(Reader r) {  // AttrValue (attr_value.proto:16-53)
  Attr a;
  int f, wt;
  while (r.tag(f, wt)) {
    switch (f) {
      case 1: {  // ListValue: only list(int) is kept
        TFS_CHECK(wt == 2, TFS_ERR_INVALID, "AttrValue.list wire type");
        a.kind = Attr::LIST;
        Reader l = r.sub();
        int f2, wt2;
        while (l.tag(f2, wt2)) {
          if (f2 == 3 && wt2 == 2) { Reader s = l.sub(); while (!s.done()) a.list_i.push_back(int64_t(s.varint())); }
          else if (f2 == 3 && wt2 == 0) a.list_i.push_back(int64_t(l.varint()));
          else l.skip(wt2);
        }
        break;
      }
      case 2: a.kind = Attr::S; a.s = r.str(); break;
      case 3: a.kind = Attr::I; a.i = int64_t(r.varint()); break;
      case 4: { a.kind = Attr::F; uint32_t u = r.fixed32(); memcpy(&a.f, &u, 4); break; }
      case 5: a.kind = Attr::B; a.b = r.varint() != 0; break;
      case 6: a.kind = Attr::TYPE; a.type = int(r.varint()); break;
      case 7: a.kind = Attr::SHAPE; a.shape = parse_shape(r.sub()); break;
      case 8: a.kind = Attr::TENSOR; a.tensor = parse_tensor(r.sub()); break;
      default: r.skip(wt);
    }
  }
  return a;
}

Node parse_node(Reader r) {  // NodeDef (graph.proto:60-112)
  Node n;
  int f, wt;
  while (r.tag(f, wt)) {
    switch (f) {
      case 1: n.name = r.str(); break;
      case 2: n.op = r.str(); break;
      case 3: n.raw_inputs.push_back(r.str()); break;
      case 5: {
        Reader e = r.sub();
        std::string key;
        Attr val;
        // A tensor attr may be parsed before we know the op; errors inside name the node later.
        int f2, wt2;
        while (e.tag(f2, wt2)) {
          if (f2 == 1 && wt2 == 2) key = e.str();
          else if (f2 == 2 && wt2 == 2) val = parse_attr(e.sub());
          else e.skip(wt2);
        }
        n.attrs[key] = std::move(val);
        break;
      }
      default: r.skip(wt);
    }
  }
  return n;
}

const Attr* get_attr(const Node& n, const char* key, Attr::Kind kind) {
  auto it = n.attrs.find(key);
  if (it == n.attrs.end() || it->second.kind != kind) return nullptr;
  return &it->second;
}

int attr_type(const Node& n, const char* key) {
  const Attr* a = get_attr(n, key, Attr::TYPE);
  TFS_CHECK(a != nullptr, TFS_ERR_INVALID, "node '" + n.name + "' (" + n.op + ") lacks type attr '" + key + "'");
  TFS_CHECK(dtype_supported(a->type), TFS_ERR_UNSUPPORTED,
            "node '" + n.name + "' (" + n.op + "): dtype code " + std::to_string(a->type) +
                " is not supported (float, double, int32, int64 only)");
  return a->type;
}

bool attr_bool(const Node& n, const char* key, bool dflt) {
  const Attr* a = get_attr(n, key, Attr::B);
  return a ? a->b : dflt;
}

int64_t attr_int(const Node& n, const char* key, int64_t dflt) {
  const Attr* a = get_attr(n, key, Attr::I);
  return a ? a->i : dflt;
}

struct OpDesc { const char* name; OpClass cls; int code; };
const OpDesc kOps[] = {
    {"Placeholder", OpClass::Placeholder, 0},
    {"Const", OpClass::Const, 0},
    {"Identity", OpClass::Unary, E_IDENTITY},
    {"Neg", OpClass::Unary, E_NEG},
    {"Relu", OpClass::Unary, E_RELU},
    {"Square", OpClass::Unary, E_SQUARE},
    {"Abs", OpClass::Unary, E_ABS},
    {"Sqrt", OpClass::Unary, E_SQRT},
    {"Reciprocal", OpClass::Unary, E_RECIP},
    {"Inv", OpClass::Unary, E_RECIP},
    {"Cast", OpClass::Unary, E_CAST},
    {"OnesLike", OpClass::Unary, E_ONES_LIKE},
    {"ZerosLike", OpClass::Unary, E_ZEROS_LIKE},
    {"Shape", OpClass::ShapeOf, 0},
    {"StridedSlice", OpClass::StridedSlice, 0},
    {"Pack", OpClass::Pack, 0},
    {"ExpandDims", OpClass::ExpandDims, 0},
    {"Tile", OpClass::Tile, 0},
    {"ArgMin", OpClass::ArgMinMax, 1},
    {"ArgMax", OpClass::ArgMinMax, 0},
    {"UnsortedSegmentSum", OpClass::SegmentSum, 0},
    {"Add", OpClass::Binary, E_ADD},
    {"BiasAdd", OpClass::Binary, E_ADD},
    {"Sub", OpClass::Binary, E_SUB},
    {"Mul", OpClass::Binary, E_MUL},
    {"Div", OpClass::Binary, E_DIV},
    {"RealDiv", OpClass::Binary, E_DIV},
    {"Minimum", OpClass::Binary, E_MIN},
    {"Maximum", OpClass::Binary, E_MAX},
    {"Sum", OpClass::Reduce, R_SUM},
    {"Min", OpClass::Reduce, R_MIN},
    {"Max", OpClass::Reduce, R_MAX},
    {"MatMul", OpClass::MatMul, 0},
    {"Fill", OpClass::Fill, 0},
};

Shape broadcast(const Shape& a, const Shape& b, const Node& n) {
  if (!a.known_rank() || !b.known_rank()) return Shape();
  const Shape& hi = a.rank >= b.rank ? a : b;
  const Shape& lo = a.rank >= b.rank ? b : a;
  Shape out; out.rank = hi.rank;
  const int off = hi.rank - lo.rank;
  for (int i = 0; i < hi.rank; ++i) {
    int64_t x = hi.d[i];
    int64_t y = i >= off ? lo.d[i - off] : 1;
    int64_t r;
    if (x == y) r = x;
    else if (x == 1) r = y;
    else if (y == 1) r = x;
    else if (x < 0) r = y;       // unknown vs k>1 -> k
    else if (y < 0) r = x;
    else fail(TFS_ERR_INVALID, "node '" + n.name + "' (" + n.op + "): incompatible shapes " + a.str() + " and " + b.str());
    out.d[i] = r;
  }
  return out;
}

std::vector<int64_t> const_ints(const TensorValue& v) {
  std::vector<int64_t> out;
  const int64_t n = v.numel();
  for (int64_t i = 0; i < n; ++i) {
    if (v.dtype == TFS_INT32) { int32_t x; memcpy(&x, v.bytes.data() + 4 * i, 4); out.push_back(x); }
    else { int64_t x; memcpy(&x, v.bytes.data() + 8 * i, 8); out.push_back(x); }
  }
  return out;
}

}  // namespace

int Graph::find(const std::string& name) const {
  std::string base = name;
  size_t c = name.rfind(':');
  if (c != std::string::npos) {
    std::string idx = name.substr(c + 1);
    bool digits = !idx.empty() && std::all_of(idx.begin(), idx.end(), [](char ch) { return ch >= '0' && ch <= '9'; });
    if (digits) {
      TFS_CHECK(idx == "0", TFS_ERR_INVALID,
                "'" + name + "': only output 0 of a node can be addressed");
      base = name.substr(0, c);
    }
  }
  auto it = by_name.find(base);
  return it == by_name.end() ? -1 : it->second;
}

std::vector<int> normalize_axes(const Node& n, int rank) {
  std::set<int> s;
  for (int a : n.axes_raw) {
    int x = a < 0 ? a + rank : a;
    TFS_CHECK(x >= 0 && x < rank, TFS_ERR_INVALID,
              "node '" + n.name + "' (" + n.op + "): reduction index " + std::to_string(a) +
                  " out of range for rank " + std::to_string(rank));
    s.insert(x);
  }
  return std::vector<int>(s.begin(), s.end());
}

namespace {
int norm_axis(const Node& n, int axis, int rank_plus, const char* what) {
  const int x = axis < 0 ? axis + rank_plus : axis;
  TFS_CHECK(x >= 0 && x < rank_plus, TFS_ERR_INVALID,
            "node '" + n.name + "' (" + n.op + "): " + what + " " + std::to_string(axis) + " out of range");
  return x;
}
}  // namespace

bool infer_int_value(const Graph& g, const Node& n, const std::vector<Shape>& in, const IntVals& vals, std::vector<int64_t>& out) {
  (void)g;
  out.clear();
  switch (n.cls) {
    case OpClass::Const:
      if ((n.dtype != TFS_INT32 && n.dtype != TFS_INT64) || n.value.shape.rank > 1 || n.value.numel() > 64) return false;
      out = const_ints(n.value);
      return true;
    case OpClass::ShapeOf:
      if (!in[0].known_rank()) return false;
      for (int i = 0; i < in[0].rank; ++i) out.push_back(in[0].d[i]);
      return true;
    case OpClass::StridedSlice: {
      // 1-D integer vectors only: x[b:e:s] or x[i] (shrink)
      if (!vals[0] || !vals[1] || !vals[2] || !vals[3]) return false;
      const std::vector<int64_t>& x = *vals[0];
      TFS_CHECK(vals[1]->size() == 1 && vals[2]->size() == 1 && vals[3]->size() == 1, TFS_ERR_UNSUPPORTED,
                "node '" + n.name + "' (StridedSlice): only slices of one-dimensional shape vectors are supported");
      const int64_t len = int64_t(x.size());
      int64_t b = (*vals[1])[0], e = (*vals[2])[0], st = (*vals[3])[0];
      TFS_CHECK(st == 1, TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (StridedSlice): only stride 1 is supported");
      if (b < 0) b += len;
      if (n.ss_shrink_mask & 1) {
        TFS_CHECK(b >= 0 && b < len, TFS_ERR_INVALID, "node '" + n.name + "' (StridedSlice): index out of range");
        out.push_back(x[size_t(b)]);
        return true;
      }
      if (e < 0) e += len;
      if (n.ss_begin_mask & 1) b = 0;
      if (n.ss_end_mask & 1) e = len;
      b = std::max<int64_t>(0, std::min(b, len));
      e = std::max<int64_t>(b, std::min(e, len));
      for (int64_t i = b; i < e; ++i) out.push_back(x[size_t(i)]);
      return true;
    }
    case OpClass::Pack:
      for (size_t k = 0; k < n.in.size(); ++k) {
        if (!vals[k] || vals[k]->size() != 1) return false;
        out.push_back((*vals[k])[0]);
      }
      return true;
    default:
      return false;
  }
}

Shape infer_shape(const Graph& g, const Node& n, const std::vector<Shape>& in, const IntVals& vals) {
  switch (n.cls) {
    case OpClass::Placeholder: {
      const Attr* a = get_attr(n, "shape", Attr::SHAPE);
      // TF 1.x legacy rule: a Placeholder whose shape attr has no dims is unconstrained
      // (this is why the reference sends shape hints for every input, core.py:52-72).
      if (!a || a->shape.rank <= 0) return Shape();
      return a->shape;
    }
    case OpClass::Const: return n.value.shape;
    case OpClass::Unary: return in[0];
    case OpClass::Binary: return broadcast(in[0], in[1], n);
    case OpClass::Reduce: {
      const Shape& x = in[0];
      if (!x.known_rank()) return Shape();
      std::vector<int> axes = normalize_axes(n, x.rank);
      Shape out; out.rank = 0;
      for (int i = 0; i < x.rank; ++i) {
        bool red = std::find(axes.begin(), axes.end(), i) != axes.end();
        if (!red) out.d[out.rank++] = x.d[i];
        else if (n.keep_dims) out.d[out.rank++] = 1;
      }
      return out;
    }
    case OpClass::MatMul: {
      const Shape &a = in[0], &b = in[1];
      if (!a.known_rank() || !b.known_rank()) { Shape s; s.rank = 2; s.d[0] = s.d[1] = -1; return s; }
      TFS_CHECK(a.rank == 2 && b.rank == 2, TFS_ERR_INVALID,
                "node '" + n.name + "' (MatMul): operands must be matrices, got " + a.str() + " and " + b.str());
      int64_t m = n.transpose_a ? a.d[1] : a.d[0];
      int64_t ka = n.transpose_a ? a.d[0] : a.d[1];
      int64_t kb = n.transpose_b ? b.d[1] : b.d[0];
      int64_t nn = n.transpose_b ? b.d[0] : b.d[1];
      TFS_CHECK(ka < 0 || kb < 0 || ka == kb, TFS_ERR_INVALID,
                "node '" + n.name + "' (MatMul): inner dimensions differ, " + a.str() + " x " + b.str());
      return Shape::of({m, nn});
    }
    case OpClass::Fill: {
      const Node& dims = g.nodes[size_t(n.in[0])];
      TFS_CHECK(dims.cls == OpClass::Const, TFS_ERR_UNSUPPORTED,
                "node '" + n.name + "' (Fill): dims must be a Const");
      Shape out; out.rank = 0;
      for (int64_t v : const_ints(dims.value)) {
        TFS_CHECK(out.rank < TFS_MAX_RANK && v >= 0, TFS_ERR_INVALID, "node '" + n.name + "' (Fill): bad dims");
        out.d[out.rank++] = v;
      }
      return out;
    }
    case OpClass::ShapeOf: {
      Shape out; out.rank = 1;
      out.d[0] = in[0].known_rank() ? in[0].rank : -1;
      return out;
    }
    case OpClass::StridedSlice:
    case OpClass::Pack: {
      // (the run-time planner insists on the value; at import an unknown one only leaves the length open)
      std::vector<int64_t> v;
      const bool known = infer_int_value(g, n, in, vals, v);
      Shape out;
      if (n.cls == OpClass::StridedSlice && (n.ss_shrink_mask & 1)) { out.rank = 0; return out; }
      out.rank = 1;
      out.d[0] = known ? int64_t(v.size()) : (n.cls == OpClass::Pack ? int64_t(n.in.size()) : -1);
      return out;
    }
    case OpClass::ExpandDims: {
      const Shape& x = in[0];
      if (!x.known_rank()) return Shape();
      TFS_CHECK(x.rank + 1 <= TFS_MAX_RANK, TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (ExpandDims): rank too large");
      const int ax = norm_axis(n, n.axis, x.rank + 1, "dim");
      Shape out; out.rank = x.rank + 1;
      for (int i = 0, j = 0; i < out.rank; ++i) out.d[i] = (i == ax) ? 1 : x.d[j++];
      return out;
    }
    case OpClass::Tile: {
      const Shape& x = in[0];
      if (!x.known_rank()) return Shape();
      Shape out; out.rank = x.rank;
      for (int i = 0; i < x.rank; ++i) out.d[i] = -1;
      if (vals[1]) {
        TFS_CHECK(int(vals[1]->size()) == x.rank, TFS_ERR_INVALID,
                  "node '" + n.name + "' (Tile): multiples has " + std::to_string(vals[1]->size()) + " entries for an input of rank " +
                      std::to_string(x.rank));
        for (int i = 0; i < x.rank; ++i) {
          const int64_t m = (*vals[1])[size_t(i)];
          TFS_CHECK(m >= -1, TFS_ERR_INVALID, "node '" + n.name + "' (Tile): negative multiple");
          out.d[i] = (m < 0 || x.d[i] < 0) ? -1 : m * x.d[i];
        }
      }
      return out;
    }
    case OpClass::ArgMinMax: {
      const Shape& x = in[0];
      if (!x.known_rank()) return Shape();
      TFS_CHECK(x.rank >= 1, TFS_ERR_INVALID, "node '" + n.name + "' (" + n.op + "): input must have at least one dimension");
      const int ax = norm_axis(n, n.axis, x.rank, "dimension");
      Shape out; out.rank = 0;
      for (int i = 0; i < x.rank; ++i)
        if (i != ax) out.d[out.rank++] = x.d[i];
      return out;
    }
    case OpClass::SegmentSum: {
      const Shape &x = in[0], &ids = in[1];
      if (!x.known_rank() || !ids.known_rank()) return Shape();
      TFS_CHECK(ids.rank == 1 && x.rank >= 1, TFS_ERR_UNSUPPORTED,
                "node '" + n.name + "' (UnsortedSegmentSum): segment_ids must be a vector over the rows of data");
      TFS_CHECK(x.d[0] < 0 || ids.d[0] < 0 || x.d[0] == ids.d[0], TFS_ERR_INVALID,
                "node '" + n.name + "' (UnsortedSegmentSum): data has " + std::to_string(x.d[0]) + " rows, segment_ids " +
                    std::to_string(ids.d[0]));
      Shape out = x;
      out.d[0] = (vals[2] && vals[2]->size() == 1) ? (*vals[2])[0] : -1;
      return out;
    }
  }
  return Shape();
}

namespace {
int desc_code_of(const Node& n) { return n.op == "ArgMin" ? 1 : 0; }
}  // namespace

Graph import_graphdef(const uint8_t* data, size_t len) {
  Graph g;
  TFS_CHECK(data != nullptr || len == 0, TFS_ERR_INVALID, "null GraphDef buffer");
  Reader r(data, len);
  int f, wt;
  while (r.tag(f, wt)) {
    if (f == 1 && wt == 2) g.nodes.push_back(parse_node(r.sub()));
    else r.skip(wt);
  }
  for (size_t i = 0; i < g.nodes.size(); ++i) {
    Node& n = g.nodes[i];
    TFS_CHECK(!n.name.empty(), TFS_ERR_INVALID, "GraphDef: node without a name");
    TFS_CHECK(g.by_name.emplace(n.name, int(i)).second, TFS_ERR_INVALID,
              "GraphDef: duplicate node name '" + n.name + "'");
  }
  // Classify, resolve inputs.
  for (Node& n : g.nodes) {
    const OpDesc* desc = nullptr;
    for (const OpDesc& d : kOps) if (n.op == d.name) desc = &d;
    TFS_CHECK(desc != nullptr, TFS_ERR_UNSUPPORTED,
              "node '" + n.name + "': op '" + n.op + "' is not supported by the B200 executor");
    n.cls = desc->cls;
    if (n.cls == OpClass::Unary || n.cls == OpClass::Binary) n.elt_op = desc->code;
    if (n.cls == OpClass::Reduce) n.red_op = desc->code;
    for (const std::string& in : n.raw_inputs) {
      TFS_CHECK(!in.empty() && in[0] != '^', TFS_ERR_UNSUPPORTED,
                "node '" + n.name + "': control input '" + in + "' is not supported");
      int idx = g.find(in);
      TFS_CHECK(idx >= 0, TFS_ERR_INVALID, "node '" + n.name + "': input '" + in + "' is not in the graph");
      n.in.push_back(idx);
    }
    size_t want = 0;
    switch (n.cls) {
      case OpClass::Placeholder: case OpClass::Const: want = 0; break;
      case OpClass::Unary: case OpClass::ShapeOf: want = 1; break;
      case OpClass::StridedSlice: want = 4; break;
      case OpClass::SegmentSum: want = 3; break;
      case OpClass::Pack: want = n.in.size(); break;
      default: want = 2;
    }
    if (n.cls == OpClass::Pack)
      TFS_CHECK(!n.in.empty() && int64_t(n.in.size()) == attr_int(n, "N", int64_t(n.in.size())), TFS_ERR_INVALID,
                "node '" + n.name + "' (Pack): attr N differs from the number of inputs");
    TFS_CHECK(n.in.size() == want, TFS_ERR_INVALID,
              "node '" + n.name + "' (" + n.op + "): expected " + std::to_string(want) + " inputs, found " +
                  std::to_string(n.in.size()));
  }
  // Topological order (also rejects cycles); GraphDef order is not guaranteed to be topological.
  std::vector<int> order, state(g.nodes.size(), 0);
  for (size_t root = 0; root < g.nodes.size(); ++root) {
    if (state[root]) continue;
    std::vector<std::pair<int, size_t>> stack{{int(root), 0}};
    state[root] = 1;
    while (!stack.empty()) {
      auto& [u, k] = stack.back();
      if (k < g.nodes[size_t(u)].in.size()) {
        int v = g.nodes[size_t(u)].in[k++];
        TFS_CHECK(state[size_t(v)] != 1, TFS_ERR_INVALID, "GraphDef: cycle through node '" + g.nodes[size_t(v)].name + "'");
        if (state[size_t(v)] == 0) { state[size_t(v)] = 1; stack.push_back({v, 0}); }
      } else {
        state[size_t(u)] = 2;
        order.push_back(u);
        stack.pop_back();
      }
    }
  }
  // dtype + static shape per node, in dependency order.
  for (int idx : order) {
    Node& n = g.nodes[size_t(idx)];
    auto in_dtype = [&](size_t k) { return g.nodes[size_t(n.in[k])].dtype; };
    switch (n.cls) {
      case OpClass::Placeholder:
        n.dtype = attr_type(n, "dtype");
        break;
      case OpClass::Const: {
        const Attr* v = get_attr(n, "value", Attr::TENSOR);
        TFS_CHECK(v != nullptr, TFS_ERR_INVALID, "node '" + n.name + "' (Const) lacks a tensor 'value'");
        n.value = v->tensor;
        n.dtype = n.value.dtype;
        auto d = n.attrs.find("dtype");
        if (d != n.attrs.end() && d->second.kind == Attr::TYPE)
          TFS_CHECK(d->second.type == n.dtype, TFS_ERR_INVALID, "node '" + n.name + "' (Const): dtype attr differs from value dtype");
        break;
      }
      case OpClass::Unary:
        if (n.elt_op == E_CAST) {
          int src = attr_type(n, "SrcT");
          n.dtype = attr_type(n, "DstT");
          TFS_CHECK(src == in_dtype(0), TFS_ERR_INVALID, "node '" + n.name + "' (Cast): SrcT differs from its input dtype");
        } else {
          n.dtype = attr_type(n, "T");
          TFS_CHECK(n.dtype == in_dtype(0), TFS_ERR_INVALID,
                    "node '" + n.name + "' (" + n.op + "): attr T is " + dtype_name(n.dtype) + " but the input is " + dtype_name(in_dtype(0)));
          if (n.elt_op == E_SQRT || n.elt_op == E_RECIP)
            TFS_CHECK(n.dtype == TFS_FLOAT || n.dtype == TFS_DOUBLE, TFS_ERR_UNSUPPORTED,
                      "node '" + n.name + "' (" + n.op + "): only float and double are supported");
        }
        break;
      case OpClass::Binary:
      case OpClass::MatMul:
        n.dtype = attr_type(n, "T");
        TFS_CHECK(n.dtype == in_dtype(0) && n.dtype == in_dtype(1), TFS_ERR_INVALID,
                  "node '" + n.name + "' (" + n.op + "): attr T is " + dtype_name(n.dtype) + " but the inputs are " +
                      dtype_name(in_dtype(0)) + " and " + dtype_name(in_dtype(1)));
        if (n.cls == OpClass::MatMul) {
          TFS_CHECK(n.dtype == TFS_FLOAT || n.dtype == TFS_DOUBLE, TFS_ERR_UNSUPPORTED,
                    "node '" + n.name + "' (MatMul): only float32 and float64 are supported");
          n.transpose_a = attr_bool(n, "transpose_a", false);
          n.transpose_b = attr_bool(n, "transpose_b", false);
        }
        break;
      case OpClass::Reduce: {
        n.dtype = attr_type(n, "T");
        TFS_CHECK(n.dtype == in_dtype(0), TFS_ERR_INVALID, "node '" + n.name + "' (" + n.op + "): attr T differs from its input dtype");
        const Node& ax = g.nodes[size_t(n.in[1])];
        TFS_CHECK(ax.cls == OpClass::Const && (ax.dtype == TFS_INT32 || ax.dtype == TFS_INT64), TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (" + n.op + "): reduction_indices must be an integer Const");
        for (int64_t a : const_ints(ax.value)) n.axes_raw.push_back(int(a));
        n.keep_dims = attr_bool(n, "keep_dims", false);
        break;
      }
      case OpClass::ShapeOf: {
        auto ot = n.attrs.find("out_type");
        n.dtype = (ot != n.attrs.end() && ot->second.kind == Attr::TYPE) ? ot->second.type : TFS_INT32;
        TFS_CHECK(n.dtype == TFS_INT32 || n.dtype == TFS_INT64, TFS_ERR_INVALID, "node '" + n.name + "' (Shape): out_type must be int32 or int64");
        break;
      }
      case OpClass::StridedSlice:
        n.dtype = in_dtype(0);
        n.ss_begin_mask = int(attr_int(n, "begin_mask", 0));
        n.ss_end_mask = int(attr_int(n, "end_mask", 0));
        n.ss_shrink_mask = int(attr_int(n, "shrink_axis_mask", 0));
        TFS_CHECK(attr_int(n, "ellipsis_mask", 0) == 0 && attr_int(n, "new_axis_mask", 0) == 0, TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (StridedSlice): ellipsis / new-axis slicing is not supported");
        break;
      case OpClass::Pack:
        n.dtype = in_dtype(0);
        for (size_t k = 1; k < n.in.size(); ++k)
          TFS_CHECK(in_dtype(k) == n.dtype, TFS_ERR_INVALID, "node '" + n.name + "' (Pack): inputs differ in dtype");
        TFS_CHECK(attr_int(n, "axis", 0) == 0, TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (Pack): only axis 0 is supported");
        break;
      case OpClass::ExpandDims:
      case OpClass::ArgMinMax: {
        const Node& ax = g.nodes[size_t(n.in[1])];
        TFS_CHECK(ax.cls == OpClass::Const && (ax.dtype == TFS_INT32 || ax.dtype == TFS_INT64) && ax.value.numel() == 1,
                  TFS_ERR_UNSUPPORTED, "node '" + n.name + "' (" + n.op + "): the axis must be an integer scalar Const");
        n.axis = int(const_ints(ax.value)[0]);
        if (n.cls == OpClass::ExpandDims) {
          n.dtype = in_dtype(0);
        } else {
          auto ot = n.attrs.find("output_type");
          n.dtype = (ot != n.attrs.end() && ot->second.kind == Attr::TYPE) ? ot->second.type : TFS_INT64;
          TFS_CHECK(n.dtype == TFS_INT32 || n.dtype == TFS_INT64, TFS_ERR_INVALID,
                    "node '" + n.name + "' (" + n.op + "): output_type must be int32 or int64");
          n.arg_min = desc_code_of(n) == 1;
        }
        break;
      }
      case OpClass::Tile:
        n.dtype = in_dtype(0);
        TFS_CHECK(in_dtype(1) == TFS_INT32 || in_dtype(1) == TFS_INT64, TFS_ERR_INVALID, "node '" + n.name + "' (Tile): multiples must be integers");
        break;
      case OpClass::SegmentSum:
        n.dtype = in_dtype(0);
        TFS_CHECK(in_dtype(1) == TFS_INT32 || in_dtype(1) == TFS_INT64, TFS_ERR_INVALID,
                  "node '" + n.name + "' (UnsortedSegmentSum): segment_ids must be int32 or int64");
        break;
      case OpClass::Fill: {
        n.dtype = attr_type(n, "T");
        TFS_CHECK(n.dtype == in_dtype(1), TFS_ERR_INVALID, "node '" + n.name + "' (Fill): attr T differs from the value dtype");
        const Node& dims = g.nodes[size_t(n.in[0])];
        TFS_CHECK(dims.cls == OpClass::Const && dims.dtype == TFS_INT32, TFS_ERR_UNSUPPORTED,
                  "node '" + n.name + "' (Fill): dims must be an int32 Const");
        break;
      }
    }
    std::vector<Shape> in_shapes;
    IntVals in_vals;
    for (int k : n.in) {
      in_shapes.push_back(g.nodes[size_t(k)].sshape);
      in_vals.push_back(g.nodes[size_t(k)].has_ival ? &g.nodes[size_t(k)].ival : nullptr);
    }
    n.has_ival = infer_int_value(g, n, in_shapes, in_vals, n.ival);
    n.sshape = infer_shape(g, n, in_shapes, in_vals);
  }
  return g;
}

}  // namespace tfs
