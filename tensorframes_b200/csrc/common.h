// Shared small types for libtfsb200: error type, dtypes, shapes.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/tfs_b200.h"

namespace tfs {

// Exception carrying a TFS_ERR_* code; never crosses the C ABI (capi.cpp catches).
struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
[[noreturn]] inline void fail(int code, const std::string& m) { throw Error(code, m); }
#define TFS_CHECK(cond, code, msg) \
  do { if (!(cond)) ::tfs::fail((code), (msg)); } while (0)

// dtype codes follow types.proto (DT_FLOAT=1, DT_DOUBLE=2, DT_INT32=3, DT_INT64=9).
inline bool dtype_supported(int dt) {
  return dt == TFS_FLOAT || dt == TFS_DOUBLE || dt == TFS_INT32 || dt == TFS_INT64;
}
inline int dtype_size(int dt) {
  switch (dt) {
    case TFS_FLOAT: case TFS_INT32: return 4;
    case TFS_DOUBLE: case TFS_INT64: return 8;
    default: return 0;
  }
}
inline const char* dtype_name(int dt) {
  switch (dt) {
    case TFS_FLOAT: return "float";
    case TFS_DOUBLE: return "double";
    case TFS_INT32: return "int32";
    case TFS_INT64: return "int64";
    case 0: return "invalid";
    default: return "unsupported";
  }
}
// Dense index 0..3 used by kernels: f32, f64, i32, i64.
inline int dtype_index(int dt) {
  switch (dt) {
    case TFS_FLOAT: return 0;
    case TFS_DOUBLE: return 1;
    case TFS_INT32: return 2;
    case TFS_INT64: return 3;
    default: return -1;
  }
}

// Shape with TensorFrames semantics: dim -1 = unknown (Shape.scala:96), rank -1 = unknown rank.
struct Shape {
  int rank = -1;
  int64_t d[TFS_MAX_RANK] = {0, 0, 0, 0, 0, 0, 0, 0};
  static Shape scalar() { Shape s; s.rank = 0; return s; }
  static Shape of(std::initializer_list<int64_t> l) {
    Shape s; s.rank = 0;
    for (auto v : l) s.d[s.rank++] = v;
    return s;
  }
  bool known_rank() const { return rank >= 0; }
  bool fully_known() const {
    if (rank < 0) return false;
    for (int i = 0; i < rank; ++i) if (d[i] < 0) return false;
    return true;
  }
  int64_t numel() const {  // requires fully_known
    int64_t n = 1;
    for (int i = 0; i < rank; ++i) n *= d[i];
    return n;
  }
  bool operator==(const Shape& o) const {
    if (rank != o.rank) return false;
    for (int i = 0; i < rank; ++i) if (d[i] != o.d[i]) return false;
    return true;
  }
  bool operator!=(const Shape& o) const { return !(*this == o); }
  std::string str() const {
    if (rank < 0) return "<unknown rank>";
    std::string s = "[";
    for (int i = 0; i < rank; ++i) {
      if (i) s += ",";
      s += d[i] < 0 ? std::string("?") : std::to_string(d[i]);
    }
    return s + "]";
  }
};

}  // namespace tfs
