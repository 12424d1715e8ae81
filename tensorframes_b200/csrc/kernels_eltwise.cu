// K1 -- fused element-wise group kernel (sm_100a).
//
// One launch evaluates every element-wise node of a graph that iterates over the
// same block (e.g. Mul -> Add -> Maximum -> Cast of BASELINE config 3) in a single
// pass over HBM: each input column is read once, each fetched column is written
// once, and no intermediate tensor exists.  TF-CPU, which the reference calls at
// impl/DebugRowOps.scala:794, materialises every intermediate.
//
// Execution model: a persistent grid; each thread owns kJ vectors of P elements
// per tile (P*sizeof(widest dtype) = 16 bytes, so every global access is a
// coalesced 128-bit -- or 64-bit for a 4-byte column next to an 8-byte one --
// transaction, kJ of them in flight per thread).  The group's program
// (program.h) is interpreted once per tile with the accumulator tile held in
// registers.  Each instruction is one flat opcode (operation, dtype, operand
// kind and operand order pre-resolved on the host), so decoding is a single
// warp-uniform jump whose cost is amortised over kJ*P elements; immediates are
// applied straight from the instruction word.  HBM-bound: the roofline is
// (bytes read + bytes written) / HBM bandwidth.
//
// Arithmetic follows the TF 1.1 CPU kernels: IEEE round-to-nearest for + - * /
// with no FMA contraction (explicit __dadd_rn/__dmul_rn...), Minimum/Maximum as
// Eigen's scalar (b < a ? b : a), Cast as C++ static_cast, integer arithmetic
// wrapping, integer division truncating toward zero and flagged when the divisor
// is zero (TF raises "Integer division by zero").
#include <type_traits>

#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;
// 8 vectors per thread and dispatch (was 4): the cost of decoding one instruction of the program is paid per tile, so a
// wider tile halves it per element; together with a real jump table for the dispatch switch (built with
// --jump-table-density, see __graft_entry__.py: nvcc's default lowers the ~150-case switch to a tree of 7-8
// compare-and-branch pairs) the C3 chain forced through this kernel went from 2875 to 5243 GB/s.
constexpr int kJ = 8;

template <int ES> struct Slot;
template <> struct Slot<8> { using type = unsigned long long; };
template <> struct Slot<4> { using type = unsigned int; };

// bit casts between the untyped register slots and the typed values (pure register moves)
template <class T> struct Bits;
template <> struct Bits<float> {
  static __device__ __forceinline__ float from(unsigned long long s) { return __uint_as_float((unsigned int)s); }
  static __device__ __forceinline__ float from(unsigned int s) { return __uint_as_float(s); }
  static __device__ __forceinline__ unsigned int to(float v) { return __float_as_uint(v); }
};
template <> struct Bits<int> {
  static __device__ __forceinline__ int from(unsigned long long s) { return (int)(unsigned int)s; }
  static __device__ __forceinline__ int from(unsigned int s) { return (int)s; }
  static __device__ __forceinline__ unsigned int to(int v) { return (unsigned int)v; }
};
template <> struct Bits<double> {
  static __device__ __forceinline__ double from(unsigned long long s) { return __longlong_as_double((long long)s); }
  static __device__ __forceinline__ double from(unsigned int) { return 0.0; }   // never instantiated meaningfully
  static __device__ __forceinline__ unsigned long long to(double v) { return (unsigned long long)__double_as_longlong(v); }
};
template <> struct Bits<long long> {
  static __device__ __forceinline__ long long from(unsigned long long s) { return (long long)s; }
  static __device__ __forceinline__ long long from(unsigned int) { return 0; }
  static __device__ __forceinline__ unsigned long long to(long long v) { return (unsigned long long)v; }
};
template <class T, class S>
__device__ __forceinline__ T from_slot(S s) { return Bits<T>::from(s); }
template <class S, class T>
__device__ __forceinline__ S to_slot(T t) { return (S)Bits<T>::to(t); }

#include "arith.inc"

// ---- global memory access -------------------------------------------------
template <int ES, int P>
__device__ __forceinline__ void load_vec(const BufRef& b, int64_t elem, int nvalid,
                                         typename Slot<ES>::type (&x)[P]) {
  using S = typename Slot<ES>::type;
  if (ES == 8 && (b.dt & 1)) {  // 8-byte elements
    const unsigned long long* p = static_cast<const unsigned long long*>(b.ptr) + elem;
    if (P == 2 && nvalid == 2) {
      ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2*>(p));
      x[0] = S(v.x);
      x[P - 1] = S(v.y);
    } else {
#pragma unroll
      for (int e = 0; e < P; ++e) x[e] = e < nvalid ? S(__ldcs(p + e)) : S(0);
    }
  } else {  // 4-byte elements
    const unsigned int* p = static_cast<const unsigned int*>(b.ptr) + elem;
    if (P == 4 && nvalid == 4) {
      uint4 v = __ldcs(reinterpret_cast<const uint4*>(p));
      x[0] = S(v.x);
      x[1 % P] = S(v.y);
      x[2 % P] = S(v.z);
      x[3 % P] = S(v.w);
    } else if (P == 2 && nvalid == 2) {
      uint2 v = __ldcs(reinterpret_cast<const uint2*>(p));
      x[0] = S(v.x);
      x[P - 1] = S(v.y);
    } else {
#pragma unroll
      for (int e = 0; e < P; ++e) x[e] = e < nvalid ? S(__ldcs(p + e)) : S(0);
    }
  }
}

// Operand broadcast along the rows (a [cell] vector against a [n, cell] block): element i reads b[i % period].
// One modulo per VECTOR (32-bit when the index allows), then either one cached vector load -- BufRef.pad is set by
// the host when the period is a multiple of the vector width and the buffer is 16-byte aligned, so a vector never
// wraps -- or per-element loads with a wrap test.
template <int ES, int P>
__device__ __forceinline__ void load_periodic(const BufRef& b, int64_t elem, int nvalid,
                                              typename Slot<ES>::type (&x)[P]) {
  using S = typename Slot<ES>::type;
  const int64_t period = b.period;
  const int64_t i0 = ((elem | period) >> 32) == 0 ? int64_t((unsigned int)elem % (unsigned int)period) : elem % period;
  const bool wide_elems = ES == 8 && (b.dt & 1);
  if (b.pad != 0 && nvalid == P) {
    if (wide_elems) {
      if (P == 2) {
        const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(static_cast<const unsigned long long*>(b.ptr) + i0));
        x[0] = S(v.x);
        x[P - 1] = S(v.y);
        return;
      }
    } else if (P == 4) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(static_cast<const unsigned int*>(b.ptr) + i0));
      x[0] = S(v.x);
      x[1 % P] = S(v.y);
      x[2 % P] = S(v.z);
      x[3 % P] = S(v.w);
      return;
    } else if (P == 2) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(static_cast<const unsigned int*>(b.ptr) + i0));
      x[0] = S(v.x);
      x[P - 1] = S(v.y);
      return;
    }
  }
#pragma unroll
  for (int e = 0; e < P; ++e) {
    S v = 0;
    if (e < nvalid) {
      int64_t i = i0 + e;
      if (i >= period) i = period >= P ? i - period : i % period;
      if (wide_elems) v = S(__ldg(static_cast<const unsigned long long*>(b.ptr) + i));
      else v = S(__ldg(static_cast<const unsigned int*>(b.ptr) + i));
    }
    x[e] = v;
  }
}

// Row mode: an operand with one (smaller) cell per row, broadcast inside the row: element i of the block reads
// b[(i / rowspan) * period + i % period].  Off the roofline path (per-row normalisations under mapRows); plain
// per-element index arithmetic, 32-bit when the index allows.
template <int ES, int P>
__device__ __forceinline__ void load_rowbcast(const BufRef& b, int64_t elem, int nvalid,
                                              typename Slot<ES>::type (&x)[P]) {
  using S = typename Slot<ES>::type;
  const bool wide_elems = ES == 8 && (b.dt & 1);
#pragma unroll
  for (int e = 0; e < P; ++e) {
    S v = 0;
    if (e < nvalid) {
      const int64_t i = elem + e;
      int64_t src;
      if (((i | b.rowspan) >> 32) == 0) {
        const unsigned int r = (unsigned int)i / (unsigned int)b.rowspan;
        src = int64_t(r) * b.period + (b.period == 1 ? 0 : int64_t(((unsigned int)i - r * (unsigned int)b.rowspan) % (unsigned int)b.period));
      } else {
        src = (i / b.rowspan) * b.period + (b.period == 1 ? 0 : (i % b.rowspan) % b.period);
      }
      if (wide_elems) v = S(__ldg(static_cast<const unsigned long long*>(b.ptr) + src));
      else v = S(__ldg(static_cast<const unsigned int*>(b.ptr) + src));
    }
    x[e] = v;
  }
}

template <int ES, int P>
__device__ __forceinline__ void store_vec(const BufRef& b, int64_t elem, int nvalid,
                                          const typename Slot<ES>::type (&x)[P]) {
  if (ES == 8 && (b.dt & 1)) {
    unsigned long long* p = static_cast<unsigned long long*>(b.ptr) + elem;
    if (P == 2 && nvalid == 2) {
      __stcs(reinterpret_cast<ulonglong2*>(p), make_ulonglong2(x[0], x[P - 1]));
    } else {
#pragma unroll
      for (int e = 0; e < P; ++e)
        if (e < nvalid) __stcs(p + e, static_cast<unsigned long long>(x[e]));
    }
  } else {
    unsigned int* p = static_cast<unsigned int*>(b.ptr) + elem;
    if (P == 4 && nvalid == 4) {
      __stcs(reinterpret_cast<uint4*>(p),
             make_uint4((unsigned)x[0], (unsigned)x[1 % P], (unsigned)x[2 % P], (unsigned)x[3 % P]));
    } else if (P == 2 && nvalid == 2) {
      __stcs(reinterpret_cast<uint2*>(p), make_uint2((unsigned)x[0], (unsigned)x[P - 1]));
    } else {
#pragma unroll
      for (int e = 0; e < P; ++e)
        if (e < nvalid) __stcs(p + e, static_cast<unsigned int>(x[e]));
    }
  }
}

// ---- typed handlers over the register tile ----------------------------------------------------
template <class T, int SUB>
__device__ __forceinline__ T bin_eval(T a, T b, bool valid, int* err) {
  if constexpr (SUB == B_ADD) return Arith<T>::add(a, b);
  else if constexpr (SUB == B_SUB) return Arith<T>::sub(a, b);
  else if constexpr (SUB == B_MUL) return Arith<T>::mul(a, b);
  else if constexpr (SUB == B_DIV) return Arith<T>::div(a, b, valid, err);
  else if constexpr (SUB == B_MIN) return Arith<T>::mn(a, b);
  else return Arith<T>::mx(a, b);
}

template <class T, int SUB, bool SWAP, int ES, int P>
__device__ __forceinline__ void do_bin_x(typename Slot<ES>::type (&acc)[kJ][P],
                                         const typename Slot<ES>::type (&x)[kJ][P], const int (&nvalid)[kJ], int* err) {
  using S = typename Slot<ES>::type;
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) {
      const T l = from_slot<T, S>(acc[j][e]);
      const T r = from_slot<T, S>(x[j][e]);
      acc[j][e] = to_slot<S, T>(SWAP ? bin_eval<T, SUB>(r, l, e < nvalid[j], err) : bin_eval<T, SUB>(l, r, e < nvalid[j], err));
    }
}

template <class T, int SUB, bool SWAP, int ES, int P>
__device__ __forceinline__ void do_bin_imm(typename Slot<ES>::type (&acc)[kJ][P], const unsigned long long imm,
                                           const int (&nvalid)[kJ], int* err) {
  using S = typename Slot<ES>::type;
  const T r = from_slot<T, unsigned long long>(imm);
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) {
      const T l = from_slot<T, S>(acc[j][e]);
      acc[j][e] = to_slot<S, T>(SWAP ? bin_eval<T, SUB>(r, l, e < nvalid[j], err) : bin_eval<T, SUB>(l, r, e < nvalid[j], err));
    }
}

template <class T, int SUB, int ES, int P>
__device__ __forceinline__ void do_unary(typename Slot<ES>::type (&acc)[kJ][P]) {
  using S = typename Slot<ES>::type;
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) {
      const T a = from_slot<T, S>(acc[j][e]);
      T r;
      if constexpr (SUB == U_NEG) r = Arith<T>::neg(a);
      else if constexpr (SUB == U_RELU) r = Arith<T>::mx(a, T(0));
      else if constexpr (SUB == U_SQUARE) r = Arith<T>::mul(a, a);
      else if constexpr (SUB == U_ABS) r = Arith<T>::abs(a);
      else if constexpr (SUB == U_SQRT) r = Arith<T>::sqrt(a);
      else r = Arith<T>::recip(a);
      acc[j][e] = to_slot<S, T>(r);
    }
}

template <class From, class To, int ES, int P>
__device__ __forceinline__ void do_cast(typename Slot<ES>::type (&acc)[kJ][P]) {
  using S = typename Slot<ES>::type;
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e)
      acc[j][e] = to_slot<S, To>(static_cast<To>(from_slot<From, S>(acc[j][e])));
}

template <int ES, int P>
__device__ __forceinline__ void copy_tile(typename Slot<ES>::type (&dst)[kJ][P], const typename Slot<ES>::type (&src)[kJ][P]) {
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e) dst[j][e] = src[j][e];
}

// 64-bit dtypes only exist in the ES == 8 instantiation
#define TFS_WIDE(stmt) if constexpr (ES == 8) { stmt; }

// the 8 opcodes of one binary sub-op with a tensor operand / with an immediate
#define TFS_BIN_CASES(SUB)                                                                             \
  case C_BIN_X + SUB * 8 + DI_F32 * 2 + 0: do_bin_x<float, SUB, false, ES, P>(acc, x, nvalid, prog.err); break;       \
  case C_BIN_X + SUB * 8 + DI_F32 * 2 + 1: do_bin_x<float, SUB, true, ES, P>(acc, x, nvalid, prog.err); break;        \
  case C_BIN_X + SUB * 8 + DI_I32 * 2 + 0: do_bin_x<int, SUB, false, ES, P>(acc, x, nvalid, prog.err); break;         \
  case C_BIN_X + SUB * 8 + DI_I32 * 2 + 1: do_bin_x<int, SUB, true, ES, P>(acc, x, nvalid, prog.err); break;          \
  case C_BIN_X + SUB * 8 + DI_F64 * 2 + 0: TFS_WIDE((do_bin_x<double, SUB, false, ES, P>(acc, x, nvalid, prog.err))) break;    \
  case C_BIN_X + SUB * 8 + DI_F64 * 2 + 1: TFS_WIDE((do_bin_x<double, SUB, true, ES, P>(acc, x, nvalid, prog.err))) break;     \
  case C_BIN_X + SUB * 8 + DI_I64 * 2 + 0: TFS_WIDE((do_bin_x<long long, SUB, false, ES, P>(acc, x, nvalid, prog.err))) break; \
  case C_BIN_X + SUB * 8 + DI_I64 * 2 + 1: TFS_WIDE((do_bin_x<long long, SUB, true, ES, P>(acc, x, nvalid, prog.err))) break;  \
  case C_BIN_IMM + SUB * 8 + DI_F32 * 2 + 0: do_bin_imm<float, SUB, false, ES, P>(acc, in.imm, nvalid, prog.err); break;     \
  case C_BIN_IMM + SUB * 8 + DI_F32 * 2 + 1: do_bin_imm<float, SUB, true, ES, P>(acc, in.imm, nvalid, prog.err); break;      \
  case C_BIN_IMM + SUB * 8 + DI_I32 * 2 + 0: do_bin_imm<int, SUB, false, ES, P>(acc, in.imm, nvalid, prog.err); break;       \
  case C_BIN_IMM + SUB * 8 + DI_I32 * 2 + 1: do_bin_imm<int, SUB, true, ES, P>(acc, in.imm, nvalid, prog.err); break;        \
  case C_BIN_IMM + SUB * 8 + DI_F64 * 2 + 0: TFS_WIDE((do_bin_imm<double, SUB, false, ES, P>(acc, in.imm, nvalid, prog.err))) break;    \
  case C_BIN_IMM + SUB * 8 + DI_F64 * 2 + 1: TFS_WIDE((do_bin_imm<double, SUB, true, ES, P>(acc, in.imm, nvalid, prog.err))) break;     \
  case C_BIN_IMM + SUB * 8 + DI_I64 * 2 + 0: TFS_WIDE((do_bin_imm<long long, SUB, false, ES, P>(acc, in.imm, nvalid, prog.err))) break; \
  case C_BIN_IMM + SUB * 8 + DI_I64 * 2 + 1: TFS_WIDE((do_bin_imm<long long, SUB, true, ES, P>(acc, in.imm, nvalid, prog.err))) break;

#define TFS_UN_CASES(SUB)                                                                   \
  case C_UNARY + SUB * 4 + DI_F32: do_unary<float, SUB, ES, P>(acc); break;                 \
  case C_UNARY + SUB * 4 + DI_I32: do_unary<int, SUB, ES, P>(acc); break;                   \
  case C_UNARY + SUB * 4 + DI_F64: TFS_WIDE((do_unary<double, SUB, ES, P>(acc))) break;     \
  case C_UNARY + SUB * 4 + DI_I64: TFS_WIDE((do_unary<long long, SUB, ES, P>(acc))) break;

#define TFS_CAST_CASES(FROM_DI, FROM_T, GUARD)                                                        \
  case C_CAST + FROM_DI * 4 + DI_F32: GUARD((do_cast<FROM_T, float, ES, P>(acc))) break;              \
  case C_CAST + FROM_DI * 4 + DI_I32: GUARD((do_cast<FROM_T, int, ES, P>(acc))) break;                \
  case C_CAST + FROM_DI * 4 + DI_F64: TFS_WIDE((do_cast<FROM_T, double, ES, P>(acc))) break;          \
  case C_CAST + FROM_DI * 4 + DI_I64: TFS_WIDE((do_cast<FROM_T, long long, ES, P>(acc))) break;
#define TFS_ALWAYS(stmt) { stmt; }

template <int ES, int P>
__global__ void __launch_bounds__(kThreads, 2)
eltwise_kernel(const __grid_constant__ EltProgram prog, const int64_t n) {
  using S = typename Slot<ES>::type;
  const int64_t n_vec = (n + P - 1) / P;
  constexpr int64_t tile_vecs = int64_t(kThreads) * kJ;
  const int64_t n_tiles = (n_vec + tile_vecs - 1) / tile_vecs;

  S acc[kJ][P];
  S x[kJ][P];
  S tmp[kMaxTmp][kJ][P];

  // The program is interpreted twice over in the source: for whole tiles (all but the block's last one) every vector is
  // complete, nvalid is the constant P, and the tail tests inside the loads, stores and divisions fold away.
  auto run_tile = [&](const int64_t tile, auto whole_tag) {
    constexpr bool WHOLE = decltype(whole_tag)::value;
    int64_t elem[kJ];
    int nvalid[kJ];
#pragma unroll
    for (int j = 0; j < kJ; ++j) {
      const int64_t v = tile * tile_vecs + int64_t(j) * kThreads + threadIdx.x;
      elem[j] = v * P;
      if constexpr (WHOLE) {
        nvalid[j] = P;
      } else {
        const int64_t left = n - elem[j];
        nvalid[j] = left >= P ? P : (left > 0 ? int(left) : 0);
      }
    }
    for (int pc = 0; pc < prog.n_instr; ++pc) {
      const Instr in = prog.ins[pc];
      switch (in.code) {
        case C_LOAD_BUF: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_vec<ES, P>(b, elem[j], nvalid[j], acc[j]);
          break;
        }
        case C_FETCH_BUF: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_vec<ES, P>(b, elem[j], nvalid[j], x[j]);
          break;
        }
        case C_LOAD_PER: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_periodic<ES, P>(b, elem[j], nvalid[j], acc[j]);
          break;
        }
        case C_FETCH_PER: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_periodic<ES, P>(b, elem[j], nvalid[j], x[j]);
          break;
        }
        case C_LOAD_ROW: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_rowbcast<ES, P>(b, elem[j], nvalid[j], acc[j]);
          break;
        }
        case C_FETCH_ROW: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) load_rowbcast<ES, P>(b, elem[j], nvalid[j], x[j]);
          break;
        }
        case C_LOAD_IMM:
#pragma unroll
          for (int j = 0; j < kJ; ++j)
#pragma unroll
            for (int e = 0; e < P; ++e) acc[j][e] = S(in.imm);
          break;
        case C_LOAD_TMP: copy_tile<ES, P>(acc, tmp[in.arg]); break;
        case C_FETCH_TMP: copy_tile<ES, P>(x, tmp[in.arg]); break;
        case C_FETCH_ACC: copy_tile<ES, P>(x, acc); break;
        case C_STTMP: copy_tile<ES, P>(tmp[in.arg], acc); break;
        case C_STORE: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) store_vec<ES, P>(b, elem[j], nvalid[j], acc[j]);
          break;
        }
        TFS_BIN_CASES(B_ADD)
        TFS_BIN_CASES(B_SUB)
        TFS_BIN_CASES(B_MUL)
        TFS_BIN_CASES(B_DIV)
        TFS_BIN_CASES(B_MIN)
        TFS_BIN_CASES(B_MAX)
        TFS_UN_CASES(U_NEG)
        TFS_UN_CASES(U_RELU)
        TFS_UN_CASES(U_SQUARE)
        TFS_UN_CASES(U_ABS)
        TFS_UN_CASES(U_SQRT)
        TFS_UN_CASES(U_RECIP)
        TFS_CAST_CASES(DI_F32, float, TFS_ALWAYS)
        TFS_CAST_CASES(DI_I32, int, TFS_ALWAYS)
        TFS_CAST_CASES(DI_F64, double, TFS_WIDE)
        TFS_CAST_CASES(DI_I64, long long, TFS_WIDE)
        default: break;
      }
    }
  };
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    if ((tile + 1) * tile_vecs * P <= n) run_tile(tile, std::true_type{});
    else run_tile(tile, std::false_type{});
  }
}

template <int ES, int P>
int launch_variant(const EltProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  // same for every device of a homogeneous box; initialised once, thread-safely
  static const int blocks_per_sm = [] {
    int b = 0;
    TFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, eltwise_kernel<ES, P>, kThreads, 0));
    return b > 0 ? b : 1;
  }();
  const int64_t n_vec = (n + P - 1) / P;
  const int64_t n_tiles = (n_vec + int64_t(kThreads) * kJ - 1) / (int64_t(kThreads) * kJ);
  int64_t grid = int64_t(dev.sm_count) * blocks_per_sm;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  eltwise_kernel<ES, P><<<dim3((unsigned)grid), dim3(kThreads), 0, stream>>>(prog, n);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace

int launch_eltwise(const EltProgram& prog, int64_t n, bool wide, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  bool aligned = true;
  for (int i = 0; i < prog.n_bufs; ++i)
    if (reinterpret_cast<uintptr_t>(prog.bufs[i].ptr) & 15) aligned = false;
  if (wide) return aligned ? launch_variant<8, 2>(prog, n, dev, stream) : launch_variant<8, 1>(prog, n, dev, stream);
  return aligned ? launch_variant<4, 4>(prog, n, dev, stream) : launch_variant<4, 1>(prog, n, dev, stream);
}

}  // namespace tfs
