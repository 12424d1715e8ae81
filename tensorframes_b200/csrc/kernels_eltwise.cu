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
// registers; the opcode switch is warp-uniform, so there is no divergence and
// its cost is amortised over kJ*P elements.  HBM-bound: the roofline is
// (bytes read + bytes written) / HBM bandwidth.
//
// Arithmetic follows the TF 1.1 CPU kernels: IEEE round-to-nearest for + - * /
// with no FMA contraction (explicit __dadd_rn/__dmul_rn...), Minimum/Maximum as
// Eigen's scalar (b < a ? b : a), Cast as C++ static_cast, integer arithmetic
// wrapping, integer division truncating toward zero and flagged when the divisor
// is zero (TF raises "Integer division by zero").
#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = 256;
constexpr int kJ = 4;

template <int ES> struct Slot;
template <> struct Slot<8> { using type = unsigned long long; };
template <> struct Slot<4> { using type = unsigned int; };

template <class T, class S>
__device__ __forceinline__ T from_slot(S s) {
  if constexpr (sizeof(T) == sizeof(S)) {
    return *reinterpret_cast<T*>(&s);
  } else {
    unsigned int lo = static_cast<unsigned int>(s);
    return *reinterpret_cast<T*>(&lo);
  }
}
template <class S, class T>
__device__ __forceinline__ S to_slot(T t) {
  if constexpr (sizeof(T) == sizeof(S)) {
    return *reinterpret_cast<S*>(&t);
  } else {
    return static_cast<S>(*reinterpret_cast<unsigned int*>(&t));
  }
}

template <class T> struct Arith;
template <> struct Arith<double> {
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b, bool, int*) { return __ddiv_rn(a, b); }
  static __device__ __forceinline__ double mn(double a, double b) { return (b < a) ? b : a; }
  static __device__ __forceinline__ double mx(double a, double b) { return (a < b) ? b : a; }
  static __device__ __forceinline__ double neg(double a) { return -a; }
  static __device__ __forceinline__ double abs(double a) { return fabs(a); }
  static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
  static __device__ __forceinline__ double recip(double a) { return __ddiv_rn(1.0, a); }
};
template <> struct Arith<float> {
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b, bool, int*) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float mn(float a, float b) { return (b < a) ? b : a; }
  static __device__ __forceinline__ float mx(float a, float b) { return (a < b) ? b : a; }
  static __device__ __forceinline__ float neg(float a) { return -a; }
  static __device__ __forceinline__ float abs(float a) { return fabsf(a); }
  static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
  static __device__ __forceinline__ float recip(float a) { return __fdiv_rn(1.0f, a); }
};
template <class I, class U> struct IntArith {
  static __device__ __forceinline__ I add(I a, I b) { return I(U(a) + U(b)); }
  static __device__ __forceinline__ I sub(I a, I b) { return I(U(a) - U(b)); }
  static __device__ __forceinline__ I mul(I a, I b) { return I(U(a) * U(b)); }
  static __device__ __forceinline__ I div(I a, I b, bool valid, int* err) {
    if (b == 0) {
      if (valid) *err = 1;
      return 0;
    }
    if (b == I(-1)) return I(U(0) - U(a));  // avoids the INT_MIN / -1 trap value
    return a / b;
  }
  static __device__ __forceinline__ I mn(I a, I b) { return (b < a) ? b : a; }
  static __device__ __forceinline__ I mx(I a, I b) { return (a < b) ? b : a; }
  static __device__ __forceinline__ I neg(I a) { return I(U(0) - U(a)); }
  static __device__ __forceinline__ I abs(I a) { return a < 0 ? I(U(0) - U(a)) : a; }
  static __device__ __forceinline__ I sqrt(I a) { return a; }   // rejected at import
  static __device__ __forceinline__ I recip(I a) { return a; }  // rejected at import
};
template <> struct Arith<int> : IntArith<int, unsigned int> {};
template <> struct Arith<long long> : IntArith<long long, unsigned long long> {};

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

template <int ES, int P>
__device__ __forceinline__ void load_periodic(const BufRef& b, int64_t elem, int nvalid,
                                              typename Slot<ES>::type (&x)[P]) {
  using S = typename Slot<ES>::type;
#pragma unroll
  for (int e = 0; e < P; ++e) {
    S v = 0;
    if (e < nvalid) {
      int64_t i = (elem + e) % b.period;
      if (ES == 8 && (b.dt & 1)) v = S(__ldg(static_cast<const unsigned long long*>(b.ptr) + i));
      else v = S(__ldg(static_cast<const unsigned int*>(b.ptr) + i));
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

// ---- typed application over the register tile -------------------------------
template <class T, int ES, int P>
__device__ __forceinline__ void apply_unary(int sub, typename Slot<ES>::type (&acc)[kJ][P]) {
  using S = typename Slot<ES>::type;
#define TFS_UN(expr)                                              \
  _Pragma("unroll") for (int j = 0; j < kJ; ++j)                  \
  _Pragma("unroll") for (int e = 0; e < P; ++e) {                 \
    T a = from_slot<T, S>(acc[j][e]);                             \
    acc[j][e] = to_slot<S, T>(expr);                              \
  }
  switch (sub) {
    case E_NEG: TFS_UN(Arith<T>::neg(a)) break;
    case E_RELU: TFS_UN(Arith<T>::mx(a, T(0))) break;
    case E_SQUARE: TFS_UN(Arith<T>::mul(a, a)) break;
    case E_ABS: TFS_UN(Arith<T>::abs(a)) break;
    case E_SQRT: TFS_UN(Arith<T>::sqrt(a)) break;
    case E_RECIP: TFS_UN(Arith<T>::recip(a)) break;
    default: break;  // E_IDENTITY
  }
#undef TFS_UN
}

template <class T, int ES, int P>
__device__ __forceinline__ void apply_binary(int sub, bool swap, typename Slot<ES>::type (&acc)[kJ][P],
                                             const typename Slot<ES>::type (&x)[kJ][P],
                                             const int (&nvalid)[kJ], int* err) {
  using S = typename Slot<ES>::type;
#define TFS_BIN_LOOP(A, B, expr)                                  \
  _Pragma("unroll") for (int j = 0; j < kJ; ++j)                  \
  _Pragma("unroll") for (int e = 0; e < P; ++e) {                 \
    T a = from_slot<T, S>(A[j][e]);                               \
    T b = from_slot<T, S>(B[j][e]);                               \
    acc[j][e] = to_slot<S, T>(expr);                              \
  }
  // the operand order is warp-uniform: branch once instead of selecting per element
#define TFS_BIN(expr)                                             \
  if (swap) { TFS_BIN_LOOP(x, acc, expr) } else { TFS_BIN_LOOP(acc, x, expr) }
  switch (sub) {
    case E_ADD: TFS_BIN(Arith<T>::add(a, b)) break;
    case E_SUB: TFS_BIN(Arith<T>::sub(a, b)) break;
    case E_MUL: TFS_BIN(Arith<T>::mul(a, b)) break;
    case E_DIV: TFS_BIN(Arith<T>::div(a, b, e < nvalid[j], err)) break;
    case E_MIN: TFS_BIN(Arith<T>::mn(a, b)) break;
    case E_MAX: TFS_BIN(Arith<T>::mx(a, b)) break;
    default: break;
  }
#undef TFS_BIN_LOOP
#undef TFS_BIN
}

template <class From, class To, int ES, int P>
__device__ __forceinline__ void apply_cast(typename Slot<ES>::type (&acc)[kJ][P]) {
  using S = typename Slot<ES>::type;
#pragma unroll
  for (int j = 0; j < kJ; ++j)
#pragma unroll
    for (int e = 0; e < P; ++e)
      acc[j][e] = to_slot<S, To>(static_cast<To>(from_slot<From, S>(acc[j][e])));
}

template <class From, int ES, int P>
__device__ __forceinline__ void cast_from(int dt2, typename Slot<ES>::type (&acc)[kJ][P]) {
  switch (dt2) {
    case DI_F32: apply_cast<From, float, ES, P>(acc); break;
    case DI_I32: apply_cast<From, int, ES, P>(acc); break;
    case DI_F64: if constexpr (ES == 8) apply_cast<From, double, ES, P>(acc); break;
    case DI_I64: if constexpr (ES == 8) apply_cast<From, long long, ES, P>(acc); break;
  }
}

template <int ES, int P>
__global__ void __launch_bounds__(kThreads, 3)
eltwise_kernel(const __grid_constant__ EltProgram prog, const int64_t n) {
  using S = typename Slot<ES>::type;
  const int64_t n_vec = (n + P - 1) / P;
  constexpr int64_t tile_vecs = int64_t(kThreads) * kJ;
  const int64_t n_tiles = (n_vec + tile_vecs - 1) / tile_vecs;

  S acc[kJ][P];
  S x[kJ][P];
  S tmp[kMaxTmp][kJ][P];

  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    int64_t elem[kJ];
    int nvalid[kJ];
#pragma unroll
    for (int j = 0; j < kJ; ++j) {
      const int64_t v = tile * tile_vecs + int64_t(j) * kThreads + threadIdx.x;
      elem[j] = v * P;
      const int64_t left = n - elem[j];
      nvalid[j] = left >= P ? P : (left > 0 ? int(left) : 0);
    }
    for (int pc = 0; pc < prog.n_instr; ++pc) {
      const Instr in = prog.ins[pc];
      // operand fetch into x (M_LOAD then moves it to acc: register moves only, and it
      // keeps both tiles statically indexed so they stay in registers)
      if (in.op == M_LOAD || in.op == M_BINARY) {
        S(&dst)[kJ][P] = x;
        switch (in.src) {
          case S_BUF: {
            const BufRef b = prog.bufs[in.arg];
#pragma unroll
            for (int j = 0; j < kJ; ++j) load_vec<ES, P>(b, elem[j], nvalid[j], dst[j]);
            break;
          }
          case S_PERIODIC: {
            const BufRef b = prog.bufs[in.arg];
#pragma unroll
            for (int j = 0; j < kJ; ++j) load_periodic<ES, P>(b, elem[j], nvalid[j], dst[j]);
            break;
          }
          case S_IMM:
#pragma unroll
            for (int j = 0; j < kJ; ++j)
#pragma unroll
              for (int e = 0; e < P; ++e) dst[j][e] = S(in.imm);
            break;
          case S_TMP:
#pragma unroll
            for (int j = 0; j < kJ; ++j)
#pragma unroll
              for (int e = 0; e < P; ++e) dst[j][e] = tmp[in.arg][j][e];
            break;
          case S_ACC:
#pragma unroll
            for (int j = 0; j < kJ; ++j)
#pragma unroll
              for (int e = 0; e < P; ++e) dst[j][e] = acc[j][e];
            break;
          default: break;
        }
      }
      switch (in.op) {
        case M_LOAD:
#pragma unroll
          for (int j = 0; j < kJ; ++j)
#pragma unroll
            for (int e = 0; e < P; ++e) acc[j][e] = x[j][e];
          break;
        case M_STORE: {
          const BufRef b = prog.bufs[in.arg];
#pragma unroll
          for (int j = 0; j < kJ; ++j) store_vec<ES, P>(b, elem[j], nvalid[j], acc[j]);
          break;
        }
        case M_STTMP:
#pragma unroll
          for (int j = 0; j < kJ; ++j)
#pragma unroll
            for (int e = 0; e < P; ++e) tmp[in.arg][j][e] = acc[j][e];
          break;
        case M_UNARY:
          switch (in.dt) {
            case DI_F32: apply_unary<float, ES, P>(in.sub, acc); break;
            case DI_I32: apply_unary<int, ES, P>(in.sub, acc); break;
            case DI_F64: if constexpr (ES == 8) apply_unary<double, ES, P>(in.sub, acc); break;
            case DI_I64: if constexpr (ES == 8) apply_unary<long long, ES, P>(in.sub, acc); break;
          }
          break;
        case M_CAST:
          switch (in.dt) {
            case DI_F32: cast_from<float, ES, P>(in.dt2, acc); break;
            case DI_I32: cast_from<int, ES, P>(in.dt2, acc); break;
            case DI_F64: if constexpr (ES == 8) cast_from<double, ES, P>(in.dt2, acc); break;
            case DI_I64: if constexpr (ES == 8) cast_from<long long, ES, P>(in.dt2, acc); break;
          }
          break;
        case M_BINARY: {
          const bool swap = (in.flags & F_SWAP) != 0;
          switch (in.dt) {
            case DI_F32: apply_binary<float, ES, P>(in.sub, swap, acc, x, nvalid, prog.err); break;
            case DI_I32: apply_binary<int, ES, P>(in.sub, swap, acc, x, nvalid, prog.err); break;
            case DI_F64: if constexpr (ES == 8) apply_binary<double, ES, P>(in.sub, swap, acc, x, nvalid, prog.err); break;
            case DI_I64: if constexpr (ES == 8) apply_binary<long long, ES, P>(in.sub, swap, acc, x, nvalid, prog.err); break;
          }
          break;
        }
        default: break;
      }
    }
  }
}

template <int ES, int P>
int launch_variant(const EltProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  static int blocks_per_sm = 0;  // same for every device of a homogeneous box
  if (blocks_per_sm == 0) {
    int b = 0;
    TFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, eltwise_kernel<ES, P>, kThreads, 0));
    blocks_per_sm = b > 0 ? b : 1;
  }
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
