// K1 fast path -- straight-line kernel for *linear chains* (sm_100a).
//
// Most map graphs over DataFrame columns are a chain: one input column, then a few unary
// ops and binary ops whose other operand is a scalar constant or another column of the same
// dtype, optionally a final Cast, one output column (README `x + 3`, BASELINE config 3
// `cast_f32(max(x*a+b, 0))`, `a + b`).  For those the general interpreter's per-instruction
// dispatch loop (kernels_eltwise.cu) is replaced by kSlots unrolled stages, each a
// warp-uniform switch over the op applied in place to the register tile: no interpreter loop,
// no loop-carried copies, ~12 instructions per element, so the kernel stays HBM-bound.
// Arithmetic is the same Arith<T> set as the interpreter (separately rounded IEEE ops).
#include "kernels.h"

namespace tfs {
namespace {

constexpr int kThreads = kChainThreads;

#include "arith.inc"

template <class T, int P> struct VecOf;
template <> struct VecOf<double, 2> { using type = double2; };
template <> struct VecOf<long long, 2> { using type = longlong2; };
template <> struct VecOf<float, 4> { using type = float4; };
template <> struct VecOf<int, 4> { using type = int4; };
template <> struct VecOf<float, 2> { using type = float2; };
template <> struct VecOf<int, 2> { using type = int2; };

template <class T, int P>
__device__ __forceinline__ void ld_tile(const T* __restrict__ base, int64_t elem, int nvalid, T (&v)[P]) {
  if (nvalid == P) {
    using V = typename VecOf<T, P>::type;
    const V t = __ldcs(reinterpret_cast<const V*>(base + elem));
    const T* tv = reinterpret_cast<const T*>(&t);
#pragma unroll
    for (int e = 0; e < P; ++e) v[e] = tv[e];
  } else {
#pragma unroll
    for (int e = 0; e < P; ++e) v[e] = e < nvalid ? __ldcs(base + elem + e) : T(1);
  }
}

template <class T, int P>
__device__ __forceinline__ void st_tile(T* __restrict__ base, int64_t elem, int nvalid, const T (&v)[P]) {
  if (nvalid == P) {
    if constexpr (P * sizeof(T) == 32) {   // widened output: two 128-bit stores
      using V = typename VecOf<T, P / 2>::type;
      V a, b;
      T* av = reinterpret_cast<T*>(&a);
      T* bv = reinterpret_cast<T*>(&b);
#pragma unroll
      for (int e = 0; e < P / 2; ++e) { av[e] = v[e]; bv[e] = v[P / 2 + e]; }
      __stcs(reinterpret_cast<V*>(base + elem), a);
      __stcs(reinterpret_cast<V*>(base + elem) + 1, b);
    } else {
      using V = typename VecOf<T, P>::type;
      V a;
      T* av = reinterpret_cast<T*>(&a);
#pragma unroll
      for (int e = 0; e < P; ++e) av[e] = v[e];
      __stcs(reinterpret_cast<V*>(base + elem), a);
    }
  } else {
#pragma unroll
    for (int e = 0; e < P; ++e)
      if (e < nvalid) __stcs(base + elem + e, v[e]);
  }
}

// one op applied in place to the register tile; OPERAND(j, e) yields the second operand
#define TFS_SLOT_SWITCH(OPERAND)                                                        \
  switch (s.op) {                                                                       \
    case CH_ADD: TFS_SLOT(Arith<T>::add(a, OPERAND))                                    \
    case CH_SUB: TFS_SLOT(Arith<T>::sub(a, OPERAND))                                    \
    case CH_RSUB: TFS_SLOT(Arith<T>::sub(OPERAND, a))                                   \
    case CH_MUL: TFS_SLOT(Arith<T>::mul(a, OPERAND))                                    \
    case CH_DIV: TFS_SLOT(Arith<T>::div(a, OPERAND, e < nvalid[j], err))                \
    case CH_RDIV: TFS_SLOT(Arith<T>::div(OPERAND, a, e < nvalid[j], err))               \
    case CH_MIN: TFS_SLOT(Arith<T>::mn(a, OPERAND))                                     \
    case CH_RMIN: TFS_SLOT(Arith<T>::mn(OPERAND, a))                                    \
    case CH_MAX: TFS_SLOT(Arith<T>::mx(a, OPERAND))                                     \
    case CH_RMAX: TFS_SLOT(Arith<T>::mx(OPERAND, a))                                    \
    case CH_NEG: TFS_SLOT(Arith<T>::neg(a))                                             \
    case CH_RELU: TFS_SLOT(Arith<T>::mx(a, T(0)))                                       \
    case CH_SQUARE: TFS_SLOT(Arith<T>::mul(a, a))                                       \
    case CH_ABS: TFS_SLOT(Arith<T>::abs(a))                                             \
    case CH_SQRT: TFS_SLOT(Arith<T>::sqrt(a))                                           \
    case CH_RECIP: TFS_SLOT(Arith<T>::recip(a))                                         \
    default: break;                                                                     \
  }
#define TFS_SLOT(expr)                                \
  _Pragma("unroll") for (int j = 0; j < kJ; ++j)      \
  _Pragma("unroll") for (int e = 0; e < P; ++e) {     \
    const T a = acc[j][e];                            \
    acc[j][e] = (expr);                               \
  }                                                   \
  break;

template <class T, int P, int kJ, bool BUFOPS>
__device__ __forceinline__ void apply_slot(const ChainSlot& s, T (&acc)[kJ][P], const int64_t (&elem)[kJ],
                                           const int (&nvalid)[kJ], int* err) {
  if (BUFOPS && s.operand_buf != nullptr) {   // second operand: another column, same flat index ...
    T b[kJ][P];
    const T* base = static_cast<const T*>(s.operand_buf);
    if (s.period != 0u) {
      // ... or a row vector broadcast down the block.  The period divides the tile span, so the position inside
      // the vector depends only on the thread's place in the tile; it is a multiple of P, so a vector never wraps.
#pragma unroll
      for (int j = 0; j < kJ; ++j) {
        const unsigned int at = (unsigned int)((j * kThreads + threadIdx.x) * P) % s.period;
        using V = typename VecOf<T, P>::type;
        const V t = __ldg(reinterpret_cast<const V*>(base + at));
        const T* tv = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int e = 0; e < P; ++e) b[j][e] = tv[e];
      }
    } else {
#pragma unroll
      for (int j = 0; j < kJ; ++j) ld_tile<T, P>(base, elem[j], nvalid[j], b[j]);
    }
    TFS_SLOT_SWITCH(b[j][e])
  } else {                          // second operand: a scalar constant (or none), kept in one register
    T imm;
    memcpy(&imm, &s.imm, sizeof(T));
    TFS_SLOT_SWITCH(imm)
  }
}
#undef TFS_SLOT
#undef TFS_SLOT_SWITCH

// BUFOPS = false: no slot reads a second column, so the operand tile is not needed and the thread carries
// twice as many vectors (8 independent 128-bit loads in flight instead of 4).
template <class T, class OutT, bool BUFOPS>
__global__ void __launch_bounds__(kThreads, 3)
chain_kernel(const __grid_constant__ ChainProgram prog, const int64_t n) {
  constexpr int P = 16 / int(sizeof(T));
  constexpr int kJ = BUFOPS ? kChainVecsPerThread : 2 * kChainVecsPerThread;
  const int64_t n_vec = (n + P - 1) / P;
  constexpr int64_t tile_vecs = int64_t(kThreads) * kJ;
  const int64_t n_tiles = (n_vec + tile_vecs - 1) / tile_vecs;
  const T* __restrict__ in = static_cast<const T*>(prog.in);
  OutT* __restrict__ out = static_cast<OutT*>(prog.out);
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    int64_t elem[kJ];
    int nvalid[kJ];
    T acc[kJ][P];
#pragma unroll
    for (int j = 0; j < kJ; ++j) {
      const int64_t v = tile * tile_vecs + int64_t(j) * kThreads + threadIdx.x;
      elem[j] = v * P;
      const int64_t left = n - elem[j];
      nvalid[j] = left >= P ? P : (left > 0 ? int(left) : 0);
    }
#pragma unroll
    for (int j = 0; j < kJ; ++j) ld_tile<T, P>(in, elem[j], nvalid[j], acc[j]);
#pragma unroll
    for (int s = 0; s < kChainSlots; ++s)
      if (s < prog.n_slots) apply_slot<T, P, kJ, BUFOPS>(prog.slots[s], acc, elem, nvalid, prog.err);
#pragma unroll
    for (int j = 0; j < kJ; ++j) {
      OutT o[P];
#pragma unroll
      for (int e = 0; e < P; ++e) o[e] = static_cast<OutT>(acc[j][e]);
      st_tile<OutT, P>(out, elem[j], nvalid[j], o);
    }
  }
}

template <class T, class OutT, bool BUFOPS>
int launch_chain_v(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  static int blocks_per_sm = 0;
  if (blocks_per_sm == 0) {
    int b = 0;
    TFS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, chain_kernel<T, OutT, BUFOPS>, kThreads, 0));
    blocks_per_sm = b > 0 ? b : 1;
  }
  constexpr int P = 16 / int(sizeof(T));
  constexpr int kJ = BUFOPS ? kChainVecsPerThread : 2 * kChainVecsPerThread;
  const int64_t n_vec = (n + P - 1) / P;
  const int64_t n_tiles = (n_vec + int64_t(kThreads) * kJ - 1) / (int64_t(kThreads) * kJ);
  int64_t grid = std::min<int64_t>(int64_t(dev.sm_count) * blocks_per_sm, n_tiles);
  if (grid < 1) grid = 1;
  chain_kernel<T, OutT, BUFOPS><<<(unsigned)grid, kThreads, 0, stream>>>(prog, n);
  TFS_CUDA(cudaGetLastError());
  return 1;
}

template <class T, class OutT>
int launch_chain_t(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  bool bufops = false;
  for (int s = 0; s < prog.n_slots; ++s) bufops = bufops || prog.slots[s].operand_buf != nullptr;
  return bufops ? launch_chain_v<T, OutT, true>(prog, n, dev, stream) : launch_chain_v<T, OutT, false>(prog, n, dev, stream);
}

template <class T>
int launch_chain_out(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  switch (prog.out_dt) {
    case DI_F32: return launch_chain_t<T, float>(prog, n, dev, stream);
    case DI_F64: return launch_chain_t<T, double>(prog, n, dev, stream);
    case DI_I32: return launch_chain_t<T, int>(prog, n, dev, stream);
    default: return launch_chain_t<T, long long>(prog, n, dev, stream);
  }
}

}  // namespace

int launch_chain(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  switch (prog.dt) {
    case DI_F32: return launch_chain_out<float>(prog, n, dev, stream);
    case DI_F64: return launch_chain_out<double>(prog, n, dev, stream);
    case DI_I32: return launch_chain_out<int>(prog, n, dev, stream);
    default: return launch_chain_out<long long>(prog, n, dev, stream);
  }
}

}  // namespace tfs
