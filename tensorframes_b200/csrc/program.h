// Structures shared between the host planner and the device kernels.
//
// K1 executes a *fused element-wise group* -- every Identity/Neg/Cast/Add/Mul/...
// node of the graph that iterates over the same block shape -- as ONE kernel.
// The group is compiled on the host into a short accumulator-machine program
// that the kernel interprets per register tile (see kernels_eltwise.cu).  This
// replaces the chain of separate Eigen kernels TF-CPU runs for the same nodes
// behind Session.run (reference: impl/DebugRowOps.scala:794), each of which
// reads and writes a full intermediate tensor.
#pragma once
#include <cstdint>

namespace tfs {

constexpr int kMaxInstr = 48;
constexpr int kMaxBufs = 12;
constexpr int kMaxTmp = 4;

// dtype index used on the device: 0 f32, 1 f64, 2 i32, 3 i64 (see common.h dtype_index).
enum : uint8_t { DI_F32 = 0, DI_F64 = 1, DI_I32 = 2, DI_I64 = 3 };

// machine opcodes
enum : uint8_t {
  M_LOAD = 0,   // acc <- src
  M_STORE = 1,  // bufs[arg] <- acc            (dt = dtype of acc)
  M_STTMP = 2,  // tmp[arg] <- acc
  M_UNARY = 3,  // acc <- un(sub, acc)         (sub = EltOp, dt)
  M_CAST = 4,   // acc <- cast<dt2>(acc as dt)
  M_BINARY = 5, // acc <- bin(sub, acc, src)   (flag SWAP: bin(sub, src, acc))
};

// operand kinds
enum : uint8_t {
  S_NONE = 0,
  S_BUF = 1,       // bufs[arg], same flat index as the output
  S_PERIODIC = 2,  // bufs[arg][flat % period]  (row-broadcast of a cell-shaped tensor, or a computed scalar)
  S_IMM = 3,       // scalar immediate (bits in imm, typed by dt)
  S_TMP = 4,       // tmp[arg]
  S_ACC = 5,       // the accumulator itself (x + x)
};

enum : uint8_t { F_SWAP = 1 };

struct Instr {
  uint8_t op;
  uint8_t sub;    // EltOp for M_UNARY / M_BINARY
  uint8_t dt;     // dtype index the op computes in
  uint8_t dt2;    // M_CAST: destination dtype index
  uint8_t src;    // operand kind
  uint8_t flags;
  uint16_t arg;   // buffer or tmp index
  uint64_t imm;
};

struct BufRef {
  void* ptr;
  int64_t period;  // S_PERIODIC: number of elements in the small tensor
  int32_t dt;      // dtype index of the buffer's elements
  int32_t pad;
};

struct EltProgram {
  int32_t n_instr;
  int32_t n_bufs;
  int32_t* err;    // device flag: set to 1 on integer division by zero
  Instr ins[kMaxInstr];
  BufRef bufs[kMaxBufs];
};

}  // namespace tfs
