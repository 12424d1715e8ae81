// Structures shared between the host planner and the device kernels.
//
// K1 executes a *fused element-wise group* -- every Identity/Neg/Cast/Add/Mul/...
// node of the graph that iterates over the same block shape -- as ONE kernel.
// The group is compiled on the host into a short accumulator-machine program
// that the kernel interprets per register tile (see kernels_eltwise.cu).  This
// replaces the chain of separate Eigen kernels TF-CPU runs for the same nodes
// behind Session.run (reference: impl/DebugRowOps.scala:794), each of which
// reads and writes a full intermediate tensor.
//
// Every instruction carries ONE flat opcode in which the operation, the dtype,
// the operand kind and the operand order are already resolved, so the device
// decodes an instruction with a single warp-uniform jump.
#pragma once
#include <cstdint>

namespace tfs {

constexpr int kMaxInstr = 48;
constexpr int kMaxBufs = 12;
constexpr int kMaxTmp = 4;

// dtype index used on the device: 0 f32, 1 f64, 2 i32, 3 i64 (see common.h dtype_index).
enum : uint8_t { DI_F32 = 0, DI_F64 = 1, DI_I32 = 2, DI_I64 = 3 };

// binary sub-ops, in opcode order
enum : uint8_t { B_ADD = 0, B_SUB = 1, B_MUL = 2, B_DIV = 3, B_MIN = 4, B_MAX = 5, B_COUNT = 6 };
// unary sub-ops
enum : uint8_t { U_NEG = 0, U_RELU = 1, U_SQUARE = 2, U_ABS = 3, U_SQRT = 4, U_RECIP = 5, U_COUNT = 6 };

// flat opcodes
enum : uint8_t {
  C_LOAD_BUF = 1,    // acc <- bufs[arg][flat]
  C_LOAD_PER = 2,    // acc <- bufs[arg][flat % period]
  C_LOAD_IMM = 3,    // acc <- imm
  C_LOAD_TMP = 4,    // acc <- tmp[arg]
  C_FETCH_BUF = 5,   // x <- bufs[arg][flat]            (operand of a following C_BIN_X)
  C_FETCH_PER = 6,   // x <- bufs[arg][flat % period]
  C_FETCH_TMP = 7,   // x <- tmp[arg]
  C_FETCH_ACC = 8,   // x <- acc                         (x + x)
  C_STORE = 9,       // bufs[arg][flat] <- acc
  C_STTMP = 10,      // tmp[arg] <- acc
  // acc <- op(acc, x)   (swap: op(x, acc));   code = C_BIN_X + sub * 8 + dt * 2 + swap
  C_BIN_X = 11,
  // acc <- op(acc, imm) (swap: op(imm, acc)); code = C_BIN_IMM + sub * 8 + dt * 2 + swap
  C_BIN_IMM = 59,
  // acc <- un(acc);                            code = C_UNARY + sub * 4 + dt
  C_UNARY = 107,
  // acc <- cast<to>(acc as from);              code = C_CAST + from * 4 + to
  C_CAST = 131,
  // row mode: a per-row operand smaller than the cell (x / reduce_sum(x) under mapRows).  Element `flat` of the
  // [n, cell...] block reads bufs[arg][(flat / rowspan) * period + flat % period]: rowspan = elements per row of the
  // block, period = elements per row of the operand (1 for a per-row scalar).
  C_LOAD_ROW = 147,
  C_FETCH_ROW = 148,
};

struct Instr {
  uint8_t code;
  uint8_t arg;      // buffer or tmp index
  uint8_t pad[6];
  uint64_t imm;     // immediate operand bits (low 32 bits for 4-byte dtypes)
};

struct BufRef {
  void* ptr;
  int64_t period;  // C_*_PER: number of elements in the small tensor; C_*_ROW: elements per row of the operand
  int32_t dt;      // dtype index of the buffer's elements
  int32_t pad;
  int64_t rowspan; // C_*_ROW: elements per row of the block being iterated
};

struct EltProgram {
  int32_t n_instr;
  int32_t n_bufs;
  int32_t* err;    // device flag: set to 1 on integer division by zero
  Instr ins[kMaxInstr];
  BufRef bufs[kMaxBufs];
};

// ---- linear-chain fast path (kernels_chain.cu) --------------------------------------------
constexpr int kChainSlots = 8;
enum : uint8_t {
  CH_NOP = 0, CH_ADD, CH_SUB, CH_RSUB, CH_MUL, CH_DIV, CH_RDIV, CH_MIN, CH_RMIN, CH_MAX, CH_RMAX,
  CH_NEG, CH_RELU, CH_SQUARE, CH_ABS, CH_SQRT, CH_RECIP,
  CH_LOAD,   // acc = operand (general variant only)
};
// where a slot's second operand comes from
enum : uint8_t {
  CK_IMM = 0,   // scalar constant in `imm`
  CK_COL,       // another column: operand_buf at the same flat index
  CK_ROWVEC,    // operand_buf holds `period` elements broadcast along the rows (x * w + b with row vectors); the
                // period divides the tile span and is a multiple of the vector width (make_chain checks)
  CK_TMP,       // the saved tile
  CK_SELF,      // the accumulator itself (x + x)
};
struct ChainSlot {
  const void* operand_buf;  // CK_COL / CK_ROWVEC
  void* store_buf;          // != null: after the op the accumulator is also written to this column (a second fetch)
  uint64_t imm;
  uint8_t op;
  uint8_t kind;
  uint8_t save;             // after the op: saved tile = accumulator
  uint8_t pad;
  uint32_t period;
};
// geometry of the chain kernel's general variant, shared with make_chain's eligibility test
constexpr int kChainThreads = 256;
constexpr int kChainVecsPerThread = 4;
struct ChainProgram {
  const void* in;
  void* out;
  int32_t* err;
  int32_t n_slots;
  uint8_t dt;       // dtype index the chain computes in
  uint8_t out_dt;   // dtype index of the output (differs when the chain ends in a Cast)
  uint8_t in_dt;    // dtype index of the input column (differs when the chain starts with a Cast; then out_dt == dt)
  uint8_t pad[1];
  ChainSlot slots[kChainSlots];
};

}  // namespace tfs
