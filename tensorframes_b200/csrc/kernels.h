// Launch wrappers for the hand-written sm_100a kernels. All launches are
// asynchronous on `stream`; the callers own error checking of the stream.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <mutex>

#include "common.h"
#include "graph.h"
#include "program.h"

namespace tfs {

inline void cuda_check(cudaError_t e, const char* what) {
  if (e != cudaSuccess)
    fail(TFS_ERR_CUDA, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define TFS_CUDA(call) ::tfs::cuda_check((call), #call)

// Function attributes (the opt-in to more than 48 KB of dynamic shared memory) belong to a device: a launcher sets them
// the first time it runs on each device.  One instance per call site.  A thread that finds the set-up under way waits
// for it: returning early would let its launch go out before the attribute is in place and be rejected.
struct PerDeviceOnce {
  std::mutex mu;
  uint64_t done = 0;
  template <class F>
  void run(int device, F&& setup) {
    const uint64_t bit = uint64_t(1) << (device & 63);
    std::lock_guard<std::mutex> lock(mu);
    if (done & bit) return;
    setup();
    done |= bit;        // only after a set-up that did not throw
  }
};

// Where a caller's pointer lives: kHostMemory (pageable or pinned host memory, or unknown to CUDA), kManagedMemory,
// or the ordinal of the device that owns the allocation.
constexpr int kHostMemory = -1, kManagedMemory = -2;
inline int pointer_location(const void* p) {
  if (p == nullptr) return kHostMemory;
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return kHostMemory;
  }
  if (a.type == cudaMemoryTypeManaged) return kManagedMemory;
  return a.type == cudaMemoryTypeDevice ? a.device : kHostMemory;
}
// true: device memory of `device` (or managed memory); false: host memory.  Memory of ANOTHER device is refused: the
// kernels (and the tensor maps of the MatMul path) address this GPU's memory only, the library moves nothing between
// devices behind the caller's back.
inline bool on_device_or_refuse(const void* p, int device, const std::string& what) {
  const int loc = pointer_location(p);
  if (loc == kHostMemory) return false;
  TFS_CHECK(loc == kManagedMemory || loc == device, TFS_ERR_INVALID,
            what + " lives on device " + std::to_string(loc) + " but the call runs on device " + std::to_string(device) +
                ": pass memory of that device, or host memory");
  return true;
}

struct DeviceInfo {
  int device = 0;
  int sm_count = 148;
};

// K1: fused element-wise group over n elements. `wide` = the program touches a
// 64-bit dtype. Returns the number of kernels launched.
int launch_eltwise(const EltProgram& prog, int64_t n, bool wide, const DeviceInfo& dev, cudaStream_t stream);

// K1 fast path: a linear chain (one input column, unary / binary-with-leaf ops, optional final cast).
int launch_chain(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream);

// K2: reduction of in[outer, inner] along axis 0 into out[inner].
// `scratch` must hold reduce_scratch_bytes(...) bytes; its first 256 bytes are a
// counter block that must be zero before the first launch (kernels re-zero it).
size_t reduce_scratch_bytes(int dtype, int64_t outer, int64_t inner, const DeviceInfo& dev);
int launch_reduce_axis0(int dtype, int red_op, const void* in, int64_t outer, int64_t inner, void* out,
                        void* scratch, const DeviceInfo& dev, cudaStream_t stream);

// Row reduction: in[outer, inner] along axis 1 into out[outer].
int launch_reduce_rows(int dtype, int red_op, const void* in, int64_t outer, int64_t inner, void* out,
                       const DeviceInfo& dev, cudaStream_t stream);

// Middle axes: in[outer, len, inner] along len into out[outer, inner].
int launch_reduce_mid(int dtype, int red_op, const void* in, int64_t outer, int64_t len, int64_t inner, void* out,
                      const DeviceInfo& dev, cudaStream_t stream);

// K4: out[M, N] = op(a) * op(b) (+ bias[N]); float32 or float64 (dtype). The SIMT kernel takes any shape.
int launch_matmul_simt(int dtype, const void* a, const void* b, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                       bool ta, bool tb, const DeviceInfo& dev, cudaStream_t stream);

// Shape-changing steps (kernels_shape.cu). Tile copies elem_size-byte elements (4 or 8).
int launch_tile(int elem_size, const void* in, const Shape& in_shape, void* out, const Shape& out_shape, const DeviceInfo& dev,
                cudaStream_t stream);
// out[o, i] = arg extreme over a of in[o, a, i]; ties -> smallest index. out_dtype: TFS_INT32 or TFS_INT64.
int launch_argminmax(int dtype, int out_dtype, bool is_min, const void* in, int64_t outer, int64_t len, int64_t inner, void* out,
                     const DeviceInfo& dev, cudaStream_t stream);
int launch_widen_ids(const int* in, int64_t n, int64_t* out, const DeviceInfo& dev, cudaStream_t stream);
// out[keys[g], :] = sums[g, :] for g < groups; keys outside [0, segments) set *err to 2.
int launch_segment_scatter(int dtype, const void* sums, const int64_t* keys, int64_t groups, int64_t cell, int64_t segments,
                           void* out, int32_t* err, const DeviceInfo& dev, cudaStream_t stream);

// tcgen05 path for x[M, K] * w[K, N] (+ bias): K % 32 == 0, N % 16 == 0, N <= 256, x not transposed.
// w is passed pre-split into tf32-exact hi / lo parts laid out [N, K] (launch_split_weights).
bool matmul_tc_supported(const void* a, const void* out, int64_t M, int64_t N, int64_t K, bool ta);
size_t matmul_tc_weight_bytes(int64_t N, int64_t K);
int launch_split_weights(const float* w, bool transposed, int64_t K, int64_t N, float* hi, float* lo, cudaStream_t stream);
int launch_matmul_tc(const float* x, const float* w_hi, const float* w_lo, const float* bias, float* out, int64_t M,
                     int64_t N, int64_t K, const DeviceInfo& dev, cudaStream_t stream);

// Stable radix partition passes (kernels_radix.cu): n (key, value) pairs ordered by `bits` (<= 11) key bits from `shift`
// up; keys are 4 or 8 bytes (xor_mask is applied before the digit is taken: the sign bit for signed order), values 4 or
// 8 bytes, or -- vals_in null, 4-byte values -- the element's own index.  keys_out may be null.  `scratch` holds
// radix_scratch_bytes().  launch_radix_sort chains ceil(bits / 11) passes from `keys_in / vals_in` through the two
// buffer pairs and returns the index of the pair that holds the result (the last pass writes keys only if want_keys).
size_t radix_scratch_bytes(const DeviceInfo& dev);
int radix_passes(int64_t n, int bits);
int launch_radix_pass(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_out, void* vals_out,
                      int64_t n, uint64_t xor_mask, int shift, int bits, void* scratch, const DeviceInfo& dev, cudaStream_t stream);
int launch_radix_sort(int key_bytes, int val_bytes, const void* keys_in, const void* vals_in, void* keys_buf[2], void* vals_buf[2],
                      int64_t n, uint64_t xor_mask, int bits, bool want_keys, void* scratch, const DeviceInfo& dev, cudaStream_t stream,
                      int* launches);

// rows ordered by slot (stable): slots[n] -> rows_out[n]; pairs_a / pairs_b are 8-byte-per-row scratch columns
// seg_lo / seg_hi (both or neither): per slot, the first position and one past the last position of its rows in rows_out
// (pre-set to 0xffffffff / 0 by the caller, updated with atomicMin / atomicMax by the last pass)
int launch_radix_rows_by_slot(const unsigned int* slots, int64_t n, int bits, unsigned long long* pairs_a, unsigned long long* pairs_b,
                              unsigned int* rows_out, unsigned int* seg_lo, unsigned int* seg_hi, void* scratch, const DeviceInfo& dev,
                              cudaStream_t stream);

// K3: keyed block reduction. Scratch layout for n_rows rows and at most max_groups groups.
struct AggLayout {
  uint64_t table_slots;
  size_t off_header, off_slot_key, off_counts, off_offsets, off_group_slot, off_tile_sums, off_perm, off_sort_temp;
  size_t off_ends;          // one past the last position of every slot's segment (rows_counted == false)
  // true: the insert pass counts the rows of every slot (atomics) and the table scan turns the counts into segment
  // offsets.  false (set by the caller before launch_agg_build; needs a row partition of two passes or more and no
  // carried fetch): nothing is counted, the last partition pass records where every slot's segment begins and ends.
  bool rows_counted;
  bool bounds_eligible;     // the row partition has two passes or more (set by agg_layout)
  size_t off_group_slot_tab, off_keys_tab;   // groups in table order, before they are sorted by key
  size_t off_row_slot[3];   // [0] the slot column written by the insert pass, [1] / [2] the slot column between passes
  size_t off_row_id[2];     // row-index columns of the partition (ping-pong)
  size_t off_pairs[2];      // packed (slot, row) columns between the passes of the row partition, 8 bytes per row
  size_t off_group_tmp_keys, off_group_tmp_slot;   // second (key, slot) column pair of the group sort
  size_t sort_temp_bytes;
  size_t bytes;
};
AggLayout agg_layout(int64_t n_rows, int64_t max_groups, const DeviceInfo& dev);
// passes 1-2: hash-insert the keys (each row records its slot, each slot counts its rows), then scan the
// table: segment offsets, dense group ids and keys in table order. header (first 16 bytes of scratch):
// [0] = occupied slots, [1] = overflow flag (more than max_groups distinct keys), [2] = number of groups.
int launch_agg_build(const int64_t* keys, int64_t n_rows, int64_t max_groups, char* scratch, const AggLayout& L,
                     const DeviceInfo& dev, cudaStream_t stream);
// pass 3: groups sorted by key -> out_keys_dev[0..n_groups) ascending (the table order depends on probe races);
// stable sort of the rows by slot -> row permutation (sets L.off_perm).
constexpr int kGroupsKeep = 0, kGroupsByKey = 1, kGroupsTableOrder = 2;   // group_order: leave as is / sort by key / table order
int launch_agg_segments(const int64_t* keys, int64_t n_rows, int64_t n_groups, char* scratch, AggLayout& L,
                        int64_t* out_keys_dev, int group_order, bool sort_rows, const DeviceInfo& dev, cudaStream_t stream);
// pass 4: out[g, :] = reduce over the rows of group g of payload[row, :].
// Scalar payloads: the values ride through their own stable slot sort and are reduced sequentially (no random gather).
// Call after launch_agg_segments(..., sort_rows = false) or before a launch_agg_segments that sorts rows.
size_t agg_carried_bytes(int64_t n_rows, int row_bytes, const AggLayout& L);   // row_bytes: 4 or 8
int launch_agg_reduce_carried(int dtype, int red_op, const void* payload, int64_t cell, int64_t n_groups, int64_t n_rows, char* scratch,
                              const AggLayout& L, void* out, char* carried, size_t carried_bytes, void* split_scratch,
                              size_t split_bytes, const DeviceInfo& dev, cudaStream_t stream);
// split_scratch (agg_split_scratch_bytes, may be null): partial rows when few large groups are cut into pieces.
size_t agg_split_scratch_bytes(int64_t cell, const DeviceInfo& dev);
int launch_agg_reduce(int dtype, int red_op, const void* payload, int64_t cell, int64_t n_groups, int64_t n_rows, char* scratch,
                      const AggLayout& L, void* out, void* split_scratch, size_t split_bytes, const DeviceInfo& dev,
                      cudaStream_t stream);

// Exchange step of aggregate: owner[i] = (mix64(key[i]) >> 32) mod world and the number of groups per owner (counts[world],
// zeroed by the call); rows gathered through a permutation (row_bytes a multiple of 4).
int launch_agg_owner(const int64_t* keys, int64_t n, int world, unsigned int* owner, unsigned long long* counts, const DeviceInfo& dev,
                     cudaStream_t stream);
int launch_gather_rows(const void* src, const unsigned int* perm, int64_t n, int64_t row_bytes, void* dst, const DeviceInfo& dev,
                       cudaStream_t stream);

// K0 (host_staging.cpp): copies of at least 64 MiB between PAGEABLE host memory and the device go through a pinned
// ring filled / drained by a few host threads; `stream` is ordered after the copy.
bool staged_copy_pays(size_t bytes);
void staged_copy(int device, void* dst, const void* src, size_t bytes, bool to_device, cudaStream_t stream);

}  // namespace tfs
