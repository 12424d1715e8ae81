// K1 fast path -- dtype dispatch of the straight-line kernel (bodies: kernels_chain.inc, one translation unit per dtype).
#include "kernels.h"

namespace tfs {

int launch_chain_f32(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream);
int launch_chain_f64(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream);
int launch_chain_i32(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream);
int launch_chain_i64(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream);

int launch_chain(const ChainProgram& prog, int64_t n, const DeviceInfo& dev, cudaStream_t stream) {
  if (n <= 0) return 0;
  switch (prog.dt) {
    case DI_F32: return launch_chain_f32(prog, n, dev, stream);
    case DI_F64: return launch_chain_f64(prog, n, dev, stream);
    case DI_I32: return launch_chain_i32(prog, n, dev, stream);
    default: return launch_chain_i64(prog, n, dev, stream);
  }
}

}  // namespace tfs
