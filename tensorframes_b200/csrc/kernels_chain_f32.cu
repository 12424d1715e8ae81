// K1 fast path, float chains: see kernels_chain.inc.
#define TFS_CHAIN_T float
#define TFS_CHAIN_ENTRY launch_chain_f32
#include "kernels_chain.inc"
