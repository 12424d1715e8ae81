// K1 fast path, int chains: see kernels_chain.inc.
#define TFS_CHAIN_T int
#define TFS_CHAIN_ENTRY launch_chain_i32
#include "kernels_chain.inc"
