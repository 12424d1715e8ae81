// K1 fast path, double chains: see kernels_chain.inc.
#define TFS_CHAIN_T double
#define TFS_CHAIN_ENTRY launch_chain_f64
#include "kernels_chain.inc"
