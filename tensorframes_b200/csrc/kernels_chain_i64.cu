// K1 fast path, long long chains: see kernels_chain.inc.
#define TFS_CHAIN_T long long
#define TFS_CHAIN_ENTRY launch_chain_i64
#include "kernels_chain.inc"
