// Options build of the chain kernel (OPT = 1): see spa_chain.cuh.  Compiled next to spa_kernels.cu into
// libspa_b200.so; holds device code only.
#include "spa_chain.cuh"

namespace spa {
ChainKernel opt_chain_kernel_for(int bf16, int kvk, int nt) { return chain_kernel_table<1>(bf16, kvk, nt); }
}  // namespace spa
