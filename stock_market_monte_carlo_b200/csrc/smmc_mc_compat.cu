// BOOTSTRAP_COMPAT instantiations of the fused kernel (the reference's RNG stream, bit for bit).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_compat(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  switch (variant) {
    case 1: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 1>(p, li, s);
    case 2: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 2>(p, li, s);
    case 8: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 8>(p, li, s);
    default: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 4>(p, li, s);
  }
}
}  // namespace smmc
