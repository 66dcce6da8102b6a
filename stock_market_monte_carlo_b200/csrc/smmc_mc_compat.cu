// BOOTSTRAP_COMPAT instantiations of the fused kernel (the reference's RNG stream, bit for bit).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_compat(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  switch (v.ilp) {  // test hook (smmc_debug_set_variant): every ILP gives identical bits
    case 2: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 2>(p, li, s);
    case 8: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 8>(p, li, s);
    default: return launch_one<SMMC_MODE_BOOTSTRAP_COMPAT, 4>(p, li, s);
  }
}
}  // namespace smmc
