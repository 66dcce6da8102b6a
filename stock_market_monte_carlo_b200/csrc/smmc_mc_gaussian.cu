// GAUSSIAN instantiations of the fused kernel (Philox4x32-10 + Box-Muller, eight months per call).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_gaussian(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  switch (variant) {
    case 1: return launch_one<SMMC_MODE_GAUSSIAN, 1>(p, li, s);
    case 2: return launch_one<SMMC_MODE_GAUSSIAN, 2>(p, li, s);
    case 4: return launch_one<SMMC_MODE_GAUSSIAN, 4>(p, li, s);
    default: return launch_one<SMMC_MODE_GAUSSIAN, 8>(p, li, s);  // all eight paths of a lane at once: +1.1 % over 4 (measured)
  }
}
}  // namespace smmc
