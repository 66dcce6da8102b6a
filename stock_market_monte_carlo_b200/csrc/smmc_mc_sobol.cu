// SOBOL_BOOTSTRAP / SOBOL_GAUSSIAN instantiations of the fused kernel (all eight paths of a lane at once).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_sobol(int mode, const KParams &p, const LaunchInfo &li, cudaStream_t s) {
  if (mode == SMMC_MODE_SOBOL_BOOTSTRAP) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8>(p, li, s);
  return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8>(p, li, s);
}
}  // namespace smmc
