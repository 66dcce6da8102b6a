// SOBOL_BOOTSTRAP / SOBOL_GAUSSIAN instantiations of the fused kernel (all eight paths of a lane at once).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_sobol(int mode, const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  if (mode == SMMC_MODE_SOBOL_BOOTSTRAP) {
#ifdef SMMC_EXPERIMENT
    if (v.sobol_unroll == 4) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, 4>(p, li, s);
    if (v.sobol_unroll == 8) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, 8>(p, li, s);
#endif
    return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, kSobolUnrollDefault>(p, li, s);
  }
  return p.tbl_bytes ? launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 1>(p, li, s)
                     : launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0>(p, li, s);
}
}  // namespace smmc
