// BOOTSTRAP instantiations of the fused kernel (Philox4x32-10, six 21/22-bit fields per call).
// The common case (32 bank-private copies per row: 128-byte rows) gets the row shift as an immediate.
#include "smmc_mc_kernel.cuh"

namespace smmc {
namespace {
template <int SHIFT>
cudaError_t launch_shift(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  switch (variant) {
    case 1: return launch_one<SMMC_MODE_BOOTSTRAP, 1, SHIFT>(p, li, s);
    case 2: return launch_one<SMMC_MODE_BOOTSTRAP, 2, SHIFT>(p, li, s);
    case 8: return launch_one<SMMC_MODE_BOOTSTRAP, 8, SHIFT>(p, li, s);
    default: return launch_one<SMMC_MODE_BOOTSTRAP, 4, SHIFT>(p, li, s);
  }
}
}  // namespace

cudaError_t launch_mc_bootstrap(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  return li.rep == 32 ? launch_shift<7>(p, li, variant, s) : launch_shift<-1>(p, li, variant, s);
}
}  // namespace smmc
