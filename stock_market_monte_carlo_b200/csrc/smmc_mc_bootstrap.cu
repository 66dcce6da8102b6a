// BOOTSTRAP instantiations of the fused kernel (Philox4x32-10 by default, Philox4x32-7 on request; six 21/22-bit fields
// per call).  The common case (32 bank-private copies per row: 128-byte rows) gets the row shift as an immediate.
#include "smmc_mc_kernel.cuh"

namespace smmc {
namespace {
template <int SHIFT, int ROUNDS>
cudaError_t launch_shift(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  switch (v.ilp) {  // test hook (smmc_debug_set_variant): every ILP gives identical bits
    case 2: return launch_one<SMMC_MODE_BOOTSTRAP, 2, SHIFT, ROUNDS>(p, li, s);
    case 4: return launch_one<SMMC_MODE_BOOTSTRAP, 4, SHIFT, ROUNDS>(p, li, s);
    default: return launch_one<SMMC_MODE_BOOTSTRAP, 8, SHIFT, ROUNDS>(p, li, s);  // +1.8 % (10 rounds) / +4.5 % (7) over 4, measured
  }
}
}  // namespace

cudaError_t launch_mc_bootstrap(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  if (p.philox_rounds == 7) return li.rep == 32 ? launch_shift<7, 7>(p, li, v, s) : launch_shift<-1, 7>(p, li, v, s);
  return li.rep == 32 ? launch_shift<7, 10>(p, li, v, s) : launch_shift<-1, 10>(p, li, v, s);
}
}  // namespace smmc
