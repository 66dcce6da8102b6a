// GAUSSIAN instantiations of the fused kernel (Philox4x32-10 by default, Philox4x32-7 on request; Box-Muller, eight months
// per call).  TBL (which pairs of a call take sin/cos from the shared-memory table) is a property of the build, not of
// the contract: every mask evaluates the same angles (smmc_device.cuh, sincos_tbl).
#include "smmc_mc_kernel.cuh"

namespace smmc {
namespace {
template <int ROUNDS>
cudaError_t launch_rounds(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  const int tbl = p.tbl_bytes ? p.gauss_tbl_mask : 0;  // chosen by the host (gauss_table_mask): only masks this build has
  switch (tbl) {
    case 0: return v.ilp == 4 ? launch_one<SMMC_MODE_GAUSSIAN, 4, -1, ROUNDS, 0>(p, li, s)
                              : launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 0>(p, li, s);
#ifdef SMMC_EXPERIMENT
    case 1: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 1>(p, li, s);
    case 3: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 3>(p, li, s);
    case 6: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 6>(p, li, s);
    case 10: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 10>(p, li, s);
    case 12: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 12>(p, li, s);
    case 15: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 15>(p, li, s);
    case (ROUNDS == 7 ? 5 : 7): return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, (ROUNDS == 7 ? 5 : 7)>(p, li, s);
    case 9: return launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, 9>(p, li, s);
#endif
    default: return v.ilp == 4 ? launch_one<SMMC_MODE_GAUSSIAN, 4, -1, ROUNDS, gauss_tbl_default(ROUNDS)>(p, li, s)
                               : launch_one<SMMC_MODE_GAUSSIAN, 8, -1, ROUNDS, gauss_tbl_default(ROUNDS)>(p, li, s);
  }
}
}  // namespace

cudaError_t launch_mc_gaussian(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  return p.philox_rounds == 7 ? launch_rounds<7>(p, li, v, s) : launch_rounds<10>(p, li, v, s);
}
}  // namespace smmc
