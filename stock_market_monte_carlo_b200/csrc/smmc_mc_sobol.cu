// SOBOL_BOOTSTRAP / SOBOL_GAUSSIAN instantiations of the fused kernel (all eight paths of a lane at once).
#include "smmc_mc_kernel.cuh"

namespace smmc {
cudaError_t launch_mc_sobol(int mode, const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  if (mode == SMMC_MODE_SOBOL_BOOTSTRAP) {
#ifdef SMMC_EXPERIMENT
    if (v.sobol_unroll == 4) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, 4>(p, li, s);
    if (v.sobol_unroll == 8) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, 8>(p, li, s);
    if (p.rep4 == 128u) {
      if (v.sobol_unroll == 100) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0xFF, kSobolUnrollDefault>(p, li, s);
      if (v.sobol_unroll == 101) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, kSobolUnrollDefault>(p, li, s);
      if (v.sobol_unroll == 102) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0x11, kSobolUnrollDefault>(p, li, s);
      if (v.sobol_unroll == 103) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0x77, kSobolUnrollDefault>(p, li, s);
    }
#endif
    // 32 table copies (rows of 128 bytes, every table up to ~1800 entries): for half of a lane's path slots the row stride is a
    // compile-time constant -- ptxas emits IMAD R, R, 0x80, R (immediate multiplier, two register operands) where the generic
    // form is IMAD R, R, R, R.  Same pipe, same count, but 2.5 % faster (5 M x 1000, same box: generic 1.433 ms, slots 0x55
    // immediate 1.397, 0x77 1.402, all 1.406, 0x11 1.435): fewer register-file reads in front of the multiplier pipe.
    if (p.rep4 == 128u) return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, kSobolBootImmSlots, kSobolUnrollDefault>(p, li, s);
    return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8, -1, 10, 0, kSobolUnrollDefault>(p, li, s);
  }
  // TBL: which of a lane's eight path slots take sin/cos from the shared-memory table (the others use MUFU.SIN/COS)
  if (!p.tbl_bytes) return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0>(p, li, s);
  switch (p.gauss_tbl_mask) {
#ifdef SMMC_EXPERIMENT
    case 0x7F: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0x7F>(p, li, s);
    case 0x77: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0x77>(p, li, s);
    case 0x6D: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0x6D>(p, li, s);
    case 0x55: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0x55>(p, li, s);
    case 0xBB: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xBB>(p, li, s);
    case 0xEE: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xEE>(p, li, s);
    case 0xDD: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xDD>(p, li, s);
    case 0x3F: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0x3F>(p, li, s);
    case 0xCF: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xCF>(p, li, s);
    case 0xB7: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xB7>(p, li, s);
    case 0xFF: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xFF>(p, li, s);
#else
    case 0xFF: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, 0xFF>(p, li, s);
#endif
    default: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8, -1, 10, kSobolTblDefault>(p, li, s);
  }
}
}  // namespace smmc
