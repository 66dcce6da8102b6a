// smmc_mc_kernel.cuh -- the fused persistent Monte Carlo kernel (template) shared by the per-mode translation units
// smmc_mc_{compat,bootstrap,gaussian,sobol}.cu (split so that the modes compile in parallel).
//
// Replaces the reference kernels mc_simulations_gpu_kernel / mc_simulations_gpu_kernel_reduceBlock
// (reference src/simulations.cu:112-247) with ONE persistent fused kernel per RNG mode:
//   * one CTA per SM, the return table staged once per CTA into shared memory by a TMA bulk copy (UBLKCP),
//     replicated 32x so that lane l only ever reads bank l (no bank conflicts on the random gather);
//   * one warp owns one logical 256-path block (lane l holds paths 32*j + l, j < 8): the compounding runs in
//     registers, the per-block mean / variance of the reference's reduceBlock layout is a pure warp-shuffle
//     reduction that reproduces the reference's reduction tree bit for bit;
//   * (count, mean, M2), the below-target count and a shared-memory privatised histogram are merged
//     warp -> CTA -> grid inside the same kernel (last-CTA-done ticket), so per-path data only touches HBM
//     when the caller asks for `finals`.
#pragma once
#include "smmc_kernels.cuh"

#include "../../include/smmc.h"

namespace smmc {

namespace {

// ---- (n, mean, M2) merge, Chan/Golub/LeVeque ------------------------------------------------------------------
struct Moments {
  double n, mean, m2;
};
__device__ __forceinline__ void merge(Moments &a, double nb, double meanb, double m2b) {
  if (nb == 0.0) return;
  const double tot = a.n + nb;
  const double delta = meanb - a.mean;
  a.mean += delta * (nb / tot);
  a.m2 += m2b + delta * delta * (a.n * nb / tot);
  a.n = tot;
}

// all 32 lanes' moments -> every lane, by a butterfly that merges in a lane-symmetric order (the lower lane is always
// "a"), so every lane ends with identical bits
__device__ __forceinline__ void butterfly_merge(Moments &c, uint32_t lane) {
#pragma unroll
  for (int st = 1; st <= 16; st <<= 1) {
    const double on = __shfl_xor_sync(0xffffffffu, c.n, st);
    const double om = __shfl_xor_sync(0xffffffffu, c.mean, st);
    const double o2 = __shfl_xor_sync(0xffffffffu, c.m2, st);
    Moments a, bb;
    if ((lane & st) == 0) {
      a = c;
      bb = {on, om, o2};
    } else {
      a = {on, om, o2};
      bb = c;
    }
    merge(a, bb.n, bb.mean, bb.m2);
    c = a;
  }
}

// the reference's 256-slot reduction tree (src/simulations.cu:157-183) on values held 8 per lane (slot 32*j+lane):
// strides 128, 64, 32 are lane-local adds, strides 16..1 are shuffles; lane 0's association order is the reference's.
__device__ __forceinline__ float tree_sum_256(const float (&s)[8]) {
  const float y0 = __fadd_rn(s[0], s[4]), y1 = __fadd_rn(s[1], s[5]);
  const float y2 = __fadd_rn(s[2], s[6]), y3 = __fadd_rn(s[3], s[7]);
  const float z0 = __fadd_rn(y0, y2), z1 = __fadd_rn(y1, y3);
  float w = __fadd_rn(z0, z1);
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1) w = __fadd_rn(w, __shfl_down_sync(0xffffffffu, w, st));
  return __shfl_sync(0xffffffffu, w, 0);
}

// ---- one sub-batch of ILP paths through all P months ------------------------------------------------------------

// ---- per-period withdrawals (the reference README's open "withdrawal strategies" item, README.md:130-133) -------------
// After the month's return: value <- max(value * q_m - w, 0), q_m = 1 - (percent taken out in month m)/100 read from a
// shared-memory schedule, w = fixed amount; 0 is absorbing (a ruined fund stays ruined).  FLOW = false compiles to
// nothing: the default kernels are unchanged.
template <bool FLOW>
struct Flow {
  const float *q;
  float w;
  __device__ __forceinline__ float operator()(float v, int m) const {
    if (!FLOW) return v;
    return fmaxf(fmaf(v, q[m], -w), 0.0f);
  }
};

// `base` = shared-window address of this lane's column of the table (smem_u32(tbl) + lane_off): the row address is
// then ONE IMAD (idx * row_bytes + base) feeding a plain LDS
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(addr));
  return r;
}

template <int ILP, bool FLOW>
__device__ __forceinline__ void run_compat(const KParams &p, const Flow<FLOW> &fl, uint32_t base, const uint32_t (&plo)[ILP],
                                           float (&tot)[ILP]) {
  uint32_t st[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) st[i] = rand_pcg(plo[i] + 1u + p.compat_seed);
  const float scale = p.idx_scale;
  const uint32_t rep4 = p.rep4;
#pragma unroll 4
  for (int m = 0; m < p.n_periods; m++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      st[i] = xorshift(st[i]);
      // reference: idx = n * (state * 2^-32f), truncated (src/simulations.cu:148); one rounding either way
      const int idx = (int)(__uint2float_rn(st[i]) * scale);
      const float r = lds_f32((uint32_t)idx * rep4 + base);
      tot[i] = fl(fmaf(tot[i], r, tot[i]), m);  // total += total * r (one FFMA, as the reference compiles)
    }
  }
}

// SOBOL_BOOTSTRAP draw: idx = mulhi(x, n) into the sorted table; row address by one IMAD.  (Tried in round 2: folding the
// row stride into the multiplier -- mulhi(x, n rep4) + base as the IMAD.HI addend, then one LOP3 to mask in the lane's
// column -- executes 7.3 instead of 8.9 instructions per path-month but runs 8 % SLOWER (1.880 vs 1.737 ms at C4): the
// LOP3 sits on the ALU pipe between the multiply and the load of every draw, the IMAD it replaces does not.)
__device__ __forceinline__ float boot_step(float tot, uint32_t w, uint32_t n, uint32_t rep4, uint32_t base) {
  const float r = lds_f32(__umulhi(w, n) * rep4 + base);  // idx = mulhi(w, n); row address by one IMAD
  return fmaf(tot, r, tot);
}

// months 6q..6q+5 of ILP paths: one Philox call each, six table gathers, six FFMAs
template <int ILP, int SHIFT, bool FLOW, int ROUNDS>
__device__ __forceinline__ void run_bootstrap(const KParams &p, const Flow<FLOW> &fl, uint32_t base,
                                              const uint32_t (&plo)[ILP], uint32_t phi, float (&tot)[ILP]) {
  const BootConsts bc = p.bc;
  const uint32_t one = p.one;
  const int nq = p.n_periods / kFieldsPerCall;
#pragma unroll 1
  for (int q = 0; q < nq; q++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4], m[6];
      stream_words<ROUNDS>(plo[i], phi, (uint32_t)q, kTagBootstrap, p.keys, w);
      field_mantissas(w, one, m);
#pragma unroll
      for (int j = 0; j < 6; j++) {
        const float r = boot_load<SHIFT>(m[j], j < 4 ? bc.c21 : bc.c22, j < 4 ? bc.k21 : bc.k22, base, bc.row_shift);
        tot[i] = fl(fmaf(tot[i], r, tot[i]), q * kFieldsPerCall + j);
      }
    }
  }
  const int rem = p.n_periods - nq * kFieldsPerCall;
  if (rem) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4], m[6];
      stream_words<ROUNDS>(plo[i], phi, (uint32_t)nq, kTagBootstrap, p.keys, w);
      field_mantissas(w, one, m);
#pragma unroll
      for (int j = 0; j < 5; j++)
        if (j < rem) {
          const float r = boot_load<SHIFT>(m[j], j < 4 ? bc.c21 : bc.c22, j < 4 ? bc.k21 : bc.k22, base, bc.row_shift);
          tot[i] = fl(fmaf(tot[i], r, tot[i]), nq * kFieldsPerCall + j);
        }
    }
  }
}

// months 8q..8q+7 of ILP paths: one Philox call each, four Box-Muller pairs.  TBL is a 4-bit mask: pair j of a call takes
// sin/cos from the shared-memory table (sincos_tbl, no MUFU) when bit j is set, from MUFU.SIN/COS otherwise -- the two
// evaluate the same angle of the same contract (smmc_device.cuh); the mix balances the XU pipe against the FMA/ALU pipes.
template <int TBL>
__device__ __forceinline__ void gauss_pair_sel(int j, uint32_t w, uint32_t one, uint32_t cell_mask, const GaussConsts gc,
                                               const char *sc_tbl, float &g0, float &g1) {
  if ((TBL >> j) & 1) gauss_pair_tbl(w, one, cell_mask, gc, sc_tbl, g0, g1);
  else gauss_pair(w, one, gc, g0, g1);
}

template <int ILP, bool FLOW, int ROUNDS, int TBL>
__device__ __forceinline__ void run_gaussian(const KParams &p, const Flow<FLOW> &fl, const char *sc_tbl,
                                             const uint32_t (&plo)[ILP], uint32_t phi, float (&tot)[ILP]) {
  const GaussConsts gc = p.gc;
  const uint32_t one = p.one, cell_mask = p.cell_mask;
  const int nq = p.n_periods / kGaussMonthsPerCall;
#pragma unroll 1
  for (int q = 0; q < nq; q++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4];
      stream_words<ROUNDS>(plo[i], phi, (uint32_t)q, kTagGaussian, p.keys, w);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float g0, g1;
        gauss_pair_sel<TBL>(j, w[j], one, cell_mask, gc, sc_tbl, g0, g1);
        const int m = q * kGaussMonthsPerCall + 2 * j;
        tot[i] = fl(__fmul_rn(fl(__fmul_rn(tot[i], g0), m), g1), m + 1);
      }
    }
  }
  const int rem = p.n_periods - nq * kGaussMonthsPerCall;
  if (rem) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4];
      stream_words<ROUNDS>(plo[i], phi, (uint32_t)nq, kTagGaussian, p.keys, w);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float g0, g1;
        gauss_pair_sel<TBL>(j, w[j], one, cell_mask, gc, sc_tbl, g0, g1);
        const int m = nq * kGaussMonthsPerCall + 2 * j;
        if (2 * j < rem) tot[i] = fl(__fmul_rn(tot[i], g0), m);
        if (2 * j + 1 < rem) tot[i] = fl(__fmul_rn(tot[i], g1), m + 1);
      }
    }
  }
}

// ---- Sobol modes: all 8 paths of the lane at once -------------------------------------------------------------------
// Point i = path + 256 = 256 h + c with c = 32 j + lane (the block base is 256-aligned), x_m(i) = XOR over the set
// bits k of gray(i) of V[m][k].  gray bits >= 8 depend on h only: that part (U_m, with the digital shift and the
// h_0 contribution to gray bit 7 folded in) is warp-uniform, computed 32 months at a time with lane = month, and
// broadcast by one SHFL per month.  gray bits 0..3 depend on the lane only (masks), bits 4..7 on (lane_4, j): the
// eight per-j variants are one or two 3-input XORs away from a common base.  ~2 instructions per path-month.
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
// base ^= v when the (loop-invariant) lane flag is set: one predicated LOP3
__device__ __forceinline__ void xor_if(uint32_t &base, uint32_t v, uint32_t flag) {
  asm("{\n .reg .pred p;\n setp.ne.u32 p, %2, 0;\n @p xor.b32 %0, %0, %1;\n}" : "+r"(base) : "r"(v), "r"(flag));
}

template <int MODE, bool FLOW, int TBL, int UNROLL>
__device__ __forceinline__ void run_sobol(const KParams &p, const Flow<FLOW> &fl, const char *tbl, uint32_t lane_off,
                                          const uint4 *vlo, uint32_t h, uint32_t lane, float (&tot)[8]) {
  const uint32_t mk0 = ((lane ^ (lane >> 1)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk1 = (((lane >> 1) ^ (lane >> 2)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk2 = (((lane >> 2) ^ (lane >> 3)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk3 = (((lane >> 3) ^ (lane >> 4)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk4 = ((lane >> 4) & 1u) ? 0xffffffffu : 0u;
  const uint32_t gh = h ^ (h >> 1);
  const int ppad = p.sobol_ppad;
  const uint32_t n = (uint32_t)p.n_returns, rep4 = p.rep4;
  const uint32_t lane_base = smem_u32(tbl) + lane_off;  // shared-window address of this lane's table column
  const uint32_t one = p.one;
  const GaussConsts gc = p.gc;
  const uint32_t sc_base = smem_u32(tbl), cell_mask = p.cell_mask;
  // warp-uniform part of month m0 + lane.  The host tabulates, per month, the XOR of the direction numbers of every 4-bit
  // group of gray(h) (six groups x 16 combinations) and the two variants of (digital shift ^ [h odd] V[7]): seven
  // independent, unconditional loads (L2-resident table) instead of 26 predicated ones behind 26 branches -- that was
  // 0.6 of the 8.9 executed instructions per path-month in round 1.  The next 32 months are fetched while the current
  // 32 are computed.
  auto block_part = [&](int m0) {
    const uint32_t *col = p.sobol_hi + m0 + lane;
    uint32_t U = __ldg(col + (h & 1u) * ppad);
#pragma unroll
    for (int g = 0; g < 6; g++) U ^= __ldg(col + (2u + 16u * g + ((gh >> (4 * g)) & 15u)) * ppad);
    return U;
  };
  // the eight words of month m (one per path of the lane) from the broadcast block part Um
  auto month_words = [&](int m, uint32_t Um, uint32_t (&x)[8]) {
    const uint4 va = vlo[m * 2], vb = vlo[m * 2 + 1];
    uint32_t base = Um;
    if (MODE == SMMC_MODE_SOBOL_GAUSSIAN) {
      // issue-bound kernel: five predicated XORs (the lane's gray bits are loop-invariant predicates) and seven explicit
      // 3-input XORs = 12 instructions; the mask form below compiles to 4 SEL + 3 LOP3 + 8 LOP3 = 15 (C4: 2.016 -> 1.936 ms
      // together with the LEA.HI table address)
      xor_if(base, va.x, mk0);
      xor_if(base, va.y, mk1);
      xor_if(base, va.z, mk2);
      xor_if(base, va.w, mk3);
      xor_if(base, vb.x, mk4);
      x[0] = base;
      x[1] = xor3(base, vb.x, vb.y);
      x[2] = xor3(base, vb.y, vb.z);
      x[3] = xor3(base, vb.x, vb.z);
      x[4] = xor3(base, vb.z, vb.w);
      x[5] = xor3(x[1], vb.z, vb.w);
      x[6] = xor3(base, vb.y, vb.w);
      x[7] = xor3(base, vb.x, vb.w);
      return;
    }
    // SOBOL_BOOTSTRAP is bound by the multiplier pipe, not by issue: the form the compiler schedules best (1.388 vs 1.395 ms)
    base ^= (va.x & mk0);
    base ^= (va.y & mk1);
    base ^= (va.z & mk2);
    base ^= (va.w & mk3);
    base ^= (vb.x & mk4);
    x[0] = base;
    x[1] = base ^ vb.x ^ vb.y;
    x[2] = base ^ vb.y ^ vb.z;
    x[3] = base ^ vb.x ^ vb.z;
    x[4] = base ^ vb.z ^ vb.w;
    x[5] = x[1] ^ vb.z ^ vb.w;
    x[6] = base ^ vb.y ^ vb.w;
    x[7] = base ^ vb.x ^ vb.w;
  };
  uint32_t U_next = p.n_periods > 0 ? block_part(0) : 0u;
#pragma unroll 1
  for (int m0 = 0; m0 < p.n_periods; m0 += 32) {
    const uint32_t U = U_next;
    if (m0 + 32 < p.n_periods) U_next = block_part(m0 + 32);
    const int mend = (p.n_periods - m0) < 32 ? (p.n_periods - m0) : 32;
    if (MODE == SMMC_MODE_SOBOL_BOOTSTRAP) {
#pragma unroll UNROLL
      for (int mi = 0; mi < mend; mi++) {
        uint32_t x[8];
        month_words(m0 + mi, __shfl_sync(0xffffffffu, U, mi), x);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          // TBL (SOBOL_BOOTSTRAP): path slots whose row stride is the constant 128 (32 table copies): IMAD with an immediate
          // multiplier instead of a register one (smmc_mc_sobol.cu)
          if ((TBL >> j) & 1) tot[j] = fl(fmaf(tot[j], lds_f32((__umulhi(x[j], n) << 7) + lane_base), tot[j]), m0 + mi);
          else tot[j] = fl(boot_step(tot[j], x[j], n, rep4, lane_base), m0 + mi);
        }
      }
    } else {
      // Box-Muller on dimension pairs (smmc_device.cuh): even month = radius word, odd month = angle word
#pragma unroll 1
      for (int mi = 0; mi + 1 < mend; mi += 2) {
        uint32_t xr[8], xa[8];
        month_words(m0 + mi, __shfl_sync(0xffffffffu, U, mi), xr);
        month_words(m0 + mi + 1, __shfl_sync(0xffffffffu, U, mi + 1), xa);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          float g0, g1;
          if ((TBL >> j) & 1) sobol_gauss_pair_tbl(xr[j], xa[j], one, gc, cell_mask, sc_base, g0, g1);  // path slot j: table
          else sobol_gauss_pair(xr[j], xa[j], gc, g0, g1);
          tot[j] = fl(__fmul_rn(fl(__fmul_rn(tot[j], g0), m0 + mi), g1), m0 + mi + 1);
        }
      }
      if (mend & 1) {  // odd horizon: the last month has no partner dimension -> inverse normal CDF of its word
        uint32_t x[8];
        month_words(m0 + mend - 1, __shfl_sync(0xffffffffu, U, mend - 1), x);
#pragma unroll
        for (int j = 0; j < 8; j++)
          tot[j] = fl(__fmul_rn(tot[j], fmaf(erfinv_py(x[j]), p.sigma_sqrt2, gc.a)), m0 + mend - 1);
      }
    }
  }
}

// ---- the fused persistent kernel ---------------------------------------------------------------------------------

// MODE: smmc_mode.  ILP: paths of a lane in flight at once.  SHIFT: log2(row bytes) as an immediate (BOOTSTRAP), -1 = from
// KParams.  FLOW: per-period withdrawals.  ROUNDS: Philox rounds (10 | 7).  TBL: Gaussian modes -- which Box-Muller pairs
// take sin/cos from the shared-memory table (GAUSSIAN: 4-bit mask over the pairs of a call; SOBOL_GAUSSIAN: 0 | 1).
// UNROLL: month-loop unroll of SOBOL_BOOTSTRAP.
template <int MODE, int ILP, int SHIFT, bool FLOW, int ROUNDS, int TBL, int UNROLL>
__global__ void __launch_bounds__(kThreads, 1) mc_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr bool kGauss = (MODE == SMMC_MODE_GAUSSIAN || MODE == SMMC_MODE_SOBOL_GAUSSIAN);
  // "table" = the bank-private return table (bootstrap modes) or the sin/cos table (Gaussian modes with TBL != 0)
  constexpr bool kHasTable = (MODE == SMMC_MODE_BOOTSTRAP_COMPAT || MODE == SMMC_MODE_BOOTSTRAP ||
                              MODE == SMMC_MODE_SOBOL_BOOTSTRAP || (kGauss && TBL != 0));
  constexpr bool kSobol = (MODE == SMMC_MODE_SOBOL_BOOTSTRAP || MODE == SMMC_MODE_SOBOL_GAUSSIAN);
  const SmemLayout L = smem_layout(p.hist_bins, (uint32_t)p.sel_targets * kSelectBins * 4u, FLOW ? (uint32_t)p.n_periods * 4u : 0u,
                                   kSobol ? p.sobol_lo_bytes : 0u, kHasTable ? p.tbl_bytes : 0u);
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
  unsigned int *hist_s = reinterpret_cast<unsigned int *>(smem + L.hist_off);  // u32 per CTA (native smem atomics), u64 globally
  double *wp = reinterpret_cast<double *>(smem + L.wp_off);
  unsigned int *sel_s = reinterpret_cast<unsigned int *>(smem + L.sel_off);
  float *flow_s = reinterpret_cast<float *>(smem + L.flow_off);
  const char *tbl = reinterpret_cast<const char *>(smem + L.tbl_off);
  const uint4 *vlo = reinterpret_cast<const uint4 *>(smem + L.sob_off);

  const int tid = threadIdx.x;
  const int n_warps = (int)(blockDim.x >> 5);  // chosen per launch (plan_launch): evens out the last wave
  const uint32_t lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler

  // ---- stage the table: one thread arms the mbarrier and issues the bulk copies, everyone waits on the phase
  if (kHasTable || kSobol) {
    if (tid == 0) {
      mbar_init(bar, 1);
      fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
      const uint32_t tb = kHasTable ? p.tbl_bytes : 0u, sb = kSobol ? p.sobol_lo_bytes : 0u;
      mbar_arrive_expect_tx(bar, tb + sb);
      constexpr uint32_t kChunk = 32768;
      for (uint32_t o = 0; o < tb; o += kChunk)
        tma_bulk_g2s(smem + L.tbl_off + o, reinterpret_cast<const char *>(p.tbl_rep) + o,
                     (tb - o) < kChunk ? (tb - o) : kChunk, bar);
      for (uint32_t o = 0; o < sb; o += kChunk)
        tma_bulk_g2s(smem + L.sob_off + o, reinterpret_cast<const char *>(p.sobol_lo) + o,
                     (sb - o) < kChunk ? (sb - o) : kChunk, bar);
    }
  }
  for (int i = tid; i < p.hist_bins + 2; i += (int)blockDim.x) hist_s[i] = 0u;
  for (int i = tid; i < p.sel_targets * kSelectBins; i += (int)blockDim.x) sel_s[i] = 0u;
  if (FLOW)
    for (int i = tid; i < p.n_periods; i += (int)blockDim.x) flow_s[i] = __ldg(p.flow_q + i);
  const Flow<FLOW> fl = {flow_s, p.flow_w};
  __syncthreads();
  if (kHasTable || kSobol) mbar_wait(bar, 0);

  const int64_t N = p.n_paths;
  const int64_t nb = (N + kBlock - 1) / kBlock;
  const int64_t total_warps = (int64_t)gridDim.x * n_warps;
  const uint32_t lane_off = (lane * 4u) & (p.rep4 - 1u);
  constexpr bool kCompat = (MODE == SMMC_MODE_BOOTSTRAP_COMPAT);
  // BOOTSTRAP: shared-window address of this lane's column, minus the float exponent bits that boot_load shifts out
  const uint32_t lane_base = smem_u32(tbl) + lane_off;  // shared-window address of this lane's table column
  const uint32_t boot_base = lane_base - (0x4B000000u << (SHIFT >= 0 ? (uint32_t)SHIFT : p.bc.row_shift));

  Moments acc = {0.0, 0.0, 0.0};
  unsigned int below_lane = 0;
  unsigned long long below_acc = 0ull;

  const bool dyn = p.block_ticket != nullptr;
  for (int64_t b = (int64_t)blockIdx.x * n_warps + warp; b < nb;) {
    const uint64_t base = p.path0 + (uint64_t)b * kBlock;  // 256-aligned: low 8 bits free for the in-block offset
    const uint32_t base_lo = (uint32_t)base, phi = (uint32_t)(base >> 32);
    const int64_t rem_paths = N - b * kBlock;
    const int live = rem_paths < kBlock ? (int)rem_paths : kBlock;

    float x[kPathsPerLane];
    if (kSobol) {
#pragma unroll
      for (int j = 0; j < kPathsPerLane; j++) x[j] = p.v0;
      run_sobol<MODE, FLOW, TBL, UNROLL>(p, fl, tbl, lane_off, vlo, (uint32_t)((base + kSobolSkip) >> 8), lane, x);
    }
#pragma unroll
    for (int s = 0; s < (kSobol ? 0 : kPathsPerLane); s += ILP) {
      uint32_t plo[ILP];
      float tot[ILP];
#pragma unroll
      for (int i = 0; i < ILP; i++) {
        plo[i] = base_lo + (uint32_t)((s + i) * 32) + lane;
        tot[i] = p.v0;
      }
      if (MODE == SMMC_MODE_BOOTSTRAP_COMPAT) run_compat<ILP, FLOW>(p, fl, lane_base, plo, tot);
      if (MODE == SMMC_MODE_BOOTSTRAP) run_bootstrap<ILP, SHIFT, FLOW, ROUNDS>(p, fl, boot_base, plo, phi, tot);
      if (MODE == SMMC_MODE_GAUSSIAN) run_gaussian<ILP, FLOW, ROUNDS, TBL>(p, fl, tbl, plo, phi, tot);
#pragma unroll
      for (int i = 0; i < ILP; i++) x[s + i] = tot[i];
    }

    // ---- per-path outputs: finals (optional), below-target count, histogram
    bool valid[kPathsPerLane];
#pragma unroll
    for (int j = 0; j < kPathsPerLane; j++) {
      valid[j] = (j * 32 + (int)lane) < live;
      if (valid[j]) {
        if (p.finals) p.finals[b * kBlock + j * 32 + lane] = x[j];
        below_lane += (x[j] < p.target) ? 1u : 0u;
        if (p.hist_bins > 0) atomicAdd(&hist_s[hist_bin(x[j], p.hist_lo, p.hist_inv_w, p.hist_bins)], 1u);
        if (p.sel_targets > 0) {  // radix-select pass: digit histogram of the keys under each target's prefix
          const uint32_t key = ordered_key(x[j]);
          const uint32_t digit = (key >> p.sel_digit_shift) & p.sel_digit_mask;
          const uint32_t pre = p.sel_prefix_shift >= 32 ? 0u : key >> p.sel_prefix_shift;
          for (int t = 0; t < p.sel_targets; t++)
            if (pre == p.sel_prefix[t]) atomicAdd(&sel_s[t * kSelectBins + digit], 1u);
        }
      }
    }
    if (p.cand_groups > 0) {  // second select pass: keep the keys under the targets' 11-bit prefixes (warp-uniform branch)
      uint32_t match = 0;
#pragma unroll
      for (int j = 0; j < kPathsPerLane; j++) {
        const uint32_t pre11 = ordered_key(x[j]) >> 21;
        bool hit = false;
        for (int g = 0; g < p.cand_groups; g++) hit = hit || (pre11 == p.cand_prefix[g]);
        if (hit && valid[j]) match |= 1u << j;
      }
      const uint32_t cnt = (uint32_t)__popc(match);
      uint32_t incl = cnt;  // inclusive warp scan of the per-lane counts: ONE global atomic per 256-path block
#pragma unroll
      for (int st = 1; st < 32; st <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, st);
        if ((int)lane >= st) incl += o;
      }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      unsigned long long base_c = 0ull;
      if (lane == 0 && total) base_c = atomicAdd(p.cand_cursor, (unsigned long long)total);
      base_c = __shfl_sync(0xffffffffu, base_c, 0) + (incl - cnt);
#pragma unroll
      for (int j = 0; j < kPathsPerLane; j++)
        if ((match >> j) & 1u) {
          if (base_c < p.cand_capacity) p.cand_keys[base_c] = ordered_key(x[j]);
          base_c++;
        }
    }

    // ---- block mean / population variance in the reference's reduceBlock layout and reduction order
    // (src/simulations.cu:224-246): dead slots contribute 0; COMPAT keeps the reference's divisor of 256 for a
    // partial last block, the native modes divide by the number of live paths.
    float s[kPathsPerLane];
#pragma unroll
    for (int j = 0; j < kPathsPerLane; j++) s[j] = valid[j] ? x[j] : 0.0f;
    const float sum = tree_sum_256(s);
    const bool full = (live == kBlock);
    const float div = (kCompat || full) ? 256.0f : (float)live;
    const float mean = __fdiv_rn(sum, div);
#pragma unroll
    for (int j = 0; j < kPathsPerLane; j++) {
      const float d = __fsub_rn(x[j], mean);
      s[j] = valid[j] ? __fmul_rn(d, d) : 0.0f;
    }
    const float sq = tree_sum_256(s);
    if (lane == 0) {
      if (p.bmeans) p.bmeans[b] = mean;
      if (p.bvars) p.bvars[b] = __fdiv_rn(sq, div);
    }
    double mean_live = (double)mean, m2_live = (double)sq;
    if (kCompat && !full) {  // global moments always describe the live paths only
      const float tmean = __fdiv_rn(sum, (float)live);
#pragma unroll
      for (int j = 0; j < kPathsPerLane; j++) {
        const float d = __fsub_rn(x[j], tmean);
        s[j] = valid[j] ? __fmul_rn(d, d) : 0.0f;
      }
      mean_live = (double)tmean;
      m2_live = (double)tree_sum_256(s);
    }
    if (below_lane > 0x7fffff00u) {  // keep the 32-bit per-lane counter far from wrapping
      below_acc += below_lane;
      below_lane = 0;
    }
    if (dyn) {  // both values are fp32 results widened above: the float2 loses nothing
      unsigned int t = 0;
      if (lane == 0) {
        p.block_moments[b] = make_float2((float)mean_live, (float)m2_live);
        t = atomicAdd(p.block_ticket, 1u);
      }
      b = total_warps + (int64_t)__shfl_sync(0xffffffffu, t, 0);
    } else {
      merge(acc, (double)live, mean_live, m2_live);
      b += total_warps;
    }
  }

  // ---- warp -> CTA -> grid
  below_acc += below_lane;
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1) below_acc += __shfl_xor_sync(0xffffffffu, below_acc, st);
  if (lane == 0) {
    wp[warp * 4 + 0] = acc.n;
    wp[warp * 4 + 1] = acc.mean;
    wp[warp * 4 + 2] = acc.m2;
    wp[warp * 4 + 3] = (double)below_acc;  // < 2^53, exact
  }
  __syncthreads();
  __shared__ bool is_last;
  if (tid == 0) {
    Moments c = {0.0, 0.0, 0.0};
    double below = 0.0;
    for (int w = 0; w < n_warps; w++) {
      merge(c, wp[w * 4 + 0], wp[w * 4 + 1], wp[w * 4 + 2]);
      below += wp[w * 4 + 3];
    }
    double *cp = p.cta_partials + (size_t)blockIdx.x * 4;
    cp[0] = c.n;
    cp[1] = c.mean;
    cp[2] = c.m2;
    cp[3] = below;
  }
  for (int i = tid; i < p.hist_bins + 2; i += (int)blockDim.x)
    if (hist_s[i]) atomicAdd(&p.hist_g[i], (unsigned long long)hist_s[i]);
  for (int i = tid; i < p.sel_targets * kSelectBins; i += (int)blockDim.x)
    if (sel_s[i]) atomicAdd(&p.sel_hist_g[i], (unsigned long long)sel_s[i]);
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = (atomicAdd(p.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // last CTA.  Ticket launches: the blocks' moments are merged in block order -- thread t takes blocks t, t + T, ...
  // sequentially, lanes by the fixed butterfly, warps in warp order -- so the result does not depend on which warp ran
  // which block.
  if (dyn) {
    Moments c = {0.0, 0.0, 0.0};
    for (int64_t i = tid; i < nb; i += (int64_t)blockDim.x) {
      const float2 m = __ldcg(p.block_moments + i);
      const int64_t rem = N - i * kBlock;
      merge(c, (double)(rem < kBlock ? rem : (int64_t)kBlock), (double)m.x, (double)m.y);
    }
    butterfly_merge(c, lane);
    __syncthreads();  // wp[] of the CTA-level merge above is free again
    if (lane == 0) {
      wp[warp * 4 + 0] = c.n;
      wp[warp * 4 + 1] = c.mean;
      wp[warp * 4 + 2] = c.m2;
    }
    __syncthreads();
  }
  // Merge the per-CTA partials in CTA order (warp 0: lane-strided sequential merge, then a fixed butterfly), emit the
  // allreduce-ready statistics vector (include/smmc.h, SMMC_STATS_HEADER)
  if (warp == 0) {
    Moments c = {0.0, 0.0, 0.0};
    double below = 0.0;
    for (unsigned g = lane; g < gridDim.x; g += 32) {
      const double *cp = p.cta_partials + (size_t)g * 4;
      merge(c, __ldcg(cp + 0), __ldcg(cp + 1), __ldcg(cp + 2));
      below += __ldcg(cp + 3);
    }
    butterfly_merge(c, lane);
#pragma unroll
    for (int st = 1; st <= 16; st <<= 1) below += __shfl_xor_sync(0xffffffffu, below, st);  // integers < 2^53: exact
    if (dyn) {  // every lane computes the same warp-order merge of the block-order partials
      c = {0.0, 0.0, 0.0};
      for (int w = 0; w < n_warps; w++) merge(c, wp[w * 4 + 0], wp[w * 4 + 1], wp[w * 4 + 2]);
    }
    if (lane == 0) {
      const double dm = c.mean - (double)p.v0;
      p.stats[0] = c.n;
      p.stats[1] = below;
      p.stats[2] = c.n * dm;               // S1 = sum (x - V0)
      p.stats[3] = c.m2 + c.n * dm * dm;   // S2 = sum (x - V0)^2
      p.stats[4] = 0.0;
      p.stats[5] = 0.0;
    }
  }
  for (int i = tid; i < p.hist_bins + 2; i += (int)blockDim.x)
    p.stats[SMMC_STATS_HEADER + i] = (double)__ldcg(&p.hist_g[i]);
}

template <int MODE, int ILP, int SHIFT, bool FLOW, int ROUNDS, int TBL, int UNROLL>
cudaError_t launch_flow(const KParams &p, const LaunchInfo &li, cudaStream_t s) {
  auto kern = mc_kernel<MODE, ILP, SHIFT, FLOW, ROUNDS, TBL, UNROLL>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)li.smem_bytes);
  if (e != cudaSuccess) return e;
  kern<<<li.grid, li.threads, li.smem_bytes, s>>>(p);
  return cudaGetLastError();
}

// the withdrawal variant is a separate instantiation: the default kernels carry no trace of it
template <int MODE, int ILP, int SHIFT = -1, int ROUNDS = 10, int TBL = 0, int UNROLL = 2>
cudaError_t launch_one(const KParams &p, const LaunchInfo &li, cudaStream_t s) {
  return p.flow_q ? launch_flow<MODE, ILP, SHIFT, true, ROUNDS, TBL, UNROLL>(p, li, s)
                  : launch_flow<MODE, ILP, SHIFT, false, ROUNDS, TBL, UNROLL>(p, li, s);
}

}  // namespace

}  // namespace smmc
