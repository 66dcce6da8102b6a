// smmc_kernels.cu -- hand-written sm_100a kernels of the Monte Carlo portfolio hot path.
//
// Replaces the reference kernels mc_simulations_gpu_kernel / mc_simulations_gpu_kernel_reduceBlock /
// compute_sum_kernel / impl_vector_add_gpu (reference src/simulations.cu:112-267, src/gpu.cu:8-14) with ONE
// persistent fused kernel per RNG mode:
//   * one CTA per SM, the return table staged once per CTA into shared memory by a TMA bulk copy (UBLKCP),
//     replicated 32x so that lane l only ever reads bank l (no bank conflicts on the random gather);
//   * one warp owns one logical 256-path block (lane l holds paths 32*j + l, j < 8): the compounding runs in
//     registers, the per-block mean / variance of the reference's reduceBlock layout is a pure warp-shuffle
//     reduction that reproduces the reference's reduction tree bit for bit;
//   * (count, mean, M2), the below-target count and a shared-memory privatised histogram are merged
//     warp -> CTA -> grid inside the same kernel (last-CTA-done ticket), so per-path data only touches HBM
//     when the caller asks for `finals`.
#include "smmc_kernels.cuh"

#include "../../include/smmc.h"

namespace smmc {

namespace {

constexpr int kThreads = 1024;
constexpr int kWarps = kThreads / 32;

// shared memory layout (dynamic):
// [0,8) mbarrier | [16, 16+8*(bins+2)) histogram | warp partials | Sobol low direction numbers | table (128-B aligned)
struct SmemLayout {
  uint32_t hist_off, wp_off, sob_off, tbl_off, total;
};
__host__ __device__ inline SmemLayout smem_layout(int bins, uint32_t sob_bytes, uint32_t tbl_bytes) {
  SmemLayout L;
  L.hist_off = 16;
  L.wp_off = L.hist_off + 8u * (uint32_t)(bins + 2);
  L.sob_off = (L.wp_off + 32u * kWarps + 127u) & ~127u;
  L.tbl_off = (L.sob_off + sob_bytes + 127u) & ~127u;
  L.total = L.tbl_off + tbl_bytes;
  return L;
}

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

template <int ILP>
__device__ __forceinline__ void run_compat(const KParams &p, const char *tbl, uint32_t lane_off, const uint32_t (&plo)[ILP],
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
      const float r = *reinterpret_cast<const float *>(tbl + (uint32_t)idx * rep4 + lane_off);
      tot[i] = fmaf(tot[i], r, tot[i]);  // total += total * r (one FFMA, as the reference compiles)
    }
  }
}

__device__ __forceinline__ float boot_step(float tot, uint32_t w, const char *tbl, uint32_t nr4, uint32_t row_mask,
                                           uint32_t lane_off) {
  const uint32_t off = (__umulhi(w, nr4) & row_mask) | lane_off;  // (idx * rep + lane%rep) * 4, idx = mulhi(w, n)
  const float r = *reinterpret_cast<const float *>(tbl + off);
  return fmaf(tot, r, tot);
}

// months 6q..6q+5 of ILP paths: one Philox call each, six table gathers, six FFMAs
template <int ILP, int SHIFT>
__device__ __forceinline__ void run_bootstrap(const KParams &p, uint32_t base, const uint32_t (&plo)[ILP],
                                              uint32_t phi, float (&tot)[ILP]) {
  const BootConsts bc = p.bc;
  const uint32_t one = p.one;
  const int nq = p.n_periods / kFieldsPerCall;
#pragma unroll 1
  for (int q = 0; q < nq; q++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4], m[6];
      stream_words(plo[i], phi, (uint32_t)q, kTagBootstrap, p.keys, w);
      field_mantissas(w, one, m);
#pragma unroll
      for (int j = 0; j < 6; j++) {
        const float r = boot_load<SHIFT>(m[j], j < 4 ? bc.c21 : bc.c22, j < 4 ? bc.k21 : bc.k22, base, bc.row_shift);
        tot[i] = fmaf(tot[i], r, tot[i]);
      }
    }
  }
  const int rem = p.n_periods - nq * kFieldsPerCall;
  if (rem) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4], m[6];
      stream_words(plo[i], phi, (uint32_t)nq, kTagBootstrap, p.keys, w);
      field_mantissas(w, one, m);
#pragma unroll
      for (int j = 0; j < 5; j++)
        if (j < rem) {
          const float r = boot_load<SHIFT>(m[j], j < 4 ? bc.c21 : bc.c22, j < 4 ? bc.k21 : bc.k22, base, bc.row_shift);
          tot[i] = fmaf(tot[i], r, tot[i]);
        }
    }
  }
}

// months 8q..8q+7 of ILP paths: one Philox call each, four Box-Muller pairs
template <int ILP>
__device__ __forceinline__ void run_gaussian(const KParams &p, const uint32_t (&plo)[ILP], uint32_t phi,
                                             float (&tot)[ILP]) {
  const GaussConsts gc = p.gc;
  const uint32_t one = p.one;
  const int nq = p.n_periods / kGaussMonthsPerCall;
#pragma unroll 1
  for (int q = 0; q < nq; q++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4];
      stream_words(plo[i], phi, (uint32_t)q, kTagGaussian, p.keys, w);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float g0, g1;
        gauss_pair(w[j], one, gc, g0, g1);
        tot[i] = __fmul_rn(__fmul_rn(tot[i], g0), g1);
      }
    }
  }
  const int rem = p.n_periods - nq * kGaussMonthsPerCall;
  if (rem) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      uint32_t w[4];
      stream_words(plo[i], phi, (uint32_t)nq, kTagGaussian, p.keys, w);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float g0, g1;
        gauss_pair(w[j], one, gc, g0, g1);
        if (2 * j < rem) tot[i] = __fmul_rn(tot[i], g0);
        if (2 * j + 1 < rem) tot[i] = __fmul_rn(tot[i], g1);
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
template <int MODE>
__device__ __forceinline__ void run_sobol(const KParams &p, const char *tbl, uint32_t lane_off, const uint4 *vlo,
                                          uint32_t h, uint32_t lane, float (&tot)[8]) {
  const uint32_t mk0 = ((lane ^ (lane >> 1)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk1 = (((lane >> 1) ^ (lane >> 2)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk2 = (((lane >> 2) ^ (lane >> 3)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk3 = (((lane >> 3) ^ (lane >> 4)) & 1u) ? 0xffffffffu : 0u;
  const uint32_t mk4 = ((lane >> 4) & 1u) ? 0xffffffffu : 0u;
  const uint32_t gh = h ^ (h >> 1);
  const int ppad = p.sobol_ppad;
  const uint32_t nr4 = p.nr4, row_mask = ~(p.rep4 - 1u);
  const float a = p.gc.a, s2 = p.sigma_sqrt2;
#pragma unroll 1
  for (int m0 = 0; m0 < p.n_periods; m0 += 32) {
    const uint32_t *col = p.sobol_hi + m0 + lane;
    uint32_t U = __ldg(col + 25 * ppad);
    if (h & 1u) U ^= __ldg(col + 24 * ppad);
    for (uint32_t g = gh; g; g &= g - 1u) U ^= __ldg(col + (__ffs((int)g) - 1) * ppad);
    const int mend = (p.n_periods - m0) < 32 ? (p.n_periods - m0) : 32;
#pragma unroll 2
    for (int mi = 0; mi < mend; mi++) {
      const uint32_t Um = __shfl_sync(0xffffffffu, U, mi);
      const uint4 va = vlo[(m0 + mi) * 2], vb = vlo[(m0 + mi) * 2 + 1];
      uint32_t base = Um ^ (va.x & mk0);
      base ^= (va.y & mk1);
      base ^= (va.z & mk2);
      base ^= (va.w & mk3);
      base ^= (vb.x & mk4);
      uint32_t x[8];
      x[0] = base;
      x[1] = base ^ vb.x ^ vb.y;
      x[2] = base ^ vb.y ^ vb.z;
      x[3] = base ^ vb.x ^ vb.z;
      x[4] = base ^ vb.z ^ vb.w;
      x[5] = x[1] ^ vb.z ^ vb.w;
      x[6] = base ^ vb.y ^ vb.w;
      x[7] = base ^ vb.x ^ vb.w;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        if (MODE == SMMC_MODE_SOBOL_BOOTSTRAP) tot[j] = boot_step(tot[j], x[j], tbl, nr4, row_mask, lane_off);
        else tot[j] = __fmul_rn(tot[j], fmaf(erfinv_py(x[j]), s2, a));
      }
    }
  }
}

// ---- the fused persistent kernel ---------------------------------------------------------------------------------

template <int MODE, int ILP, int SHIFT>
__global__ void __launch_bounds__(kThreads, 1) mc_kernel(const __grid_constant__ KParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr bool kHasTable = (MODE == SMMC_MODE_BOOTSTRAP_COMPAT || MODE == SMMC_MODE_BOOTSTRAP ||
                              MODE == SMMC_MODE_SOBOL_BOOTSTRAP);
  constexpr bool kSobol = (MODE == SMMC_MODE_SOBOL_BOOTSTRAP || MODE == SMMC_MODE_SOBOL_GAUSSIAN);
  const SmemLayout L = smem_layout(p.hist_bins, kSobol ? p.sobol_lo_bytes : 0u, kHasTable ? p.tbl_bytes : 0u);
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
  unsigned long long *hist_s = reinterpret_cast<unsigned long long *>(smem + L.hist_off);
  double *wp = reinterpret_cast<double *>(smem + L.wp_off);
  const char *tbl = reinterpret_cast<const char *>(smem + L.tbl_off);
  const uint4 *vlo = reinterpret_cast<const uint4 *>(smem + L.sob_off);

  const int tid = threadIdx.x;
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
  for (int i = tid; i < p.hist_bins + 2; i += kThreads) hist_s[i] = 0ull;
  __syncthreads();
  if (kHasTable || kSobol) mbar_wait(bar, 0);

  const int64_t N = p.n_paths;
  const int64_t nb = (N + kBlock - 1) / kBlock;
  const int64_t total_warps = (int64_t)gridDim.x * kWarps;
  const uint32_t lane_off = (lane * 4u) & (p.rep4 - 1u);
  constexpr bool kCompat = (MODE == SMMC_MODE_BOOTSTRAP_COMPAT);
  // BOOTSTRAP: shared-window address of this lane's column, minus the float exponent bits that boot_load shifts out
  const uint32_t boot_base = smem_u32(tbl) + lane_off - (0x4B000000u << (SHIFT >= 0 ? (uint32_t)SHIFT : p.bc.row_shift));

  Moments acc = {0.0, 0.0, 0.0};
  unsigned int below_lane = 0;
  unsigned long long below_acc = 0ull;

  for (int64_t b = (int64_t)blockIdx.x * kWarps + warp; b < nb; b += total_warps) {
    const uint64_t base = p.path0 + (uint64_t)b * kBlock;  // 256-aligned: low 8 bits free for the in-block offset
    const uint32_t base_lo = (uint32_t)base, phi = (uint32_t)(base >> 32);
    const int64_t rem_paths = N - b * kBlock;
    const int live = rem_paths < kBlock ? (int)rem_paths : kBlock;

    float x[kPathsPerLane];
    if (kSobol) {
#pragma unroll
      for (int j = 0; j < kPathsPerLane; j++) x[j] = p.v0;
      run_sobol<MODE>(p, tbl, lane_off, vlo, (uint32_t)((base + kSobolSkip) >> 8), lane, x);
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
      if (MODE == SMMC_MODE_BOOTSTRAP_COMPAT) run_compat<ILP>(p, tbl, lane_off, plo, tot);
      if (MODE == SMMC_MODE_BOOTSTRAP) run_bootstrap<ILP, SHIFT>(p, boot_base, plo, phi, tot);
      if (MODE == SMMC_MODE_GAUSSIAN) run_gaussian<ILP>(p, plo, phi, tot);
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
        if (p.hist_bins > 0) atomicAdd(&hist_s[hist_bin(x[j], p.hist_lo, p.hist_inv_w, p.hist_bins)], 1ull);
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
    merge(acc, (double)live, mean_live, m2_live);

    if (below_lane > 0x7fffff00u) {  // keep the 32-bit per-lane counter far from wrapping
      below_acc += below_lane;
      below_lane = 0;
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
    for (int w = 0; w < kWarps; w++) {
      merge(c, wp[w * 4 + 0], wp[w * 4 + 1], wp[w * 4 + 2]);
      below += wp[w * 4 + 3];
    }
    double *cp = p.cta_partials + (size_t)blockIdx.x * 4;
    cp[0] = c.n;
    cp[1] = c.mean;
    cp[2] = c.m2;
    cp[3] = below;
  }
  for (int i = tid; i < p.hist_bins + 2; i += kThreads)
    if (hist_s[i]) atomicAdd(&p.hist_g[i], hist_s[i]);
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = (atomicAdd(p.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // last CTA: merge the per-CTA partials in CTA order (warp 0: lane-strided sequential merge, then a fixed
  // butterfly), emit the allreduce-ready statistics vector (include/smmc.h, SMMC_STATS_HEADER)
  if (warp == 0) {
    Moments c = {0.0, 0.0, 0.0};
    double below = 0.0;
    for (unsigned g = lane; g < gridDim.x; g += 32) {
      const double *cp = p.cta_partials + (size_t)g * 4;
      merge(c, __ldcg(cp + 0), __ldcg(cp + 1), __ldcg(cp + 2));
      below += __ldcg(cp + 3);
    }
#pragma unroll
    for (int st = 1; st <= 16; st <<= 1) {
      const double on = __shfl_xor_sync(0xffffffffu, c.n, st);
      const double om = __shfl_xor_sync(0xffffffffu, c.mean, st);
      const double o2 = __shfl_xor_sync(0xffffffffu, c.m2, st);
      const double ob = __shfl_xor_sync(0xffffffffu, below, st);
      // merge in a lane-symmetric order so every lane ends with identical bits: lower lane is always "a"
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
      below += ob;
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
  for (int i = tid; i < p.hist_bins + 2; i += kThreads)
    p.stats[SMMC_STATS_HEADER + i] = (double)__ldcg(&p.hist_g[i]);
}

// ---- table preparation -----------------------------------------------------------------------------------------
__global__ void prep_table_kernel(const float *__restrict__ src, int n, int rep, float *__restrict__ dst,
                                  uint32_t dst_floats) {
  for (uint32_t o = blockIdx.x * blockDim.x + threadIdx.x; o < dst_floats; o += gridDim.x * blockDim.x) {
    const uint32_t row = o / (uint32_t)rep;
    dst[o] = row < (uint32_t)n ? src[row] * 0.01f : 0.0f;  // reference src/simulations.cu:126: returns[i] * float(0.01)
  }
}

// stable rank sort, O(n^2): n is a table length (~1e3), once per call
__global__ void sort_table_kernel(const float *__restrict__ src, int n, float *__restrict__ dst) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float v = src[i];
    int rank = 0;
    for (int j = 0; j < n; j++) {
      const float u = __ldg(src + j);
      rank += (u < v || (u == v && j < i)) ? 1 : 0;
    }
    dst[rank] = v;
  }
}

// ---- introspection kernels ----------------------------------------------------------------------------------------
__global__ void debug_draws_kernel(int mode, const __grid_constant__ KParams p, int n_paths,
                                   const uint32_t *__restrict__ sobol_full, const uint32_t *__restrict__ shift,
                                   uint32_t *out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_paths) return;
  const uint64_t path = p.path0 + (uint64_t)t;
  const uint32_t plo = (uint32_t)path, phi = (uint32_t)(path >> 32);
  uint32_t *o = out + (size_t)t * p.n_periods;
  if (mode == SMMC_MODE_BOOTSTRAP_COMPAT) {
    uint32_t st = rand_pcg(plo + 1u + p.compat_seed);
    for (int m = 0; m < p.n_periods; m++) {
      st = xorshift(st);
      o[m] = (uint32_t)(int)(__uint2float_rn(st) * p.idx_scale);
    }
  } else if (mode == SMMC_MODE_BOOTSTRAP) {  // the index as the simulation kernel computes it (fma.rz floor)
    uint32_t w[4], mt[6];
    for (int m = 0; m < p.n_periods; m++) {
      const int j = m % 6;
      if (j == 0) {
        stream_words(plo, phi, (uint32_t)(m / 6), kTagBootstrap, p.keys, w);
        field_mantissas(w, p.one, mt);
      }
      const float y = __fmaf_rz(__uint_as_float(mt[j]), j < 4 ? p.bc.c21 : p.bc.c22, j < 4 ? p.bc.k21 : p.bc.k22);
      o[m] = __float_as_uint(y) - 0x4B000000u;
    }
  } else if (mode == SMMC_MODE_GAUSSIAN) {  // even month: radius field, odd month: the hashed angle word
    uint32_t w[4];
    for (int m = 0; m < p.n_periods; m++) {
      const int j = m % 8;
      if (j == 0) stream_words(plo, phi, (uint32_t)(m / 8), kTagGaussian, p.keys, w);
      o[m] = (j & 1) ? w[j >> 1] * kAngleHash : (w[j >> 1] & 0xFFFFFu);
    }
  } else {  // Sobol modes: plain Gray-code evaluation over all 32 direction numbers of dimension m
    const uint64_t i = path + kSobolSkip;
    const uint32_t g = (uint32_t)(i ^ (i >> 1));
    for (int m = 0; m < p.n_periods; m++) {
      uint32_t x = shift[m];
      for (uint32_t gg = g; gg; gg &= gg - 1u) x ^= sobol_full[m * 32 + (__ffs((int)gg) - 1)];
      o[m] = mode == SMMC_MODE_SOBOL_BOOTSTRAP ? __umulhi(x, (uint32_t)p.n_returns) : x;
    }
  }
}

__global__ void debug_philox_kernel(const uint32_t *ctr, PhiloxKeys keys, int n, uint32_t *out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  uint32_t w[4];
  philox4x32_10(ctr[4 * t + 0], ctr[4 * t + 1], ctr[4 * t + 2], ctr[4 * t + 3], keys, w);
  for (int j = 0; j < 4; j++) out[4 * t + j] = w[j];
}

// ---- reduce_mean / vector_add ------------------------------------------------------------------------------------
// single-pass sum: float4 grid-stride loads, fp64 accumulation, block partials + last-block-done merge
__global__ void __launch_bounds__(512) reduce_sum_kernel(const float *__restrict__ v, int64_t n, double *partials,
                                                         unsigned int *ticket, double *sum_out) {
  double acc = 0.0;
  const int64_t n4 = n >> 2;
  const float4 *v4 = reinterpret_cast<const float4 *>(v);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 q = __ldg(v4 + i);
    acc += ((double)q.x + (double)q.y) + ((double)q.z + (double)q.w);
  }
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc += (double)v[i];
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, st);
  __shared__ double ws[16];
  __shared__ bool last;
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += ws[w];
    partials[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double t = 0.0;
    for (unsigned g = 0; g < gridDim.x; g++) t += __ldcg(partials + g);
    *sum_out = t;
  }
}

__global__ void vector_add_kernel(float *__restrict__ out, const float *__restrict__ a, const float *__restrict__ b,
                                  int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = a[i] + b[i];
}

template <int MODE, int ILP, int SHIFT = -1>
cudaError_t launch_one(const KParams &p, const LaunchInfo &li, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(mc_kernel<MODE, ILP, SHIFT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)li.smem_bytes);
  if (e != cudaSuccess) return e;
  mc_kernel<MODE, ILP, SHIFT><<<li.grid, li.threads, li.smem_bytes, s>>>(p);
  return cudaGetLastError();
}

template <int MODE>
cudaError_t launch_mode(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  switch (variant) {
    case 1: return launch_one<MODE, 1>(p, li, s);
    case 2: return launch_one<MODE, 2>(p, li, s);
    case 8: return launch_one<MODE, 8>(p, li, s);
    case 4:
    default: return launch_one<MODE, 4>(p, li, s);
  }
}

// BOOTSTRAP: the common case (32 bank-private copies per row: 128-byte rows) gets the row shift as an immediate
cudaError_t launch_bootstrap(const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  if (li.rep != 32) return launch_mode<SMMC_MODE_BOOTSTRAP>(p, li, variant, s);
  switch (variant) {
    case 1: return launch_one<SMMC_MODE_BOOTSTRAP, 1, 7>(p, li, s);
    case 2: return launch_one<SMMC_MODE_BOOTSTRAP, 2, 7>(p, li, s);
    case 8: return launch_one<SMMC_MODE_BOOTSTRAP, 8, 7>(p, li, s);
    case 4:
    default: return launch_one<SMMC_MODE_BOOTSTRAP, 4, 7>(p, li, s);
  }
}

}  // namespace

bool plan_launch(int mode, int n_returns, int n_periods, int hist_bins, int n_sms, size_t smem_optin, int variant,
                 LaunchInfo *li) {
  (void)variant;
  const bool sobol = (mode == SMMC_MODE_SOBOL_BOOTSTRAP || mode == SMMC_MODE_SOBOL_GAUSSIAN);
  const uint32_t sob_bytes = sobol ? (uint32_t)(((n_periods + 31) & ~31) * 32) : 0u;
  const bool has_table = (mode == SMMC_MODE_BOOTSTRAP_COMPAT || mode == SMMC_MODE_BOOTSTRAP ||
                          mode == SMMC_MODE_SOBOL_BOOTSTRAP);
  li->grid = n_sms;
  li->threads = kThreads;
  li->rep = 0;
  uint32_t tbl_bytes = 0;
  if (has_table) {
    int rep = 32;
    for (; rep >= 1; rep >>= 1) {
      tbl_bytes = (uint32_t)((((size_t)(n_returns + 1) * rep * 4) + 15) & ~(size_t)15);
      if (smem_layout(hist_bins, sob_bytes, tbl_bytes).total <= smem_optin) break;
    }
    if (rep < 1) return false;
    li->rep = rep;
  }
  li->smem_bytes = smem_layout(hist_bins, sob_bytes, tbl_bytes).total;
  return li->smem_bytes <= smem_optin;
}

void launch_prep_table(const float *returns_dev, int n, int rep, float *dst, uint32_t dst_floats, cudaStream_t s) {
  const int threads = 256;
  const int blocks = (int)((dst_floats + threads - 1) / threads);
  prep_table_kernel<<<blocks > 1184 ? 1184 : blocks, threads, 0, s>>>(returns_dev, n, rep, dst, dst_floats);
}

void launch_sort_table(const float *src, int n, float *dst, cudaStream_t s) {
  sort_table_kernel<<<(n + 127) / 128 > 1184 ? 1184 : (n + 127) / 128, 128, 0, s>>>(src, n, dst);
}

cudaError_t launch_mc(int mode, const KParams &p, const LaunchInfo &li, int variant, cudaStream_t s) {
  switch (mode) {
    case SMMC_MODE_SOBOL_BOOTSTRAP: return launch_one<SMMC_MODE_SOBOL_BOOTSTRAP, 8>(p, li, s);
    case SMMC_MODE_SOBOL_GAUSSIAN: return launch_one<SMMC_MODE_SOBOL_GAUSSIAN, 8>(p, li, s);
    case SMMC_MODE_BOOTSTRAP_COMPAT: return launch_mode<SMMC_MODE_BOOTSTRAP_COMPAT>(p, li, variant, s);
    case SMMC_MODE_BOOTSTRAP: return launch_bootstrap(p, li, variant, s);
    case SMMC_MODE_GAUSSIAN: return launch_mode<SMMC_MODE_GAUSSIAN>(p, li, variant, s);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_debug_draws(int mode, const KParams &p, int n_paths, const uint32_t *sobol_full,
                               const uint32_t *shift, uint32_t *out_dev, cudaStream_t s) {
  if (n_paths <= 0) return cudaSuccess;
  debug_draws_kernel<<<(n_paths + 127) / 128, 128, 0, s>>>(mode, p, n_paths, sobol_full, shift, out_dev);
  return cudaGetLastError();
}

cudaError_t launch_debug_philox(const uint32_t *ctr_dev, PhiloxKeys keys, int n, uint32_t *out_dev, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  debug_philox_kernel<<<(n + 127) / 128, 128, 0, s>>>(ctr_dev, keys, n, out_dev);
  return cudaGetLastError();
}

cudaError_t launch_reduce_sum(const float *v_dev, int64_t n, double *partials_dev, unsigned int *ticket_dev,
                              double *sum_dev, int n_sms, cudaStream_t s) {
  reduce_sum_kernel<<<n_sms * 4, 512, 0, s>>>(v_dev, n, partials_dev, ticket_dev, sum_dev);
  return cudaGetLastError();
}

cudaError_t launch_vector_add(float *out, const float *a, const float *b, int64_t n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  vector_add_kernel<<<(int)blocks, 256, 0, s>>>(out, a, b, n);
  return cudaGetLastError();
}

}  // namespace smmc
