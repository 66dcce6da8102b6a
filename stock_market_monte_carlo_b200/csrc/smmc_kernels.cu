// smmc_kernels.cu -- small kernels around the Monte Carlo path (table preparation, introspection, reduce_mean,
// vector_add), the launch planner and the mode dispatch.  The fused simulation kernel itself is the template in
// smmc_mc_kernel.cuh, instantiated per RNG mode in smmc_mc_{compat,bootstrap,gaussian,sobol}.cu.
//
// Replaces compute_sum_kernel / impl_vector_add_gpu (reference src/simulations.cu:249-267, src/gpu.cu:8-14).
#include <cstring>

#include "smmc_kernels.cuh"

#include "../../include/smmc.h"

namespace smmc {

// per-mode launchers (smmc_mc_*.cu)
cudaError_t launch_mc_compat(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s);
cudaError_t launch_mc_bootstrap(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s);
cudaError_t launch_mc_gaussian(const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s);
cudaError_t launch_mc_sobol(int mode, const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s);

namespace {

// ---- table preparation -----------------------------------------------------------------------------------------
__global__ void prep_table_kernel(const float *__restrict__ src, int n, int rep, float *__restrict__ dst,
                                  uint32_t dst_floats) {
  for (uint32_t o = blockIdx.x * blockDim.x + threadIdx.x; o < dst_floats; o += gridDim.x * blockDim.x) {
    const uint32_t row = o / (uint32_t)rep;
    dst[o] = row < (uint32_t)n ? src[row] * 0.01f : 0.0f;  // reference src/simulations.cu:126: returns[i] * float(0.01)
  }
}

// stable rank sort, O(n^2): n is a table length (~1e3), once per call.  Ranks come from the order-preserving integer keys
// (ordered_key: a total order, -0 < +0, NaNs beyond +-inf), so every rank 0..n-1 is assigned exactly once even for a table
// with NaNs (device-pointer callers; the host-pointer path rejects non-finite returns in validate()).
// One warp per element: the lanes share the scan over the table (coalesced 128-byte loads), the counts are summed by
// shuffles.  (One thread per element took 18 us at n = 1200 -- 1200 dependent-latency loads -- against a 1.4 ms kernel.)
__global__ void sort_table_kernel(const float *__restrict__ src, int n, float *__restrict__ dst) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
    const float v = __ldg(src + i);
    const uint32_t kv = ordered_key(v);
    int rank = 0;
    for (int j = lane; j < n; j += 32) {
      const uint32_t ku = ordered_key(__ldg(src + j));
      rank += (ku < kv || (ku == kv && j < i)) ? 1 : 0;
    }
#pragma unroll
    for (int st = 16; st >= 1; st >>= 1) rank += __shfl_xor_sync(0xffffffffu, rank, st);
    if (lane == 0) dst[rank] = v;
  }
}

// scale * (cos, sin) of the cell centres 2 pi (i + 1/2) / kSinCosCells, evaluated in double and rounded once (sincos_tbl).
// scale = sqrt(-k) = std/100 * sqrt(2 ln 2): the Box-Muller radius of a table pair is then sqrt(-lg2 u) with the negation
// folded into the MUFU operand -- the FMUL by k of the MUFU pairs is not needed (gauss_pair_tbl).
__global__ void sincos_table_kernel(float *__restrict__ dst, double scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kSinCosCells) return;
  double sv, cv;
  sincospi(2.0 * ((double)i + 0.5) / (double)kSinCosCells, &sv, &cv);
  const float c = (float)(scale * cv), sn = (float)(scale * sv);
  const uint32_t slot = sincos_slot((uint32_t)i);  // bank swizzle, see sincos_tbl
  dst[2 * slot] = c;
  dst[2 * slot + 1] = sn;
  float *lin = dst + 2 * kSinCosCells;  // the linear copy GAUSSIAN reads
  lin[2 * i] = c;
  lin[2 * i + 1] = sn;
}

// ---- trajectory mode ------------------------------------------------------------------------------------------------
// One thread per path, the same device functions and the same operation order as the statistics kernel, so that the
// last value of a trajectory is bit-identical to that path's final in mc_kernel.  tbl = scaled table ([n+1], entry n
// a zero pad; sorted for SOBOL_BOOTSTRAP), raw = the same in percent (returns_pct output).
__global__ void trajectory_kernel(int mode, const __grid_constant__ KParams p, int n_paths, const float *__restrict__ tbl,
                                  const float *__restrict__ raw, const uint32_t *__restrict__ sobol_full,
                                  const uint32_t *__restrict__ shift, float *__restrict__ values,
                                  float *__restrict__ rets) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_paths) return;
  const int P = p.n_periods;
  const uint64_t path = p.path0 + (uint64_t)t;
  const uint32_t plo = (uint32_t)path, phi = (uint32_t)(path >> 32);
  float *v = values + (size_t)t * (P + 1);
  float *rp = rets ? rets + (size_t)t * P : nullptr;
  float tot = p.v0;
  v[0] = tot;
  const int n = p.n_returns;
  // withdrawals (same formula and order as Flow<true> in smmc_mc_kernel.cuh)
  auto flow = [&](float x, int m) { return p.flow_q ? fmaxf(fmaf(x, __ldg(p.flow_q + m), -p.flow_w), 0.0f) : x; };
  if (mode == SMMC_MODE_BOOTSTRAP_COMPAT) {
    uint32_t st = rand_pcg(plo + 1u + p.compat_seed);
    for (int m = 0; m < P; m++) {
      st = xorshift(st);
      const int idx = (int)(__uint2float_rn(st) * p.idx_scale);
      tot = flow(fmaf(tot, tbl[idx], tot), m);
      v[m + 1] = tot;
      if (rp) rp[m] = idx < n ? raw[idx] : 0.0f;
    }
  } else if (mode == SMMC_MODE_BOOTSTRAP) {
    uint32_t w[4], mt[6];
    for (int m = 0; m < P; m++) {
      const int j = m % kFieldsPerCall;
      if (j == 0) {
        stream_words_rt(p.philox_rounds, plo, phi, (uint32_t)(m / kFieldsPerCall), kTagBootstrap, p.keys, w);
        field_mantissas(w, p.one, mt);
      }
      const float y = __fmaf_rz(__uint_as_float(mt[j]), j < 4 ? p.bc.c21 : p.bc.c22, j < 4 ? p.bc.k21 : p.bc.k22);
      const uint32_t idx = __float_as_uint(y) - 0x4B000000u;
      tot = flow(fmaf(tot, tbl[idx], tot), m);
      v[m + 1] = tot;
      if (rp) rp[m] = raw[idx];
    }
  } else if (mode == SMMC_MODE_GAUSSIAN) {
    uint32_t w[4];
    float g[2] = {1.0f, 1.0f};
    for (int m = 0; m < P; m++) {
      const int j = m % kGaussMonthsPerCall;
      if (j == 0) stream_words_rt(p.philox_rounds, plo, phi, (uint32_t)(m / kGaussMonthsPerCall), kTagGaussian, p.keys, w);
      if ((j & 1) == 0) {  // the same evaluation of sin/cos as the statistics kernel's pair j/2 (table or MUFU)
        if ((p.gauss_tbl_mask >> (j >> 1)) & 1) gauss_pair_tbl(w[j >> 1], p.one, p.cell_mask, p.gc, reinterpret_cast<const char *>(p.tbl_rep), g[0], g[1]);
        else gauss_pair(w[j >> 1], p.one, p.gc, g[0], g[1]);
      }
      const float gm = g[j & 1];
      tot = flow(__fmul_rn(tot, gm), m);
      v[m + 1] = tot;
      if (rp) rp[m] = __fmul_rn(__fsub_rn(gm, 1.0f), 100.0f);
    }
  } else {  // Sobol modes: plain Gray-code evaluation over all 32 direction numbers of dimension m
    const uint64_t i = path + kSobolSkip;
    const uint32_t gray = (uint32_t)(i ^ (i >> 1));
    auto word = [&](int m) {
      uint32_t x = shift[m];
      for (uint32_t gg = gray; gg; gg &= gg - 1u) x ^= sobol_full[m * 32 + (__ffs((int)gg) - 1)];
      return x;
    };
    float g[2] = {1.0f, 1.0f};
    for (int m = 0; m < P; m++) {
      const uint32_t x = word(m);
      if (mode == SMMC_MODE_SOBOL_BOOTSTRAP) {
        const uint32_t idx = __umulhi(x, (uint32_t)n);
        tot = flow(fmaf(tot, tbl[idx], tot), m);
        if (rp) rp[m] = raw[idx];
      } else {  // Box-Muller on dimension pairs; an unpaired last month takes the inverse CDF (smmc_device.cuh)
        if ((m & 1) == 0) {
          if (m + 1 < P) {
            if ((p.gauss_tbl_mask >> ((uint32_t)(path >> 5) & 7u)) & 1)  // the statistics kernel's choice for this path slot
              sobol_gauss_pair_tbl(x, word(m + 1), p.one, p.gc, reinterpret_cast<const char *>(p.tbl_rep), g[0], g[1]);
            else sobol_gauss_pair(x, word(m + 1), p.gc, g[0], g[1]);
          } else {
            g[0] = fmaf(erfinv_py(x), p.sigma_sqrt2, p.gc.a);
          }
        }
        const float gm = g[m & 1];
        tot = flow(__fmul_rn(tot, gm), m);
        if (rp) rp[m] = __fmul_rn(__fsub_rn(gm, 1.0f), 100.0f);
      }
      v[m + 1] = tot;
    }
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
        stream_words_rt(p.philox_rounds, plo, phi, (uint32_t)(m / 6), kTagBootstrap, p.keys, w);
        field_mantissas(w, p.one, mt);
      }
      const float y = __fmaf_rz(__uint_as_float(mt[j]), j < 4 ? p.bc.c21 : p.bc.c22, j < 4 ? p.bc.k21 : p.bc.k22);
      o[m] = __float_as_uint(y) - 0x4B000000u;
    }
  } else if (mode == SMMC_MODE_GAUSSIAN) {  // even month: radius field, odd month: the hashed angle word
    uint32_t w[4];
    for (int m = 0; m < p.n_periods; m++) {
      const int j = m % 8;
      if (j == 0) stream_words_rt(p.philox_rounds, plo, phi, (uint32_t)(m / 8), kTagGaussian, p.keys, w);
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

__global__ void debug_philox_kernel(const uint32_t *ctr, PhiloxKeys keys, int rounds, int n, uint32_t *out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  uint32_t w[4];
  if (rounds == 7) philox4x32<7>(ctr[4 * t + 0], ctr[4 * t + 1], ctr[4 * t + 2], ctr[4 * t + 3], keys, w);
  else philox4x32<10>(ctr[4 * t + 0], ctr[4 * t + 1], ctr[4 * t + 2], ctr[4 * t + 3], keys, w);
  for (int j = 0; j < 4; j++) out[4 * t + j] = w[j];
}

// ---- radix select over collected candidate keys ---------------------------------------------------------------------
// The second select replay appends the ordered keys under the targets' 11-bit prefixes to a buffer (mc_kernel epilogue);
// the remaining key bits are resolved here over that buffer -- an HBM-bound pass of 4 B per key instead of a third
// simulation replay.  Same histogram layout as the replay passes: hist[t][digit] over the keys matching prefix[t].
struct CandParams {
  int targets, digit_shift, prefix_shift;
  uint32_t digit_mask;
  uint32_t prefix[kSelectMaxTargets];
};
__global__ void __launch_bounds__(512) select_candidates_kernel(const uint32_t *__restrict__ keys,
                                                                const unsigned long long *__restrict__ count_dev,
                                                                unsigned long long capacity, CandParams cp,
                                                                unsigned long long *__restrict__ hist) {
  extern __shared__ unsigned int cand_s[];  // [targets][kSelectBins]
  for (int i = threadIdx.x; i < cp.targets * kSelectBins; i += blockDim.x) cand_s[i] = 0u;
  __syncthreads();
  unsigned long long n = *count_dev;
  if (n > capacity) n = capacity;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = __ldcs(keys + i);
    const uint32_t digit = (key >> cp.digit_shift) & cp.digit_mask;
    const uint32_t pre = cp.prefix_shift >= 32 ? 0u : key >> cp.prefix_shift;
    for (int t = 0; t < cp.targets; t++)
      if (pre == cp.prefix[t]) atomicAdd(&cand_s[t * kSelectBins + digit], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cp.targets * kSelectBins; i += blockDim.x)
    if (cand_s[i]) atomicAdd(&hist[i], (unsigned long long)cand_s[i]);
}

// block -> grid stage of the sum: warp shuffle, block partial, last-block-done merge in block order (deterministic bits)
__device__ __forceinline__ void reduce_sum_epilogue(double acc, double *partials, unsigned int *ticket, double *sum_out) {
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, st);
  __shared__ double ws[32];
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

// ---- reduce_mean / vector_add ------------------------------------------------------------------------------------
// Single-pass sum of n floats (replaces the reference's multi-launch strided compute_sum_kernel, src/simulations.cu:249-341).
// HBM-bound: 4 B per element and nothing else.  Streaming 16-byte loads (ld.global.cs: read once, evict first), four
// independent loads and fp64 accumulators per thread per iteration (148 x 4 CTAs x 512 threads x 64 B = 19 MB in flight),
// any 4-byte-aligned pointer (scalar head up to the first 16-byte boundary), block partials + last-block-done merge in
// block order (deterministic bits).
__global__ void __launch_bounds__(512) reduce_sum_kernel(const float *__restrict__ v, int64_t n, double *partials,
                                                         unsigned int *ticket, double *sum_out) {
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t head = (int64_t)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(v) & 15u)) & 15u) >> 2);
  if (head > n) head = n;
#ifndef SMMC_RED_UNROLL
#define SMMC_RED_UNROLL 2
#endif
#ifndef SMMC_RED_LD
#define SMMC_RED_LD __ldcs
#endif
  constexpr int U = SMMC_RED_UNROLL;
  double a[U];
#pragma unroll
  for (int k = 0; k < U; k++) a[k] = 0.0;
  if (gtid < head) a[0] += (double)v[gtid];
  const float4 *v4 = reinterpret_cast<const float4 *>(v + head);
  const int64_t n4 = (n - head) >> 2;
  int64_t i = gtid;
  for (; i + (U - 1) * stride < n4; i += U * stride) {
    float4 q[U];
#pragma unroll
    for (int k = 0; k < U; k++) q[k] = SMMC_RED_LD(v4 + i + k * stride);
#pragma unroll
    for (int k = 0; k < U; k++) a[k] += ((double)q[k].x + (double)q[k].y) + ((double)q[k].z + (double)q[k].w);
  }
  for (; i < n4; i += stride) {
    const float4 q = SMMC_RED_LD(v4 + i);
    a[0] += ((double)q.x + (double)q.y) + ((double)q.z + (double)q.w);
  }
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < U; k++) acc += a[k];
  for (int64_t t = head + (n4 << 2) + gtid; t < n; t += stride) acc += (double)v[t];
  reduce_sum_epilogue(acc, partials, ticket, sum_out);
}

// The same sum for large inputs, B200-style: every CTA streams its 32 KB chunks through a 3-stage shared-memory ring
// filled by TMA bulk copies (cp.async.bulk + mbarrier complete_tx; UBLKCP in SASS), 2 CTAs per SM = 192 KB of loads in
// flight per SM that cost no registers and no LSU address instructions; the threads read the landed chunk with
// conflict-free LDS.128 and accumulate in fp64.  Chunk c of the 16-byte-aligned middle part goes to CTA c % grid; the
// unaligned head (< 4 floats) and the tail (< one chunk) are read with plain loads.
constexpr int kRedThreads = 512;
constexpr int kRedStages = 3;
constexpr int kRedChunkFloats = 8192;  // 32 KB
constexpr size_t kRedSmemBytes = 128 + (size_t)kRedStages * kRedChunkFloats * 4;

__global__ void __launch_bounds__(kRedThreads, 2) reduce_sum_tma_kernel(const float *__restrict__ v, int64_t n, double *partials,
                                                                        unsigned int *ticket, double *sum_out) {
  extern __shared__ __align__(128) unsigned char red_smem[];
  uint64_t *bar = reinterpret_cast<uint64_t *>(red_smem);
  float4 *buf = reinterpret_cast<float4 *>(red_smem + 128);
  const int tid = threadIdx.x;
  const int64_t G = gridDim.x, b = blockIdx.x;
  int64_t head = (int64_t)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(v) & 15u)) & 15u) >> 2);
  if (head > n) head = n;
  const float *mid = v + head;
  const int64_t n_chunks = (n - head) / kRedChunkFloats;
  const int64_t my = n_chunks > b ? (n_chunks - b + G - 1) / G : 0;  // chunks b, b + G, ...
  if (tid == 0) {
    for (int s = 0; s < kRedStages; s++) mbar_init(bar + s, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    for (int64_t k = 0; k < my && k < kRedStages; k++) {
      mbar_arrive_expect_tx(bar + k, kRedChunkFloats * 4);
      tma_bulk_g2s(buf + k * (kRedChunkFloats / 4), mid + (b + k * G) * kRedChunkFloats, kRedChunkFloats * 4, bar + k);
    }
  }
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  for (int64_t k = 0; k < my; k++) {
    const int s = (int)(k % kRedStages);
    mbar_wait(bar + s, (uint32_t)((k / kRedStages) & 1));
    const float4 *src = buf + s * (kRedChunkFloats / 4);
    const float4 q0 = src[tid], q1 = src[tid + kRedThreads], q2 = src[tid + 2 * kRedThreads], q3 = src[tid + 3 * kRedThreads];
    a0 += ((double)q0.x + (double)q0.y) + ((double)q0.z + (double)q0.w);
    a1 += ((double)q1.x + (double)q1.y) + ((double)q1.z + (double)q1.w);
    a2 += ((double)q2.x + (double)q2.y) + ((double)q2.z + (double)q2.w);
    a3 += ((double)q3.x + (double)q3.y) + ((double)q3.z + (double)q3.w);
    __syncthreads();  // every thread has read stage s: it can be refilled
    if (tid == 0 && k + kRedStages < my) {
      mbar_arrive_expect_tx(bar + s, kRedChunkFloats * 4);
      tma_bulk_g2s(buf + s * (kRedChunkFloats / 4), mid + (b + (k + kRedStages) * G) * kRedChunkFloats, kRedChunkFloats * 4,
                   bar + s);
    }
  }
  double acc = (a0 + a1) + (a2 + a3);
  // head and tail (fewer than 4 + 8192 floats in total): plain loads, grid-strided
  const int64_t gtid = b * kRedThreads + tid, stride = G * kRedThreads;
  if (gtid < head) acc += (double)v[gtid];
  for (int64_t t = head + n_chunks * kRedChunkFloats + gtid; t < n; t += stride) acc += (double)v[t];
  reduce_sum_epilogue(acc, partials, ticket, sum_out);
}

// out = a + b (reference src/gpu.cu:8-14, one element per thread of 1024-thread blocks).  HBM-bound, 12 B per element:
// 16-byte streaming loads / stores (read once, written once: evict-first), two independent quads per thread per
// iteration, grid = 4 CTAs per SM; pointers that are not all 16-byte aligned take the scalar loop.
__global__ void __launch_bounds__(512) vector_add_kernel(float *__restrict__ out, const float *__restrict__ a,
                                                         const float *__restrict__ b, int64_t n, int vec) {
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t done = 0;
  if (vec) {
    const float4 *a4 = reinterpret_cast<const float4 *>(a), *b4 = reinterpret_cast<const float4 *>(b);
    float4 *o4 = reinterpret_cast<float4 *>(out);
    const int64_t n4 = n >> 2;
    int64_t i = gtid;
    for (; i + stride < n4; i += 2 * stride) {
      const float4 x0 = __ldcs(a4 + i), y0 = __ldcs(b4 + i), x1 = __ldcs(a4 + i + stride), y1 = __ldcs(b4 + i + stride);
      __stcs(o4 + i, make_float4(x0.x + y0.x, x0.y + y0.y, x0.z + y0.z, x0.w + y0.w));
      __stcs(o4 + i + stride, make_float4(x1.x + y1.x, x1.y + y1.y, x1.z + y1.z, x1.w + y1.w));
    }
    for (; i < n4; i += stride) {
      const float4 x = __ldcs(a4 + i), y = __ldcs(b4 + i);
      __stcs(o4 + i, make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w));
    }
    done = n4 << 2;
  }
  for (int64_t i = done + gtid; i < n; i += stride) out[i] = a[i] + b[i];
}

}  // namespace

// Warps per CTA: every warp walks whole 256-path blocks, so a launch takes ceil(nb / (grid * warps)) waves and the
// last wave may be nearly empty (5 M paths = 19 531 blocks = 4.12 waves of 148 x 32 warps: 82 % efficient).  Fewer
// warps per CTA can fill the last wave (27 warps: 4.89 of 5 waves); fewer resident warps cost some latency
// hiding, hence the mild penalty.  Large launches (tens of waves) keep 32.
int plan_warps(int64_t n_paths, int n_sms, bool ticket) {
  const int64_t nb = (n_paths + kBlock - 1) / kBlock;
  if (nb <= 0) return kWarps;
  int best = kWarps;
  double best_score = 0.0;
  // ticket launches hand blocks out dynamically, so only the warp count per scheduler matters: multiples of 4 keep the four
  // sub-partitions even (measured at 5 M x 1000 Sobol: 28 warps 1.385 ms, 27: 1.439, 30: 1.443, 32: 1.444, 24: 1.405)
  for (int w = kWarps; w >= (ticket ? 24 : 20); w -= (ticket ? 4 : 1)) {
    const int64_t slots = (int64_t)n_sms * w;
    const int64_t waves = (nb + slots - 1) / slots;
    const double eff = (double)nb / (double)(slots * waves);
    const double score = eff * (0.55 + 0.45 * (double)w / kWarps);
    if (score > best_score * 1.02) {  // prefer more warps unless the gain is clear
      best_score = score;
      best = w;
    }
  }
  return best;
}

// Blocks are handed out by ticket when the launch is more than one wave and fewer than kTicketMaxWaves waves of 32-warp
// CTAs: there the tail (warps that run dry while others still hold a block) is a visible share of the kernel (5 M x 1000:
// -5 %; 20 M x 360: -2 %; 100 M x 360, 82 waves: +-0.1 %, kept static).
bool plan_ticket(int64_t n_paths, int n_sms, int ticket_override) {
  if (ticket_override == 0 || ticket_override == 1) return ticket_override == 1;
  const int64_t nb = (n_paths + kBlock - 1) / kBlock;
  const int64_t per_wave = (int64_t)n_sms * kWarps;
  return nb > per_wave && nb < (int64_t)kTicketMaxWaves * per_wave;
}

bool plan_launch(int mode, int n_returns, int n_periods, int hist_bins, int sel_targets, bool withdrawals, bool sincos,
                 int64_t n_paths, int n_sms, size_t smem_optin, int warps_override, int ticket_override, LaunchInfo *li) {
  const uint32_t sel_bytes = (uint32_t)sel_targets * kSelectBins * 4u;
  if (withdrawals && (size_t)n_periods * 4u > smem_optin) return false;  // (size_t: n_periods * 4 must not wrap)
  const uint32_t flow_bytes = withdrawals ? (uint32_t)n_periods * 4u : 0u;
  const bool sobol = (mode == SMMC_MODE_SOBOL_BOOTSTRAP || mode == SMMC_MODE_SOBOL_GAUSSIAN);
  const uint32_t sob_bytes = sobol ? (uint32_t)(((n_periods + 31) & ~31) * 32) : 0u;
  const bool has_table = (mode == SMMC_MODE_BOOTSTRAP_COMPAT || mode == SMMC_MODE_BOOTSTRAP ||
                          mode == SMMC_MODE_SOBOL_BOOTSTRAP);
  li->grid = n_sms;
  li->ticket = plan_ticket(n_paths, n_sms, ticket_override);
  li->threads = 32 * ((warps_override >= 1 && warps_override <= kWarps) ? warps_override : plan_warps(n_paths, n_sms, li->ticket));
  li->rep = 0;
  uint32_t tbl_bytes = 0;
  if (has_table) {
    int rep = 32;
    for (; rep >= 1; rep >>= 1) {
      tbl_bytes = (uint32_t)((((size_t)(n_returns + 1) * rep * 4) + 15) & ~(size_t)15);
      if (smem_layout(hist_bins, sel_bytes, flow_bytes, sob_bytes, tbl_bytes).total <= smem_optin) break;
    }
    if (rep < 1) return false;
    li->rep = rep;
  } else if (sincos && (mode == SMMC_MODE_GAUSSIAN || mode == SMMC_MODE_SOBOL_GAUSSIAN)) {
    tbl_bytes = kSinCosBytes;
  }
  li->smem_bytes = smem_layout(hist_bins, sel_bytes, flow_bytes, sob_bytes, tbl_bytes).total;
  return li->smem_bytes <= smem_optin;
}

void launch_prep_table(const float *returns_dev, int n, int rep, float *dst, uint32_t dst_floats, cudaStream_t s) {
  const int threads = 256;
  const int blocks = (int)((dst_floats + threads - 1) / threads);
  prep_table_kernel<<<blocks > 1184 ? 1184 : blocks, threads, 0, s>>>(returns_dev, n, rep, dst, dst_floats);
}

void launch_sort_table(const float *src, int n, float *dst, cudaStream_t s) {
  const int blocks = (n + 7) / 8;  // 8 warps per block, one element per warp
  sort_table_kernel<<<blocks > 1184 ? 1184 : blocks, 256, 0, s>>>(src, n, dst);
}

cudaError_t launch_mc(int mode, const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s) {
  switch (mode) {
    case SMMC_MODE_SOBOL_BOOTSTRAP:
    case SMMC_MODE_SOBOL_GAUSSIAN: return launch_mc_sobol(mode, p, li, v, s);
    case SMMC_MODE_BOOTSTRAP_COMPAT: return launch_mc_compat(p, li, v, s);
    case SMMC_MODE_BOOTSTRAP: return launch_mc_bootstrap(p, li, v, s);
    case SMMC_MODE_GAUSSIAN: return launch_mc_gaussian(p, li, v, s);
    default: return cudaErrorInvalidValue;
  }
}

void launch_sincos_table(float *dst, double scale, cudaStream_t s) { sincos_table_kernel<<<kSinCosCells / 256, 256, 0, s>>>(dst, scale); }

cudaError_t launch_select_candidates(const uint32_t *keys_dev, const unsigned long long *count_dev, unsigned long long capacity,
                                     int targets, int digit_shift, uint32_t digit_mask, int prefix_shift,
                                     const uint32_t *prefix, unsigned long long *hist_dev, int n_sms, cudaStream_t s) {
  CandParams cp;
  cp.targets = targets;
  cp.digit_shift = digit_shift;
  cp.prefix_shift = prefix_shift;
  cp.digit_mask = digit_mask;
  for (int t = 0; t < kSelectMaxTargets; t++) cp.prefix[t] = prefix[t];
  const size_t smem = (size_t)targets * kSelectBins * sizeof(unsigned int);
  cudaError_t e = cudaFuncSetAttribute(select_candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  select_candidates_kernel<<<n_sms * 2, 512, smem, s>>>(keys_dev, count_dev, capacity, cp, hist_dev);
  return cudaGetLastError();
}

cudaError_t launch_debug_draws(int mode, const KParams &p, int n_paths, const uint32_t *sobol_full,
                               const uint32_t *shift, uint32_t *out_dev, cudaStream_t s) {
  if (n_paths <= 0) return cudaSuccess;
  debug_draws_kernel<<<(n_paths + 127) / 128, 128, 0, s>>>(mode, p, n_paths, sobol_full, shift, out_dev);
  return cudaGetLastError();
}

cudaError_t launch_trajectories(int mode, const KParams &p, int n_paths, const float *tbl, const float *raw,
                                const uint32_t *sobol_full, const uint32_t *shift, float *values_dev, float *rets_dev,
                                cudaStream_t s) {
  if (n_paths <= 0) return cudaSuccess;
  trajectory_kernel<<<(n_paths + 63) / 64, 64, 0, s>>>(mode, p, n_paths, tbl, raw, sobol_full, shift, values_dev, rets_dev);
  return cudaGetLastError();
}

cudaError_t launch_debug_philox(const uint32_t *ctr_dev, PhiloxKeys keys, int rounds, int n, uint32_t *out_dev, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  debug_philox_kernel<<<(n + 127) / 128, 128, 0, s>>>(ctr_dev, keys, rounds, n, out_dev);
  return cudaGetLastError();
}

namespace {
bool reduce_uses_tma(int64_t n, int n_sms, int which) {
  // >= 4 chunks per CTA: the TMA-pipelined kernel (2 CTAs per SM); smaller inputs: plain streaming loads
  const bool big = n >= (int64_t)n_sms * 2 * 4 * kRedChunkFloats;
  return which == 2 || (which != 1 && big);
}
}  // namespace

const char *reduce_sum_kernel_name(int64_t n, int n_sms, int which) {
  return reduce_uses_tma(n, n_sms, which) ? "smmc::reduce_sum_tma_kernel" : "smmc::reduce_sum_kernel";
}

cudaError_t launch_reduce_sum(const float *v_dev, int64_t n, double *partials_dev, unsigned int *ticket_dev,
                              double *sum_dev, int n_sms, int which, cudaStream_t s) {
  if (reduce_uses_tma(n, n_sms, which)) {
    cudaError_t e = cudaFuncSetAttribute(reduce_sum_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRedSmemBytes);
    if (e != cudaSuccess) return e;
    reduce_sum_tma_kernel<<<n_sms * 2, kRedThreads, kRedSmemBytes, s>>>(v_dev, n, partials_dev, ticket_dev, sum_dev);
  } else {
    reduce_sum_kernel<<<n_sms * 4, 512, 0, s>>>(v_dev, n, partials_dev, ticket_dev, sum_dev);
  }
  return cudaGetLastError();
}

// dst[i] += src[i] over a statistics vector (a few hundred doubles): folds the first part of a split run into the second on
// the device so that the rank path all-reduces ONE vector
__global__ void add_f64_kernel(double *__restrict__ dst, const double *__restrict__ src, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] += src[i];
}
cudaError_t launch_add_f64(double *dst, const double *src, int n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  add_f64_kernel<<<(n + 255) / 256, 256, 0, s>>>(dst, src, n);
  return cudaGetLastError();
}

cudaError_t launch_vector_add(float *out, const float *a, const float *b, int64_t n, int n_sms, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int vec = ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15u) == 0;
  int64_t blocks = (n / (vec ? 4 : 1) + 511) / 512;
  if (blocks > (int64_t)n_sms * 4) blocks = (int64_t)n_sms * 4;
  if (blocks < 1) blocks = 1;
  vector_add_kernel<<<(int)blocks, 512, 0, s>>>(out, a, b, n, vec);
  return cudaGetLastError();
}

}  // namespace smmc
