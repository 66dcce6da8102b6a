/*
 * smmc_oracle.c -- CPU restatement of the Monte Carlo portfolio hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for the CUDA kernels in stock_market_monte_carlo_b200/csrc.  It is imported
 * only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; the product
 * (libsmmc.so) never links or calls it, and has no CPU fallback.
 *
 * What it restates (citations are /root/reference paths):
 *   - rand_pcg / xorshift / float index map / fmaf compounding of the reference GPU kernels
 *     (src/simulations.cu:62-75, :140-151, :214-222)                        -> orc_rand_pcg, orc_xorshift, compat path
 *   - the reduceBlock block mean / population variance with its exact reduction tree and its partial-block
 *     quirks (src/simulations.cu:157-183, :224-246)                          -> block_stats()
 *   - host statistics of the callers: mean/std (examples/benchmark_mc_gpu.cpp:7-28), strict `<` count below
 *     target (:30-41)                                                         -> global moments, below_count
 *   - Gaussian returns r ~ N(mean, std) in percent, fund*(100+r)/100 (src/simulations.cpp:14-16, :41-55) as
 *     the per-month growth factor (1 + mean/100) + (std/100) z
 *   - Sobol points: Gray-code XOR of direction numbers (include/sobol/sobol.cpp:13776-13811, :13883-13888)
 * What has no reference counterpart and is restated from its published algorithm:
 *   - Philox4x32-10 (Salmon et al., SC'11; Random123 1.x known-answer vectors, checked in tests/)
 *   - the histogram bin rule and the RNG stream contract, which are defined by this build.
 *
 * Pinning: orc_* is checked against (a) the reference's own sources compiled into oracle/_ref (CPU functions,
 * Burkardt i4_sobol, and -- on the GPU box -- the reference .cu kernels compiled for sm_100a), (b) Random123
 * known-answer vectors, (c) closed-form moments.  See tests/test_oracle_*.py.
 *
 * Build: oracle/Makefile  (gcc -O2 -ffp-contract=off -mfma -fopenmp -shared -fPIC).
 * -ffp-contract=off matters: every fused multiply-add below is an explicit fmaf() that mirrors one FFMA.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/smmc.h"

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------------------------- */
/* reference PRNG pieces (src/simulations.cu:62-75)                                                          */

uint32_t orc_rand_pcg(uint32_t rng_state) {
  /* the LCG advance on reference line 64 modifies a by-value copy and is dead code */
  uint32_t state = rng_state;
  uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
  return (word >> 22u) ^ word;
}

uint32_t orc_xorshift(uint32_t y) {
  y = y ^ (y << 11);
  y = y ^ (y >> 7);
  return y ^ (y >> 12);
}

/* reference index map, src/simulations.cu:148: int idx = n * (state * float(2.3283064e-10)) */
int32_t orc_compat_index(uint32_t state, int32_t n_returns) {
  float u = (float)state * 2.3283064e-10f; /* u32 -> f32 round-to-nearest, then exact scale by 2^-32 */
  float x = (float)n_returns * u;
  return (int32_t)x; /* truncation; equals n_returns when state >= 2^32 - 128 */
}

/* ------------------------------------------------------------------------------------------------------- */
/* Philox4x32-10                                                                                             */

#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

/* Philox4x32-R: R = 10 is the default of every native mode; R = 7 (smmc_config.philox_rounds) is the smallest round count
 * Salmon et al. report as Crush-resistant; Random123 publishes known-answer vectors for both (checked in tests/). */
void orc_philox4x32(const uint32_t ctr[4], const uint32_t key[2], int rounds, uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < rounds; r++) {
    uint64_t p0 = (uint64_t)PHILOX_M0 * c0;
    uint64_t p1 = (uint64_t)PHILOX_M1 * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += PHILOX_W0; k1 += PHILOX_W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { orc_philox4x32(ctr, key, 10, out); }

/* stream contract of the native modes:
 * counter = (q, path_lo, domain tag, path_hi), key = (seed_lo, seed_hi); call q serves months 6q..6q+5
 * (bootstrap: one 21/22-bit field per month) or 8q..8q+7 (gaussian: one 32-bit word per Box-Muller pair). */
#define TAG_BOOTSTRAP 0x424F4F54u /* "BOOT" */
#define TAG_GAUSSIAN 0x47415553u  /* "GAUS" */
#define TAG_SOBOL_SHIFT 0x534F424Cu /* "SOBL" */

static void stream_words_r(uint64_t path, uint32_t q, uint32_t tag, uint64_t seed, int rounds, uint32_t w[4]) {
  uint32_t ctr[4] = {q, (uint32_t)path, tag, (uint32_t)(path >> 32)};
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  orc_philox4x32(ctr, key, rounds, w);
}
static void stream_words(uint64_t path, uint32_t q, uint32_t tag, uint64_t seed, uint32_t w[4]) {
  stream_words_r(path, q, tag, seed, 10, w);
}
static int rounds_of(const smmc_config *cfg) { return cfg->philox_rounds == 7 ? 7 : 10; }

/* six bit fields of one call: f0..f3 = low 21 bits of w0..w3; f4 = w0[31:21]:w1[31:21]; f5 = w2[31:21]:w3[31:21] */
static uint32_t field_of(const uint32_t w[4], int j) {
  if (j < 4) return w[j] & 0x1FFFFFu;
  return j == 4 ? (((w[0] >> 21) << 11) | (w[1] >> 21)) : (((w[2] >> 21) << 11) | (w[3] >> 21));
}

/* bootstrap index of field j: floor(f n / 2^bits) in exact integer arithmetic.  (The device gets the same
 * integer from one round-toward-zero FMA; see boot_load in csrc/smmc_device.cuh.) */
static uint32_t boot_index(const uint32_t w[4], int j, uint32_t n) {
  return (uint32_t)(((uint64_t)field_of(w, j) * n) >> (j < 4 ? 21 : 22));
}

/* ------------------------------------------------------------------------------------------------------- */
/* Gaussian growth factors from one 32-bit word (Box-Muller pair)                                            */

typedef struct gauss_consts {
  float a;  /* 1 + mean/100                         */
  float k;  /* -2 ln2 sigma^2  (sigma = std/100)    */
} gauss_consts;

static gauss_consts make_gauss_consts(float mean_pct, float std_pct) {
  gauss_consts g;
  float mu = mean_pct * 0.01f;
  float sigma = std_pct * 0.01f;
  g.a = 1.0f + mu;
  g.k = (float)(-2.0 * 0.69314718055994530942 * (double)sigma * (double)sigma);
  return g;
}

static float as_float(uint32_t bits) {
  float f;
  memcpy(&f, &bits, sizeof f);
  return f;
}

/* word w -> growth factors of two consecutive months (DESIGN.md "Gaussian draws"):
 *   radius u = (f + 1/2)/2^20, f = w[19:0];  angle v = 2 pi int32(w K)/2^32, K = 0x9E3779B1 (bijective hash, so
 *   that the offset of the 4096-direction grid behind each radius is equidistributed).  The float constants are the device's,
 * digit for digit.  The device uses MUFU lg2/sqrt/sin/cos approximations where this uses libm, so Gaussian
 * finals agree to a stated tolerance, not bit-for-bit (tests state it). */
static void gauss_pair(uint32_t w, gauss_consts gc, float *g0, float *g1) {
  float mr = as_float((w & 0xFFFFFu) | 0x3F800000u);
  float u = fmaf(mr, 8.0f, -7.99999952316284179688f);
  float v = (float)(int32_t)(w * 0x9E3779B1u) * 1.4629180792671596e-09f;
  float t = log2f(u) * gc.k;
  float rs = sqrtf(t);
  float c = cosf(v), s = sinf(v);
  *g0 = fmaf(rs, c, gc.a);
  *g1 = fmaf(rs, s, gc.a);
}

/* test hook: growth factors of n words (g[2i], g[2i+1]) */
void orc_gauss_pairs(const uint32_t *w, int64_t n, float mean_pct, float std_pct, float *g) {
  gauss_consts gc = make_gauss_consts(mean_pct, std_pct);
  for (int64_t i = 0; i < n; i++) gauss_pair(w[i], gc, &g[2 * i], &g[2 * i + 1]);
}

/* ------------------------------------------------------------------------------------------------------- */
/* Sobol                                                                                                     */

/* direction numbers V[d][k], d < n_dims, k < 32, scaled so that bit 31 is the first binary digit
 * (V = v_burkardt << 2 for the 30-bit i4_sobol table).  Loaded by the caller from data/sobol_joe_kuo.bin. */
static const uint32_t *g_sobol_v = NULL;
static int32_t g_sobol_dims = 0;

void orc_sobol_set_table(const uint32_t *v, int32_t n_dims) {
  g_sobol_v = v;
  g_sobol_dims = n_dims;
}

/* x(point, dim) = XOR_{k : bit k of gray(point)} V[dim][k]   (include/sobol/sobol.cpp:13883-13888 unrolled) */
uint32_t orc_sobol_point(uint64_t point, int32_t dim) {
  uint64_t g = point ^ (point >> 1);
  uint32_t x = 0;
  for (int k = 0; k < 32 && (g >> k); k++)
    if ((g >> k) & 1u) x ^= g_sobol_v[(size_t)dim * 32 + k];
  return x;
}

/* Sobol point index of path p is p + SOBOL_SKIP: skips the all-zero point 0 and keeps every 256-path block aligned
 * to a 256-point block of the sequence (DESIGN.md "Sobol"). */
#define SOBOL_SKIP 256u

/* digital shift of dimension `dim` (0 when seed == 0, i.e. plain Sobol) */
static uint32_t sobol_shift(int32_t dim, uint64_t seed) {
  if (seed == 0) return 0;
  uint32_t w[4];
  stream_words((uint64_t)(dim >> 2), 0, TAG_SOBOL_SHIFT, seed, w);
  return w[dim & 3];
}

/* SOBOL_GAUSSIAN growth factors of months 2k, 2k+1 from the words of dimensions 2k (radius) and 2k+1 (angle)
 * (device: sobol_gauss_pair in csrc/smmc_device.cuh):
 *   u = (x_r + 1/2)/2^32 as fma(float(x_r), 2^-32, 2^-33);  v = 2 pi int32(x_a)/2^32;
 *   (g0, g1) = a + sqrt(log2(u) k) (cos v, sin v).  Constants are the device's, digit for digit; libm here, MUFU there. */
static void sobol_gauss_pair(uint32_t xr, uint32_t xa, gauss_consts gc, float *g0, float *g1) {
  float u = fmaf((float)xr, 2.3283064365386963e-10f, 1.1641532182693481e-10f);
  float v = (float)(int32_t)xa * 1.4629180792671596e-09f;
  float t = log2f(u) * gc.k;
  float rs = sqrtf(t);
  *g0 = fmaf(rs, cosf(v), gc.a);
  *g1 = fmaf(rs, sinf(v), gc.a);
}

/* test hook: growth factors of n (radius word, angle word) pairs (g[2i], g[2i+1]) */
void orc_sobol_gauss_pairs(const uint32_t *xr, const uint32_t *xa, int64_t n, float mean_pct, float std_pct, float *g) {
  gauss_consts gc = make_gauss_consts(mean_pct, std_pct);
  for (int64_t i = 0; i < n; i++) sobol_gauss_pair(xr[i], xa[i], gc, &g[2 * i], &g[2 * i + 1]);
}

/* inverse normal CDF used by SOBOL_GAUSSIAN for the unpaired last month of an odd horizon: single-precision restatement, see DESIGN.md "Sobol Gaussian".
 * Central/tail split of Giles' erfinv approximation ("Approximating the erfinv function", GPU Gems 2012),
 * evaluated with explicit fmaf so that host and device agree up to the log2/sqrt approximations. */
static float normal_icdf(uint32_t x32) { /* returns z / sqrt(2) */
  /* u in (0,1), symmetric about 0.5:  u = (x + 0.5) * 2^-32;  y = (2u - 1)(1 - 2^-23): |y| <= 1 - 2^-23, so
   * 1 - y^2 >= 2^-22 for every word (no clamp);  z = sqrt(2) * erfinv(y).  Constants are the device's, digit for digit. */
  float y = fmaf((float)(int32_t)(x32 ^ 0x80000000u), 4.6566123179654e-10f, 2.3283064365386963e-10f);
  float t = fmaf(-y, y, 1.0f);
  float w = fmaf(log2f(t), -0.69314718055994530942f, -2.5f); /* w - 2.5 of Giles' central branch, one FFMA */
  float p;
  if (w < 2.5f) {
    p = 2.81022636e-08f;
    p = fmaf(p, w, 3.43273939e-07f);
    p = fmaf(p, w, -3.5233877e-06f);
    p = fmaf(p, w, -4.39150654e-06f);
    p = fmaf(p, w, 0.00021858087f);
    p = fmaf(p, w, -0.00125372503f);
    p = fmaf(p, w, -0.00417768164f);
    p = fmaf(p, w, 0.246640727f);
    p = fmaf(p, w, 1.50140941f);
  } else {
    w = sqrtf(w + 2.5f) - 3.0f;
    p = -0.000200214257f;
    p = fmaf(p, w, 0.000100950558f);
    p = fmaf(p, w, 0.00134934322f);
    p = fmaf(p, w, -0.00367342844f);
    p = fmaf(p, w, 0.00573950773f);
    p = fmaf(p, w, -0.0076224613f);
    p = fmaf(p, w, 0.00943887047f);
    p = fmaf(p, w, 1.00167406f);
    p = fmaf(p, w, 2.83297682f);
  }
  return p * y; /* z = sqrt(2) * p * y; the caller folds sqrt(2) * sigma into one constant, as the device does */
}

/* ------------------------------------------------------------------------------------------------------- */
/* one path                                                                                                  */

typedef struct run_consts {
  const smmc_config *cfg;
  const float *tbl; /* returns * 0.01f, with one zero pad entry at [n] (compat one-past-the-end read) */
  const float *raw; /* the same entries in percent (trajectory mode's returns_pct); sorted like tbl */
  gauss_consts gc;
  float sigma_sqrt2; /* (float)(sqrt(2) * (double)(std * 0.01f)) */
  const uint32_t *shift; /* per-month Sobol digital shift */
  const float *flow_q;   /* per-month keep-fraction 1 - pct/100 of the withdrawal schedule, NULL = no withdrawals */
  float flow_w;          /* fixed amount per period */
} run_consts;

/* draws[m] (optional) receives the table index (bootstrap modes) or the raw word (gaussian modes) of month m */
/* withdrawals (include/smmc.h, ABI 2): after month m's return, value <- max(value * q_m - w, 0), one FFMA + one max,
 * as the device's Flow<true> computes it */
static float apply_flow(const run_consts *rc, float v, int m) {
  if (!rc->flow_q) return v;
  return fmaxf(fmaf(v, rc->flow_q[m], -rc->flow_w), 0.0f);
}

static float *build_flow(const smmc_config *cfg) {
  if (cfg->withdraw_fixed == 0.0f && cfg->withdraw_pct == 0.0f && !cfg->withdraw_pct_schedule) return NULL;
  int P = cfg->n_periods;
  float *q = (float *)malloc((size_t)(P > 0 ? P : 1) * sizeof(float));
  for (int m = 0; m < P; m++) {
    float pct = cfg->withdraw_pct_schedule ? cfg->withdraw_pct_schedule[m] : cfg->withdraw_pct;
    q[m] = 1.0f - pct * 0.01f;
  }
  return q;
}

/* values[0..P] (optional): V0 and the value after every month; rets[m] (optional): that month's return in percent */
static float simulate_path_full(const run_consts *rc, uint64_t path, uint32_t *draws, float *values, float *rets) {
  const smmc_config *cfg = rc->cfg;
  const int P = cfg->n_periods;
  const int n = cfg->n_returns;
  float total = cfg->initial_capital;
  if (values) values[0] = total;
  switch (cfg->mode) {
    case SMMC_MODE_BOOTSTRAP_COMPAT: {
      uint32_t state = orc_rand_pcg((uint32_t)path + 1u + (uint32_t)cfg->seed);
      for (int m = 0; m < P; m++) {
        state = orc_xorshift(state);
        int32_t idx = orc_compat_index(state, n);
        if (draws) draws[m] = (uint32_t)idx;
        total = apply_flow(rc, fmaf(total, rc->tbl[idx], total), m); /* total += total * r, one FFMA (reference SASS) */
        if (values) values[m + 1] = total;
        if (rets) rets[m] = idx < n ? rc->raw[idx] : 0.0f;
      }
    } break;
    case SMMC_MODE_BOOTSTRAP: {
      uint32_t w[4];
      for (int m = 0; m < P; m++) {
        int j = m % 6;
        if (j == 0) stream_words_r(path, (uint32_t)(m / 6), TAG_BOOTSTRAP, cfg->seed, rounds_of(cfg), w);
        uint32_t idx = boot_index(w, j, (uint32_t)n);
        if (draws) draws[m] = idx;
        total = apply_flow(rc, fmaf(total, rc->tbl[idx], total), m);
        if (values) values[m + 1] = total;
        if (rets) rets[m] = rc->raw[idx];
      }
    } break;
    case SMMC_MODE_GAUSSIAN: {
      uint32_t w[4];
      float g[2];
      for (int m = 0; m < P; m++) {
        int j = m % 8;
        if (j == 0) stream_words_r(path, (uint32_t)(m / 8), TAG_GAUSSIAN, cfg->seed, rounds_of(cfg), w);
        if ((j & 1) == 0) gauss_pair(w[j >> 1], rc->gc, &g[0], &g[1]);
        if (draws) draws[m] = (j & 1) ? w[j >> 1] * 0x9E3779B1u : (w[j >> 1] & 0xFFFFFu); /* radius field | hashed angle word */
        total = apply_flow(rc, total * g[j & 1], m);
        if (values) values[m + 1] = total;
        if (rets) rets[m] = (g[j & 1] - 1.0f) * 100.0f;
      }
    } break;
    case SMMC_MODE_SOBOL_BOOTSTRAP: {
      for (int m = 0; m < P; m++) {
        uint32_t x = orc_sobol_point(path + SOBOL_SKIP, m) ^ rc->shift[m];
        uint32_t idx = (uint32_t)(((uint64_t)x * (uint32_t)n) >> 32);
        if (draws) draws[m] = idx;
        total = apply_flow(rc, fmaf(total, rc->tbl[idx], total), m);
        if (values) values[m + 1] = total;
        if (rets) rets[m] = rc->raw[idx];
      }
    } break;
    case SMMC_MODE_SOBOL_GAUSSIAN: {
      /* Box-Muller on dimension pairs: months 2k (radius word) and 2k+1 (angle word); an unpaired last month of an
       * odd horizon takes the inverse normal CDF of its word */
      float g[2] = {1.0f, 1.0f};
      for (int m = 0; m < P; m++) {
        uint32_t x = orc_sobol_point(path + SOBOL_SKIP, m) ^ rc->shift[m];
        if (draws) draws[m] = x;
        if ((m & 1) == 0) {
          if (m + 1 < P) {
            uint32_t xa = orc_sobol_point(path + SOBOL_SKIP, m + 1) ^ rc->shift[m + 1];
            sobol_gauss_pair(x, xa, rc->gc, &g[0], &g[1]);
          } else {
            g[0] = fmaf(normal_icdf(x), rc->sigma_sqrt2, rc->gc.a);
          }
        }
        float gm = g[m & 1];
        total = apply_flow(rc, total * gm, m);
        if (values) values[m + 1] = total;
        if (rets) rets[m] = (gm - 1.0f) * 100.0f;
      }
    } break;
    default: break;
  }
  return total;
}

static float simulate_path(const run_consts *rc, uint64_t path, uint32_t *draws) {
  return simulate_path_full(rc, path, draws, NULL, NULL);
}

/* ------------------------------------------------------------------------------------------------------- */
/* block statistics: the reference's reduction tree (src/simulations.cu:157-183) on 256 slots                */

static float tree_sum_256(float *s) {
  for (int t = 0; t < 128; t++) s[t] += s[t + 128];
  for (int t = 0; t < 64; t++) s[t] += s[t + 64];
  /* warpReduce: lanes 0..31, strides 32,16,8,4,2,1; every lane executes, lane t reads slot t+stride which at
   * stride 32 may be a slot >= 64 holding an already-consumed value -- only slot 0's dependency cone matters,
   * and that cone only ever reads slots whose values are "live" partial sums. */
  for (int stride = 32; stride >= 1; stride >>= 1) {
    float tmp[32];
    for (int t = 0; t < 32; t++) tmp[t] = s[t] + s[t + stride];
    /* lanes run in lockstep: all reads of a step happen before its writes */
    for (int t = 0; t < 32; t++) s[t] = tmp[t];
  }
  return s[0];
}

/* x[0..live) live finals of one logical block.  compat != 0 reproduces the reference's partial-block behaviour:
 * dead slots contribute 0 to both sums and the divisor stays 256 (src/simulations.cu:205-207, :232, :241-242). */
static void block_stats(const float *x, int live, int compat, float *mean_out, float *var_out, double *m2_live,
                        double *mean_live) {
  float s[SMMC_BLOCK];
  for (int t = 0; t < SMMC_BLOCK; t++) s[t] = t < live ? x[t] : 0.0f;
  float sum = tree_sum_256(s);
  float div = (compat || live == SMMC_BLOCK) ? 256.0f : (float)live;
  float mean = sum / div;
  for (int t = 0; t < SMMC_BLOCK; t++) {
    if (t < live) {
      float d = x[t] - mean;
      s[t] = d * d;
    } else {
      s[t] = 0.0f;
    }
  }
  float sq = tree_sum_256(s);
  *mean_out = mean;
  *var_out = sq / div;
  if (live == SMMC_BLOCK || !compat) {
    *mean_live = (double)mean;
    *m2_live = (double)sq;
  } else {
    /* partial compat block: the global moments use the true mean of the live paths */
    float tmean = sum / (float)live;
    for (int t = 0; t < SMMC_BLOCK; t++) {
      if (t < live) {
        float d = x[t] - tmean;
        s[t] = d * d;
      } else {
        s[t] = 0.0f;
      }
    }
    *mean_live = (double)tmean;
    *m2_live = (double)tree_sum_256(s);
  }
}

/* histogram bin rule (defined by this build, DESIGN.md "Histogram"): bins+2 counters */
int32_t orc_hist_bin(float x, float lo, float inv_w, int32_t bins) {
  float t = (x - lo) * inv_w;
  if (!(t >= 0.0f)) return 0;
  if (t >= (float)bins) return bins + 1;
  return 1 + (int32_t)t;
}

float orc_hist_inv_width(float lo, float hi, int32_t bins) { return (float)((double)bins / ((double)hi - (double)lo)); }

/* ------------------------------------------------------------------------------------------------------- */

static int cmp_float(const void *a, const void *b) {
  float x = *(const float *)a, y = *(const float *)b;
  return (x > y) - (x < y);
}

/* tbl[i] = returns[i] * 0.01f (src/simulations.cu:126) plus one zero pad entry at [n] (compat one-past-the-end
 * read).  SOBOL_BOOTSTRAP samples the SORTED table: idx = floor(u n) is then the empirical quantile function,
 * which is what makes a low-discrepancy u meaningful; the resampling distribution is unchanged. */
static float *build_table(const smmc_config *cfg) {
  const int n = cfg->n_returns;
  if (n <= 0 || !cfg->returns) return NULL;
  float *tbl = (float *)malloc(((size_t)n + 1) * sizeof(float));
  for (int i = 0; i < n; i++) tbl[i] = cfg->returns[i] * 0.01f;
  tbl[n] = 0.0f;
  if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP) qsort(tbl, (size_t)n, sizeof(float), cmp_float);
  return tbl;
}

static int validate(const smmc_config *cfg) {
  if (!cfg || cfg->n_simulations < 0 || cfg->n_periods < 0) return 0;
  int boot = cfg->mode == SMMC_MODE_BOOTSTRAP_COMPAT || cfg->mode == SMMC_MODE_BOOTSTRAP ||
             cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP;
  if (boot && (cfg->n_returns <= 0 || !cfg->returns)) return 0;
  if ((cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP || cfg->mode == SMMC_MODE_SOBOL_GAUSSIAN) &&
      (!g_sobol_v || cfg->n_periods > g_sobol_dims))
    return 0;
  if (cfg->hist_bins < 0 || cfg->hist_bins > SMMC_MAX_HIST_BINS) return 0;
  return 1;
}

/* Same contract as smmc_run (include/smmc.h) with host buffers.  Returns 0 on success. */
int orc_run(const smmc_config *cfg, smmc_result *out) {
  if (!validate(cfg) || !out) return 1;
  const int64_t N = cfg->n_simulations;
  const int P = cfg->n_periods;
  const int compat = cfg->mode == SMMC_MODE_BOOTSTRAP_COMPAT;
  const int bins = cfg->hist_bins;
  const int64_t nb = (N + SMMC_BLOCK - 1) / SMMC_BLOCK;

  run_consts rc;
  memset(&rc, 0, sizeof rc);
  rc.cfg = cfg;
  float *tbl = build_table(cfg);
  rc.tbl = tbl;
  rc.gc = make_gauss_consts(cfg->gauss_mean, cfg->gauss_std);
  rc.sigma_sqrt2 = (float)(1.41421356237309504880 * (double)(cfg->gauss_std * 0.01f));
  uint32_t *shift = (uint32_t *)calloc((size_t)(P > 0 ? P : 1), sizeof(uint32_t));
  if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP || cfg->mode == SMMC_MODE_SOBOL_GAUSSIAN)
    for (int m = 0; m < P; m++) shift[m] = sobol_shift(m, cfg->seed);
  rc.shift = shift;
  float *flowq = build_flow(cfg);
  rc.flow_q = flowq;
  rc.flow_w = cfg->withdraw_fixed;

  const float inv_w = bins > 0 ? orc_hist_inv_width(cfg->hist_lo, cfg->hist_hi, bins) : 0.0f;
  uint64_t *hist = (uint64_t *)calloc((size_t)bins + 2, sizeof(uint64_t));
  uint64_t below = 0;
  /* per-block (n, mean, M2) kept so the fp64 merge runs in block order regardless of threading */
  double *bmean = (double *)malloc((size_t)(nb > 0 ? nb : 1) * sizeof(double));
  double *bm2 = (double *)malloc((size_t)(nb > 0 ? nb : 1) * sizeof(double));

#pragma omp parallel
  {
    uint64_t *lhist = (uint64_t *)calloc((size_t)bins + 2, sizeof(uint64_t));
    uint64_t lbelow = 0;
#pragma omp for schedule(dynamic, 16)
    for (int64_t b = 0; b < nb; b++) {
      float x[SMMC_BLOCK];
      int64_t first = b * SMMC_BLOCK;
      int live = (int)((N - first) < SMMC_BLOCK ? (N - first) : SMMC_BLOCK);
      for (int t = 0; t < live; t++) {
        x[t] = simulate_path(&rc, cfg->stream_offset + (uint64_t)(first + t), NULL);
        if (out->finals) out->finals[first + t] = x[t];
        if (x[t] < cfg->target) lbelow++;
        if (bins > 0) lhist[orc_hist_bin(x[t], cfg->hist_lo, inv_w, bins)]++;
      }
      float mean, var;
      block_stats(x, live, compat, &mean, &var, &bm2[b], &bmean[b]);
      if (out->block_means) out->block_means[b] = mean;
      if (out->block_vars) out->block_vars[b] = var;
    }
#pragma omp critical
    {
      below += lbelow;
      for (int i = 0; i < bins + 2; i++) hist[i] += lhist[i];
    }
    free(lhist);
  }

  /* Chan et al. pairwise merge of (n, mean, M2) in fp64, block order */
  double cn = 0.0, cmean = 0.0, cm2 = 0.0;
  for (int64_t b = 0; b < nb; b++) {
    int64_t first = b * SMMC_BLOCK;
    double nbk = (double)((N - first) < SMMC_BLOCK ? (N - first) : SMMC_BLOCK);
    double tot = cn + nbk;
    double delta = bmean[b] - cmean;
    cmean += delta * (nbk / tot);
    cm2 += bm2[b] + delta * delta * (cn * nbk / tot);
    cn = tot;
  }
  out->count = (uint64_t)N;
  out->below_count = below;
  out->mean = cmean;
  out->m2 = cm2;
  out->device_ms = 0.0f;
  out->gpu_launches = 0;
  if (out->hist && bins > 0) memcpy(out->hist, hist, ((size_t)bins + 2) * sizeof(uint64_t));
  free(hist);
  free(bmean);
  free(bm2);
  free(shift);
  free(flowq);
  free(tbl);
  return 0;
}

/* Same contract as smmc_debug_draws. */
int orc_draws(const smmc_config *cfg, int32_t n_paths, uint32_t *outbuf) {
  if (!validate(cfg) || !outbuf) return 1;
  run_consts rc;
  memset(&rc, 0, sizeof rc);
  rc.cfg = cfg;
  const int P = cfg->n_periods;
  float *tbl = build_table(cfg);
  rc.tbl = tbl;
  rc.gc = make_gauss_consts(cfg->gauss_mean, cfg->gauss_std);
  rc.sigma_sqrt2 = (float)(1.41421356237309504880 * (double)(cfg->gauss_std * 0.01f));
  uint32_t *shift = (uint32_t *)calloc((size_t)(P > 0 ? P : 1), sizeof(uint32_t));
  if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP || cfg->mode == SMMC_MODE_SOBOL_GAUSSIAN)
    for (int m = 0; m < P; m++) shift[m] = sobol_shift(m, cfg->seed);
  rc.shift = shift;
  float *flowq = build_flow(cfg);
  rc.flow_q = flowq;
  rc.flow_w = cfg->withdraw_fixed;
  for (int32_t p = 0; p < n_paths; p++) simulate_path(&rc, cfg->stream_offset + (uint64_t)p, outbuf + (size_t)p * P);
  free(shift);
  free(flowq);
  free(tbl);
  return 0;
}

/* Same contract as smmc_trajectories: values [n_paths][P+1], returns_pct [n_paths][P] (nullable).  Reference:
 * mc_simulations_keepdata stores many_updates()'s P+1 values per path (src/simulations.cpp:139-202, :24-39). */
int orc_trajectories(const smmc_config *cfg, int32_t n_paths, float *values, float *returns_pct) {
  if (!validate(cfg) || (n_paths > 0 && !values)) return 1;
  run_consts rc;
  memset(&rc, 0, sizeof rc);
  rc.cfg = cfg;
  const int P = cfg->n_periods;
  float *tbl = build_table(cfg);
  rc.tbl = tbl;
  float *raw = NULL;
  if (tbl) { /* percent values in the same (possibly sorted) order as tbl */
    raw = (float *)malloc(((size_t)cfg->n_returns + 1) * sizeof(float));
    memcpy(raw, cfg->returns, (size_t)cfg->n_returns * sizeof(float));
    raw[cfg->n_returns] = 0.0f;
    if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP) qsort(raw, (size_t)cfg->n_returns, sizeof(float), cmp_float);
  }
  rc.raw = raw;
  rc.gc = make_gauss_consts(cfg->gauss_mean, cfg->gauss_std);
  rc.sigma_sqrt2 = (float)(1.41421356237309504880 * (double)(cfg->gauss_std * 0.01f));
  uint32_t *shift = (uint32_t *)calloc((size_t)(P > 0 ? P : 1), sizeof(uint32_t));
  if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP || cfg->mode == SMMC_MODE_SOBOL_GAUSSIAN)
    for (int m = 0; m < P; m++) shift[m] = sobol_shift(m, cfg->seed);
  rc.shift = shift;
  float *flowq = build_flow(cfg);
  rc.flow_q = flowq;
  rc.flow_w = cfg->withdraw_fixed;
  for (int32_t p = 0; p < n_paths; p++)
    simulate_path_full(&rc, cfg->stream_offset + (uint64_t)p, NULL, values + (size_t)p * (P + 1),
                       returns_pct ? returns_pct + (size_t)p * P : NULL);
  free(shift);
  free(flowq);
  free(raw);
  free(tbl);
  return 0;
}

/* Order statistics the way the reference takes them: fully sort the finals, read the value at each rank
 * (update_quartiles, examples/visualize_returns_gpu.cpp:83-110: std::sort, then v[0], v[N/4], v[N/2], v[N/4+N/2], v[N-1]). */
int orc_quantiles(const float *finals, int64_t n, int32_t n_targets, const uint64_t *ranks, float *out) {
  if (!finals || n <= 0 || !ranks || !out) return 1;
  float *v = (float *)malloc((size_t)n * sizeof(float));
  memcpy(v, finals, (size_t)n * sizeof(float));
  qsort(v, (size_t)n, sizeof(float), cmp_float);
  int rc = 0;
  for (int t = 0; t < n_targets; t++) {
    if (ranks[t] >= (uint64_t)n) { rc = 1; break; }
    out[t] = v[ranks[t]];
  }
  free(v);
  return rc;
}

/* ranks of {min, Q1, median, Q3, max} as the reference indexes them (examples/visualize_returns_gpu.cpp:84-86, :109) */
void orc_quartile_ranks(int64_t n, uint64_t ranks[5]) {
  int64_t q1 = n / 4, q2 = n / 2, q3 = q1 + q2;
  ranks[0] = 0;
  ranks[1] = (uint64_t)q1;
  ranks[2] = (uint64_t)q2;
  ranks[3] = (uint64_t)q3;
  ranks[4] = (uint64_t)(n - 1);
}

/* host restatements of the two small reference entry points, for their parity tests */
float orc_reduce_mean(const float *v, int64_t n) { /* reference returns sum/n of fp32 block sums; we compare with tolerance */
  double s = 0.0;
  for (int64_t i = 0; i < n; i++) s += v[i];
  return (float)(s / (double)n);
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
