// smmc_device.cuh -- device building blocks of the Monte Carlo portfolio path (sm_100a).
//
// Arithmetic contracts are stated in DESIGN.md and restated on the CPU in oracle/smmc_oracle.c; reference
// citations (/root/reference) are given per function.
#pragma once
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

namespace smmc {

constexpr int kBlock = 256;          // logical block = THREADS_PER_BLOCK of the reference (src/simulations.cu:17)
constexpr int kPathsPerLane = 8;     // one warp owns one logical block: lane l holds paths 32*j + l, j < 8

// ---- reference stream (BOOTSTRAP_COMPAT) ----------------------------------------------------------------

// PCG RXS-M-XS output hash used as per-path seed (reference src/simulations.cu:62-67; the LCG advance on its
// line 64 acts on a by-value copy and is dead code)
__device__ __forceinline__ uint32_t rand_pcg(uint32_t s) {
  uint32_t word = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
  return (word >> 22u) ^ word;
}

// reference src/simulations.cu:69-75
__device__ __forceinline__ uint32_t xorshift(uint32_t y) {
  y ^= y << 11;
  y ^= y >> 7;
  return y ^ (y >> 12);
}

// ---- Philox4x32-10 ------------------------------------------------------------------------------------------
// Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3" (SC'11).  Round keys are
// precomputed on the host (they depend on the seed only) and live in the kernel parameter constant bank.
struct PhiloxKeys {
  uint32_t k0[10];
  uint32_t k1[10];
};

constexpr uint32_t kPhiloxM0 = 0xD2511F53u;
constexpr uint32_t kPhiloxM1 = 0xCD9E8D57u;
constexpr uint32_t kPhiloxW0 = 0x9E3779B9u;
constexpr uint32_t kPhiloxW1 = 0xBB67AE85u;

__device__ __forceinline__ void philox_round(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3, uint32_t k0,
                                             uint32_t k1) {
  const uint64_t p0 = (uint64_t)kPhiloxM0 * c0;  // IMAD.WIDE.U32
  const uint64_t p1 = (uint64_t)kPhiloxM1 * c2;
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;  // LOP3 (3-input xor)
  const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
  c1 = (uint32_t)p1;
  c3 = (uint32_t)p0;
  c0 = n0;
  c2 = n2;
}

// ROUNDS = 10 is Philox4x32-10, the default of every native mode; ROUNDS = 7 is Philox4x32-7, the smallest round count the
// SC'11 paper reports as Crush-resistant (its table 2) and for which Random123 ships known-answer vectors -- an explicit
// opt-in of the stream contract (smmc_config.philox_rounds), never chosen silently.
template <int ROUNDS>
__device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const PhiloxKeys &k,
                                           uint32_t (&out)[4]) {
  static_assert(ROUNDS >= 1 && ROUNDS <= 10, "PhiloxKeys holds ten round keys");
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) philox_round(c0, c1, c2, c3, k.k0[r], k.k1[r]);
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              const PhiloxKeys &k, uint32_t (&out)[4]) {
  philox4x32<10>(c0, c1, c2, c3, k, out);
}

// Stream contract of the native modes: counter = (q, path_lo, tag, path_hi), key = seed.
// One Philox call = 128 bits serves SIX bootstrap months (call q -> months 6q..6q+5, one 21/22-bit field each)
// or EIGHT Gaussian months (call q -> months 8q..8q+7, one 32-bit word per Box-Muller pair).  The wide
// multiplies of Philox run at a quarter of the issue rate on sm_100a (profiles/r01_pipe_microbench.txt), so no
// bits are wasted.  The per-lane quantity (path_lo) sits in counter word 1 so that in rounds 1-3 one of the two
// 32x32->64 multiplies has warp-uniform operands (q, tag, path_hi and the key are uniform over a 256-path block):
// only 16 of the 20 multiplies of a call are per-lane work.
constexpr uint32_t kTagBootstrap = 0x424F4F54u;   // "BOOT"
constexpr uint32_t kTagGaussian = 0x47415553u;    // "GAUS"
constexpr uint32_t kTagSobolShift = 0x534F424Cu;  // "SOBL"

template <int ROUNDS = 10>
__device__ __forceinline__ void stream_words(uint32_t path_lo, uint32_t path_hi, uint32_t q, uint32_t tag,
                                             const PhiloxKeys &k, uint32_t (&w)[4]) {
  philox4x32<ROUNDS>(q, path_lo, tag, path_hi, k, w);
}
// run-time round count (trajectory / introspection kernels: one thread per path, not performance-critical)
__device__ __forceinline__ void stream_words_rt(int rounds, uint32_t path_lo, uint32_t path_hi, uint32_t q, uint32_t tag,
                                                const PhiloxKeys &k, uint32_t (&w)[4]) {
  if (rounds == 7) philox4x32<7>(q, path_lo, tag, path_hi, k, w);
  else philox4x32<10>(q, path_lo, tag, path_hi, k, w);
}

// ---- MUFU wrappers ------------------------------------------------------------------------------------------
__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sin_approx(float x) {
  float y;
  asm("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float cos_approx(float x) {
  float y;
  asm("cos.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- six bit fields of one Philox call ------------------------------------------------------------------------
//   f0..f3 = low 21 bits of w0..w3;  f4 = w0[31:21] : w1[31:21];  f5 = w2[31:21] : w3[31:21]  (22 bits)
// A field is used as the mantissa of a float in [1,2): m = 1 + f 2^-23 (one LOP3: (w & mask) | 0x3F800000).
constexpr int kFieldsPerCall = 6;
constexpr uint32_t kOne = 0x3F800000u;

__device__ __forceinline__ uint32_t field_of(const uint32_t (&w)[4], int j) {
  if (j < 4) return w[j] & 0x1FFFFFu;
  return j == 4 ? (((w[0] >> 21) << 11) | (w[1] >> 21)) : (((w[2] >> 21) << 11) | (w[3] >> 21));
}
// mantissa form of the six fields; `one` is 0x3F800000 held in a register so that (w & mask) | one is ONE LOP3
__device__ __forceinline__ void field_mantissas(const uint32_t (&w)[4], uint32_t one, uint32_t (&m)[6]) {
#pragma unroll
  for (int j = 0; j < 4; j++) m[j] = (w[j] & 0x1FFFFFu) | one;
  m[4] = (((w[0] >> 10) & 0x3FF800u) | (w[1] >> 21)) | one;
  m[5] = (((w[2] >> 10) & 0x3FF800u) | (w[3] >> 21)) | one;
}

// ---- bootstrap index (native modes) ---------------------------------------------------------------------------
// idx = floor(f * n / 2^bits), f a `bits`-wide field: unbiased to n / 2^21.  Computed without an integer multiply:
//   y = fma.rz(m, c, 2^23 - c),  m = 1 + f 2^-23,  c = n * 2^(23 - bits)   (c and 2^23 - c are exact: n < 2^21)
// m * c is exact inside the FMA and the sum lands in [2^23, 2^24), where the float spacing is 1, so round-toward-
// zero IS the floor: the bit pattern of y is 0x4B000000 + idx.  The byte address of the bank-private table row is
// then one shift-add (LEA): (bits(y) << log2(row bytes)) + base, with the shifted-out exponent folded into `base`.
struct BootConsts {
  float c21, k21;      // 4n, 2^23 - 4n
  float c22, k22;      // 2n, 2^23 - 2n
  uint32_t row_shift;  // log2(rep * 4)
};

template <int SHIFT>
__device__ __forceinline__ float boot_load(uint32_t m, float c, float k, uint32_t base, uint32_t row_shift) {
  const float y = __fmaf_rz(__uint_as_float(m), c, k);
  const uint32_t sh = SHIFT >= 0 ? (uint32_t)SHIFT : row_shift;
  const uint32_t addr = (__float_as_uint(y) << sh) + base;
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(addr));
  return r;
}

// ---- Gaussian growth factors (Box-Muller) -------------------------------------------------------------------
// r ~ N(mean, std) percent, fund *= (100 + r)/100 (reference src/simulations.cpp:14-16, :41-55), evaluated as
// g = (1 + mean/100) + (std/100) z with (z0, z1) = sqrt(-2 ln u) (cos v, sin v).
//
// One 32-bit word w serves one Box-Muller pair = two months:
//   radius  u = (f + 1/2) / 2^20  in (0,1),  f = w[19:0]        (never 0 or 1: no special cases)
//   angle   v = 2 pi * int32(w * K) / 2^32  in [-pi, pi],  K = 0x9E3779B1 (odd: w -> w K is a bijection mod 2^32)
// For a given radius field the 4096 words that share it map to a uniform grid of 4096 directions whose offset is
// frac(K f / 2^20): a hash of the radius bits, equidistributed over any radius interval.  Averages of a
// trigonometric polynomial of degree < 4096 over each grid equal the continuous average (all joint moments of
// (z0, z1) up to that order are those of independent normals), and the hashed offsets make the angle effectively
// continuous for non-smooth functionals (2-D chi-square and polar-bin tests in tests/test_oracle_pins.py; taking
// the angle from the raw word, or from its top 12 bits only, fails them).
//   (std/100) sqrt(-2 ln u) = sqrt(lg2(u) * k),  k = -2 ln2 (std/100)^2.
struct GaussConsts {
  float a;  // 1 + mean/100
  float k;  // -2 ln2 sigma^2
};

constexpr int kGaussMonthsPerCall = 8;
constexpr uint32_t kAngleHash = 0x9E3779B1u;

__device__ __forceinline__ void gauss_pair(uint32_t w, uint32_t one, const GaussConsts gc, float &g0, float &g1) {
  const float mr = __uint_as_float((w & 0xFFFFFu) | one);            // 1 + f 2^-23
  const float u = fmaf(mr, 8.0f, -7.99999952316284179688f);          // (f + 1/2) 2^-20, exact
  const float v = __int2float_rn((int)(w * kAngleHash)) * 1.4629180792671596e-09f;  // 2 pi 2^-32
  const float t = lg2_approx(u) * gc.k;
  const float rs = sqrt_approx(t);
  g0 = fmaf(rs, cos_approx(v), gc.a);
  g1 = fmaf(rs, sin_approx(v), gc.a);
}

// ---- sin / cos of the Box-Muller angle from a shared-memory table (no MUFU) -------------------------------------------
// The angle is 2 pi h / 2^32 for a 32-bit word h (mod 2 pi the same angle as the signed form above).  h = cell : low with a
// 13-bit cell index and a 19-bit remainder:  theta = theta_c + beta,  theta_c = 2 pi (cell + 1/2) / 2^13 (cell centre),
// beta = 2 pi (low - 2^18) / 2^32 in [-pi/2^13, pi/2^13).  The table holds (cos theta_c, sin theta_c) rounded from double;
// one first-order rotation  cos = c - s beta,  sin = s + c beta  leaves a truncation error beta^2/2 <= 7.4e-8 (2.5e-8 on
// average): the size of one float rounding near 1 and ~7 times smaller than the absolute error of MUFU.SIN/COS (2^-20.9),
// so the two evaluations of the same contract differ by less than either differs from libm (the oracle).  beta comes from
// the mantissa trick: m = 1 + low 2^-23 (one LOP3), beta = m kap + kap0 with kap = 2 pi 2^23 / 2^32, kap0 = -(1 + 2^-5) kap.
// Cost: SHF + LOP3 (address) + LDS.64 + LOP3 + 3 FFMA instead of I2FP + FMUL + FMUL.RZ + 2 MUFU: the XU pipe (4 lanes per
// clock per sub-partition, the binding pipe of the Gaussian modes) is relieved of half its work.  64 KB of shared memory.
constexpr int kSinCosCells = 8192;
constexpr uint32_t kSinCosBytes = kSinCosCells * 8u;   // one copy; the device buffer holds two: [swizzled][linear]
constexpr uint32_t kSinCosCellMask = 0xFFF80000u;       // the 13 cell bits of the angle word
constexpr float kSinCosKap = (float)(3.14159265358979323846 / 256.0);                       // 2 pi 2^23 / 2^32
constexpr float kSinCosKap0 = (float)(-(1.0 + 1.0 / 32.0) * 3.14159265358979323846 / 256.0);  // -(1 + 2^-5) kap

// Table layout: cell c lives at slot c ^ (c >> 9), i.e. the top four cell bits are XOR-ed into the bank-selecting bits.
// In SOBOL_GAUSSIAN the 32 lanes of a warp hold points that differ only in the TOP five bits of the word (the direction
// numbers of index bits 0..4), so a linear layout would put all 32 lanes in ONE bank (measured: 3.6x slower than MUFU);
// with the swizzle the 16 lanes of each LDS.64 half-warp phase hit 16 different 8-byte bank pairs: conflict-free.
// GAUSSIAN's hashed words are random either way.  Byte offset = ((h >> 16) ^ (h >> 25)) & 0xFFF8: two SHF + one LOP3.
__host__ __device__ __forceinline__ uint32_t sincos_slot(uint32_t cell) { return cell ^ (cell >> 9); }

// GAUSSIAN reads the LINEAR copy of the table (second half of the buffer, sincos_table_kernel): its hashed words are
// random, a swizzle buys nothing there, and the linear offset is cheaper -- hc = h & cell_mask (LOP3, the mask passed as an
// opaque value so that the shift is not commuted in front of it), address = base + (hc >> 16) (one LEA.HI), mantissa
// (h & ~cell_mask) | one (one LOP3): 2 address instructions instead of 3 (SASS checked: LOP3, LEA.HI, LOP3, LDS.64 [R+UR]).
template <bool SWZ>
__device__ __forceinline__ void sincos_tbl(uint32_t h, uint32_t one, uint32_t cell_mask, const char *sc_tbl, float &cv, float &sv) {
  float2 cs;
  uint32_t mant;
  if (SWZ) {
    cs = *reinterpret_cast<const float2 *>(sc_tbl + (((h >> 16) ^ (h >> 25)) & 0xFFF8u));
    mant = (h & 0x7FFFFu) | one;
  } else {
    const uint32_t hc = h & cell_mask;
    cs = *reinterpret_cast<const float2 *>(sc_tbl + (hc >> 16));
    mant = (h ^ hc) | one;
  }
  const float beta = fmaf(__uint_as_float(mant), kSinCosKap, kSinCosKap0);
  cv = fmaf(-cs.y, beta, cs.x);
  sv = fmaf(cs.x, beta, cs.y);
}

// The swizzled copy with the table given as a 32-bit shared-window address, the form the simulation kernel uses: slot << 3 =
// ((h ^ (h >> 9)) & cell_mask) >> 16, so the address is SHF, LOP3 ((h ^ s) & mask, the mask opaque), LEA.HI (base + (t >> 16)):
// the base add rides on the shift (the pointer form costs SHF, SHF, LOP3, IADD3; SOBOL_GAUSSIAN is issue-bound).
__device__ __forceinline__ void sincos_tbl_swz_u32(uint32_t h, uint32_t one, uint32_t cell_mask, uint32_t sc_base, float &cv,
                                                   float &sv) {
  float c, s;
  const uint32_t t = (h ^ (h >> 9)) & cell_mask;
  asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c), "=f"(s) : "r"(sc_base + (t >> 16)));
  const float beta = fmaf(__uint_as_float((h & 0x7FFFFu) | one), kSinCosKap, kSinCosKap0);
  cv = fmaf(-s, beta, c);
  sv = fmaf(c, beta, s);
}

__device__ __forceinline__ void gauss_pair_tbl(uint32_t w, uint32_t one, uint32_t cell_mask, const GaussConsts gc, const char *sc_tbl,
                                               float &g0, float &g1) {
  const float mr = __uint_as_float((w & 0xFFFFFu) | one);
  const float u = fmaf(mr, 8.0f, -7.99999952316284179688f);
  float cv, sv;
  sincos_tbl<false>(w * kAngleHash, one, cell_mask, sc_tbl, cv, sv);  // the table is pre-scaled by sqrt(-k)
  const float rs = sqrt_approx(-lg2_approx(u));
  g0 = fmaf(rs, cv, gc.a);
  g1 = fmaf(rs, sv, gc.a);
}

// ---- SOBOL_GAUSSIAN: Box-Muller on dimension pairs --------------------------------------------------------------
// Months 2k and 2k+1 (Sobol dimensions 2k, 2k+1) form one Box-Muller pair: the (digitally shifted) word of the even
// dimension is the radius uniform, the word of the odd dimension the angle:
//   u = (x_r + 1/2) / 2^32   (float: fma(float(x_r), 2^-32, 2^-33) in (0, 1]; u = 1 gives radius 0)
//   v = 2 pi int32(x_a) / 2^32   in [-pi, pi]
//   (g_2k, g_2k+1) = a + sqrt(lg2(u) k) (cos v, sin v),   a = 1 + mean/100,  k = -2 ln2 (std/100)^2.
// The transform is a smooth map of the unit square, so the pair keeps the low-discrepancy structure (Okten & Goncu,
// "Generating low-discrepancy sequences from the normal distribution: Box-Muller or inverse transform?", 2011); measured
// on the C4 integrand the randomised-QMC error of mean and second moment equals the inverse-CDF construction's
// (DESIGN.md section 3.4).  7 instructions per path-month instead of 17, 2 MUFU instead of 1.  An odd last month has no
// partner dimension and takes the inverse normal CDF below.
__device__ __forceinline__ void sobol_gauss_pair(uint32_t xr, uint32_t xa, const GaussConsts gc, float &g0, float &g1) {
  const float u = fmaf(__uint2float_rn(xr), 2.3283064365386963e-10f, 1.1641532182693481e-10f);  // 2^-32, 2^-33
  const float v = __int2float_rn((int32_t)xa) * 1.4629180792671596e-09f;                          // 2 pi 2^-32
  const float rs = sqrt_approx(lg2_approx(u) * gc.k);
  g0 = fmaf(rs, cos_approx(v), gc.a);
  g1 = fmaf(rs, sin_approx(v), gc.a);
}

// pointer form (trajectory kernel: the table is read from global memory there); same arithmetic, same bits
__device__ __forceinline__ void sobol_gauss_pair_tbl(uint32_t xr, uint32_t xa, uint32_t one, const GaussConsts gc,
                                                     const char *sc_tbl, float &g0, float &g1) {
  const float u = fmaf(__uint2float_rn(xr), 2.3283064365386963e-10f, 1.1641532182693481e-10f);  // 2^-32, 2^-33
  float cv, sv;
  sincos_tbl<true>(xa, one, 0u, sc_tbl, cv, sv);
  const float rs = sqrt_approx(-lg2_approx(u));
  g0 = fmaf(rs, cv, gc.a);
  g1 = fmaf(rs, sv, gc.a);
}

__device__ __forceinline__ void sobol_gauss_pair_tbl(uint32_t xr, uint32_t xa, uint32_t one, const GaussConsts gc,
                                                     uint32_t cell_mask, uint32_t sc_base, float &g0, float &g1) {
  const float u = fmaf(__uint2float_rn(xr), 2.3283064365386963e-10f, 1.1641532182693481e-10f);  // 2^-32, 2^-33
  float cv, sv;
  sincos_tbl_swz_u32(xa, one, cell_mask, sc_base, cv, sv);
  const float rs = sqrt_approx(-lg2_approx(u));
  g0 = fmaf(rs, cv, gc.a);
  g1 = fmaf(rs, sv, gc.a);
}

// ---- inverse normal CDF (SOBOL_GAUSSIAN, unpaired last month of an odd horizon) -----------------------------------
// Giles, "Approximating the erfinv function" (GPU Computing Gems, 2012), single-precision central/tail split.
// Returns p*y with z = sqrt(2) * p * y; the caller folds sqrt(2)*sigma into its FFMA.
//   y = 2u - 1,  u = (x + 1/2) / 2^32, scaled by (1 - 2^-23) so that |y| <= 1 - 2^-23 and 1 - y^2 >= 2^-22 > 0 for every
//       32-bit word (no clamp, no special case; the scaling moves z by < 1.2e-7 relative);
//   w - 2.5 = fma(lg2(1 - y^2), -ln 2, -2.5) feeds the central polynomial directly (one FFMA instead of FMUL + FADD).
// `xs` is the Sobol word with its top bit flipped (x ^ 0x80000000) read as a signed integer.
constexpr float kIcdfScale = 4.6566123179654e-10f;   // 2^-31 (1 - 2^-23)
constexpr float kIcdfOffset = 2.3283064365386963e-10f;  // 2^-32
__device__ __forceinline__ float erfinv_py_signed(int32_t xs) {
  const float y = fmaf(__int2float_rn(xs), kIcdfScale, kIcdfOffset);
  const float t = fmaf(-y, y, 1.0f);
  float w = fmaf(lg2_approx(t), -0.69314718055994530942f, -2.5f);
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
    w = sqrt_approx(w + 2.5f) - 3.0f;
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
  return p * y;
}
__device__ __forceinline__ float erfinv_py(uint32_t x32) { return erfinv_py_signed((int32_t)(x32 ^ 0x80000000u)); }

// ---- histogram bin rule (DESIGN.md "Histogram") ------------------------------------------------------------
__device__ __forceinline__ int hist_bin(float x, float lo, float inv_w, int bins) {
  const float t = (x - lo) * inv_w;
  if (!(t >= 0.0f)) return 0;
  if (t >= (float)bins) return bins + 1;
  return 1 + (int)t;
}

// ---- order-preserving key of a float (radix select over the finals) -------------------------------------------
// key(a) < key(b)  <=>  a < b for all non-NaN floats (-0 sorts just below +0); the inverse is key_to_float.
__host__ __device__ __forceinline__ uint32_t ordered_key(float x) {
#ifdef __CUDA_ARCH__
  const uint32_t b = __float_as_uint(x);
#else
  uint32_t b;
  memcpy(&b, &x, 4);
#endif
  return b ^ ((b & 0x80000000u) ? 0xFFFFFFFFu : 0x80000000u);
}

// ---- TMA bulk copy + mbarrier (table staging) ----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk async copy (UBLKCP), completion signalled on the mbarrier; bytes % 16 == 0, 16-B aligned
__device__ __forceinline__ void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

}  // namespace smmc
