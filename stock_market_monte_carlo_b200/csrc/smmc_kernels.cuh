// smmc_kernels.cuh -- kernel parameter block and launch interface between smmc_kernels.cu (device code) and
// smmc_api.cu (host side of the C ABI).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "smmc_device.cuh"

namespace smmc {

constexpr int kSelectMaxTargets = 8;  // = SMMC_SELECT_MAX_TARGETS
constexpr int kSelectBins = 2048;      // = SMMC_SELECT_BINS (11-bit digits: 11 + 11 + 10 bits = 3 passes)

struct KParams {
  int64_t n_paths;       // N of this launch
  uint64_t path0;        // global id of path 0 (multiple of 256)
  int32_t n_periods;     // P
  int32_t n_returns;     // n
  uint32_t compat_seed;  // added to the PCG seed input in BOOTSTRAP_COMPAT
  // return table, scaled by 0.01f and replicated `rep` times per entry ([n+1][rep], entry n is a zero pad);
  // Gaussian modes: the (cos, sin) table of sincos_tbl() instead ([kSinCosCells][2], tbl_bytes = kSinCosBytes; 0 = MUFU sin/cos)
  const float *tbl_rep;
  uint32_t tbl_bytes;  // bytes to stage (multiple of 16)
  uint32_t rep4;       // rep * 4: bytes per table row
  float idx_scale;     // compat: float(n) * 2^-32
  BootConsts bc;       // BOOTSTRAP: fma.rz index constants and the row shift
  uint32_t one;        // 0x3F800000, passed as a value the compiler cannot fold: keeps (w & mask) | one a single LOP3
  uint32_t cell_mask;  // kSinCosCellMask, opaque for the same reason (sincos_tbl<false>)
  PhiloxKeys keys;
  int32_t philox_rounds;  // 10 (default) or 7 (smmc_config.philox_rounds); the simulation kernel takes it as a template parameter
  GaussConsts gc;
  float sigma_sqrt2;  // SOBOL_GAUSSIAN: sqrt(2) * std/100
  float v0;
  float target;
  float hist_lo, hist_inv_w;
  int32_t hist_bins;
  // outputs (device pointers, nullable)
  float *finals;
  float *bmeans;
  float *bvars;
  // workspace
  double *cta_partials;       // [grid][4] = n, mean, M2, below
  unsigned long long *hist_g;  // [bins+2], zeroed before launch
  unsigned int *ticket;        // zeroed before launch
  // dynamic block assignment (launches of few waves): a warp's first block is static, every further one comes from
  // *block_ticket (zeroed before launch).  Which warp ran a block then depends on timing, so the blocks' (mean, M2) go to
  // block_moments[b] and the last CTA merges them in BLOCK order: every output bit stays a function of the launch plan.
  unsigned int *block_ticket;  // null = static assignment (warp w of CTA c walks b = c W + w, += grid W)
  float2 *block_moments;       // [nb]
  double *stats;               // [SMMC_STATS_HEADER + bins + 2]
  // Sobol (DESIGN.md "Sobol"): point index of path p is p + 256; month m is dimension m
  const uint32_t *sobol_lo;  // [P_pad][8]   direction numbers of index bits 0..7, month-major (staged to smem)
  const uint32_t *sobol_hi;  // [98][P_pad]  rows 0..1: digital shift per month (0 = plain), row 1 with the bit-7 direction
                             //              number folded in (h odd); rows 2 + 16 g + c: XOR of the direction numbers
                             //              of index bits 8 + 4 g + {set bits of c}, g < 6, c < 16
  int32_t sobol_ppad;        // P rounded up to a multiple of 32
  uint32_t sobol_lo_bytes;   // P_pad * 32
  // order-statistic search pass (smmc_select_pass): a radix-select digit histogram of the finals' ordered keys,
  // restricted per target to the key prefix found so far.  sel_targets == 0: not a search pass.
  int32_t sel_targets;          // histograms kept (1 in the first pass: no prefix yet)
  int32_t sel_digit_shift;      // digit = (key >> sel_digit_shift) & sel_digit_mask
  uint32_t sel_digit_mask;      // bins - 1
  int32_t sel_prefix_shift;     // prefix = key >> sel_prefix_shift; 32 = first pass (every key matches)
  uint32_t sel_prefix[kSelectMaxTargets];
  unsigned long long *sel_hist_g;  // [sel_targets][bins], zeroed before launch
  // candidate collection (second select pass): the ordered keys whose top 11 bits equal one of cand_prefix[] are appended
  // to cand_keys (one atomicAdd on *cand_cursor per 256-path block); the remaining 21 key bits are then resolved over
  // that buffer (select_candidates_kernel) instead of by a third replay.  Keys beyond cand_capacity are counted, not stored.
  int32_t cand_groups;                     // 0 = do not collect
  uint32_t cand_prefix[kSelectMaxTargets];  // distinct 11-bit prefixes
  uint32_t *cand_keys;
  unsigned long long *cand_cursor;  // zeroed before launch
  unsigned long long cand_capacity;
  int32_t gauss_tbl_mask;  // trajectory kernel: which Box-Muller pairs take sin/cos from the table (the mc kernel's TBL)
  // per-period withdrawals (null = none): value <- max(value * flow_q[m] - flow_w, 0) after month m's return
  const float *flow_q;  // [n_periods] keep-fractions 1 - pct_m/100 (device)
  float flow_w;         // fixed amount per period
};

// CTA shape of the simulation kernel and its dynamic shared-memory layout (shared by the planner and the kernel)
#ifndef SMMC_MAX_THREADS
#define SMMC_MAX_THREADS 1024
#endif
constexpr int kThreads = SMMC_MAX_THREADS;
constexpr int kWarps = kThreads / 32;

// shared memory layout (dynamic):
// [0,8) mbarrier | [16, 16+4*(bins+2)) histogram (u32 per CTA: a 64-bit shared-memory atomicAdd is a CAS loop, which the
// 32 lanes of a warp hitting a handful of bins turn into 2.7 % of the C2 kernel; counts per CTA stay < 2^32 because
// validate() caps a call at 2^38 paths) | warp partials | select-pass digit histograms (u32) |
// Sobol low direction numbers | table (128-B aligned)
struct SmemLayout {
  uint32_t hist_off, wp_off, sel_off, flow_off, sob_off, tbl_off, total;
};
// tbl_bytes: the return table (bootstrap modes) or the sin/cos table (Gaussian modes) -- a mode has one or the other
__host__ __device__ inline SmemLayout smem_layout(int bins, uint32_t sel_bytes, uint32_t flow_bytes, uint32_t sob_bytes,
                                                  uint32_t tbl_bytes) {
  SmemLayout L;
  L.hist_off = 16;
  L.wp_off = (L.hist_off + 4u * (uint32_t)(bins + 2) + 7u) & ~7u;
  L.sel_off = (L.wp_off + 32u * kWarps + 15u) & ~15u;
  L.flow_off = (L.sel_off + sel_bytes + 15u) & ~15u;
  L.sob_off = (L.flow_off + flow_bytes + 127u) & ~127u;
  L.tbl_off = (L.sob_off + sob_bytes + 127u) & ~127u;
  L.total = L.tbl_off + tbl_bytes;
  return L;
}

// Build-time choices that do not touch the contract (every choice computes the same draws):
constexpr int kGaussTblDefault = 5;      // GAUSSIAN: pairs 0 and 2 of a call take sin/cos from the shared-memory table
// ... per Philox round count: with 7 rounds the multiplier pipe has more room and three table pairs per call are best
// (100 M x 360 with the pre-scaled table: mask 5 14.78 ms, 9 14.64, 15 14.63, 7 14.56; with 10 rounds mask 5 17.11, 9 17.32,
// 7 17.46, 15 17.39; profiles/r02_variants_*.txt)
// SOBOL_GAUSSIAN: path slots (of a lane's eight) whose pairs take sin/cos from the table; slots 3 and 7 use MUFU.SIN/COS.
// With every pair from the table the kernel is issue-bound (84 % issue, XU 56 %): a MUFU pair is 3 instructions shorter,
// and a quarter of them balances the two (5 M x 1000: 0xFF 1.935 ms, 0x77 1.847, 0xBB 1.862, 0x6D 1.873, 0x55 1.887, 0x7F 1.950)
constexpr int kSobolTblDefault = 0x77;
constexpr int kSobolBootImmSlots = 0x55;  // SOBOL_BOOTSTRAP with 128-byte rows: path slots whose row stride is an immediate
constexpr int gauss_tbl_default(int rounds) { return rounds == 7 ? 7 : kGaussTblDefault; }
constexpr int kSobolUnrollDefault = 2;   // SOBOL_BOOTSTRAP month-loop unroll

// Test / development hook (smmc_debug_set_variant): selects among the compiled kernel variants.  Zero / negative = default.
// Never read from the environment: a drop-in library must not change kernels because of an environment variable.
struct Variant {
  int ilp = 0;           // paths per lane in flight (COMPAT / BOOTSTRAP: 2, 4, 8; GAUSSIAN: 4, 8)
  int warps = 0;         // warps per CTA (1..32), 0 = plan_warps
  int reduce = 0;        // reduce_mean kernel: 1 = plain streaming loads, 2 = TMA ring, 0 = by size
  int gauss_tbl = -1;    // GAUSSIAN: table mask (0 = all MUFU), -1 = default
  int sobol_tbl = -1;    // SOBOL_GAUSSIAN: 0 = MUFU sin/cos, 1 = table for every path slot, 2..255 = slot mask, -1 = default
  int sobol_unroll = 0;  // SOBOL_BOOTSTRAP month unroll (experiment builds), 0 = default
  int finals_chunks = 0; // host-pointer finals pipeline: number of chunks (0 = automatic, 1 = one monolithic launch)
  int finals_threads = 0; // ... and its host copy-out threads (0 = automatic)
  int select_candidates = -1;  // 0 = the third select pass always replays (no candidate buffer), -1 / 1 = default
  int ticket = -1;       // block assignment: 0 = static, 1 = ticket, -1 = by launch size (plan_ticket)
  int verbose = 0;       // 1 = smmc_comm_init_rank reports its communicator on stderr (one line per rank); silent otherwise
  int host_merge = 0;    // 1 = smmc_run_multi / smmc_quantiles merge their shards on the host as if NCCL could not be loaded
};

constexpr int kTicketMaxWaves = 48;  // launches of fewer waves of 256-path blocks take the ticket (the tail matters there)
constexpr int kSobolHiRows = 98;   // rows of KParams::sobol_hi
constexpr uint32_t kSobolSkip = 256;  // first Sobol point used; keeps 256-path blocks aligned to 256-point blocks

struct LaunchInfo {
  int grid;
  int threads;
  size_t smem_bytes;
  int rep;  // table replication factor chosen
  bool ticket;  // 256-path blocks handed to warps by an atomic ticket (KParams::block_ticket) instead of statically
};

// Chooses replication / shared-memory layout for `mode`; returns false if the table cannot fit.
// Also picks the warps per CTA for `n_paths` (warps_override in 1..32 forces it; 0 = automatic).
// sincos: Gaussian modes stage the sin/cos table (kSinCosBytes) where the bootstrap modes stage the return table.
bool plan_launch(int mode, int n_returns, int n_periods, int hist_bins, int sel_targets, bool withdrawals, bool sincos,
                 int64_t n_paths, int n_sms, size_t smem_optin, int warps_override, int ticket_override, LaunchInfo *li);

// Builds the scaled, replicated table: dst[(i*rep + c)] = src[i] * 0.01f, row n = 0, tail padding = 0.
void launch_prep_table(const float *returns_dev, int n, int rep, float *dst, uint32_t dst_floats, cudaStream_t s);
// dst = src sorted ascending (stable rank sort; SOBOL_BOOTSTRAP samples the empirical quantile function)
void launch_sort_table(const float *src, int n, float *dst, cudaStream_t s);

cudaError_t launch_mc(int mode, const KParams &p, const LaunchInfo &li, const Variant &v, cudaStream_t s);
// fills dst[kSinCosCells][2] with (cos, sin) of the cell centres, correctly rounded from double (once per device)
void launch_sincos_table(float *dst, double scale, cudaStream_t s);  // scale = sqrt(-GaussConsts::k)

// introspection kernels (parity tests)
// sobol_full: [n_dims][32] direction numbers (device), only read in the Sobol modes
cudaError_t launch_debug_draws(int mode, const KParams &p, int n_paths, const uint32_t *sobol_full,
                               const uint32_t *shift, uint32_t *out_dev, cudaStream_t s);
// trajectory mode: values_dev [n_paths][P+1], rets_dev [n_paths][P] (nullable); tbl/raw: [n+1] scaled / percent
cudaError_t launch_trajectories(int mode, const KParams &p, int n_paths, const float *tbl, const float *raw,
                                const uint32_t *sobol_full, const uint32_t *shift, float *values_dev, float *rets_dev,
                                cudaStream_t s);
cudaError_t launch_debug_philox(const uint32_t *ctr_dev, PhiloxKeys keys, int rounds, int n, uint32_t *out_dev, cudaStream_t s);

// digit histograms of the candidate keys collected by a select pass: hist[t][digit] over the keys whose top
// (32 - prefix_shift) bits equal prefix[t]; n_keys is read from *count_dev (clamped to capacity) on the device
cudaError_t launch_select_candidates(const uint32_t *keys_dev, const unsigned long long *count_dev, unsigned long long capacity,
                                     int targets, int digit_shift, uint32_t digit_mask, int prefix_shift,
                                     const uint32_t *prefix, unsigned long long *hist_dev, int n_sms, cudaStream_t s);

// small reference entry points
// which: 0 = by size, 1 = plain streaming loads, 2 = TMA ring (test hook)
cudaError_t launch_reduce_sum(const float *v_dev, int64_t n, double *partials_dev, unsigned int *ticket_dev,
                              double *sum_dev, int n_sms, int which, cudaStream_t s);
// name of the kernel launch_reduce_sum dispatches for (n, which) -- reported by the bench
const char *reduce_sum_kernel_name(int64_t n, int n_sms, int which);
cudaError_t launch_vector_add(float *out, const float *a, const float *b, int64_t n, int n_sms, cudaStream_t s);
cudaError_t launch_add_f64(double *dst, const double *src, int n, cudaStream_t s);  // dst[i] += src[i], statistics vectors

}  // namespace smmc
