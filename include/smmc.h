/*
 * smmc.h -- C ABI of the B200-native Monte Carlo portfolio hot path.
 *
 * This is the drop-in boundary underneath the reference's C++ entry points
 * (reference: include/stock_market_monte_carlo/simulations.h:69-86, gpu.h:1-2;
 * definitions src/simulations.cu:269-341, :661-697, src/gpu.cu:17-47).  The C++
 * symbols of the reference (`mc_simulations_gpu`, `mc_simulations_gpu_reduceBlock`,
 * `reduce_mean_gpu`, `vector_add`, `vector_add_gpu`) are exported by the same
 * shared library as thin marshalling wrappers over the functions declared here
 * (see include/stock_market_monte_carlo/simulations.h in this repo).
 *
 * Plain pointers and sizes only: no C++ types, no torch types.  All functions
 * return an smmc_status; none of them calls exit() or throws.
 *
 * There is NO CPU fallback: every compute entry point needs a CUDA device of
 * compute capability 10.x and fails with SMMC_ERR_NO_DEVICE / SMMC_ERR_CUDA otherwise.
 */
#ifndef SMMC_H_
#define SMMC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMMC_ABI_VERSION 3
#define SMMC_BLOCK 256          /* logical block = THREADS_PER_BLOCK of the reference (src/simulations.cu:17) */
#define SMMC_MAX_HIST_BINS 4096

typedef enum smmc_status {
  SMMC_OK = 0,
  SMMC_ERR_INVALID_ARG = 1,   /* bad config (message via smmc_last_error) */
  SMMC_ERR_NO_DEVICE = 2,     /* no CUDA device / wrong architecture */
  SMMC_ERR_CUDA = 3,          /* a CUDA runtime call failed */
  SMMC_ERR_UNSUPPORTED = 4    /* valid request this build cannot serve (e.g. table too large for shared memory) */
} smmc_status;

/* How each monthly return is produced. */
typedef enum smmc_mode {
  /* Reference stream, bit-for-bit: seed rand_pcg(path+1) (src/simulations.cu:62-67,:140), xorshift per month
   * (:69-75,:146), float index map (:148), fmaf compounding (:149).  32-bit path ids: stream_offset+N < 2^31. */
  SMMC_MODE_BOOTSTRAP_COMPAT = 0,
  /* Philox4x32-10 counter stream, six 21/22-bit fields per call, index = floor(field * n_returns / 2^bits)
   * (unbiased to n/2^21; n_returns < 2^21). */
  SMMC_MODE_BOOTSTRAP = 1,
  /* Philox4x32-10 counter stream + Box-Muller (one 32-bit word per pair); r ~ N(gauss_mean, gauss_std) in percent
   * (reference CPU semantics: sample_returns_gaussian, src/simulations.cpp:41-55). */
  SMMC_MODE_GAUSSIAN = 2,
  /* Sobol (Joe-Kuo direction numbers as vendored in reference include/sobol/sobol.cpp), point = path id + 256,
   * dimension = month; optional Philox-keyed digital shift when seed != 0.
   * SOBOL_BOOTSTRAP: index = mulhi(word, n) into the SORTED table (empirical quantile function).
   * SOBOL_GAUSSIAN: Box-Muller on dimension pairs -- month 2k takes its radius from the word of dimension 2k, month
   * 2k+1 its angle from dimension 2k+1; the unpaired last month of an odd horizon takes the inverse normal CDF.
   * LIMITS (validated, SMMC_ERR_INVALID_ARG): n_periods <= 1111 (the dimensions of the vendored table) and 32-bit point
   * indices, stream_offset + n_simulations + 256 <= 2^32.  The reference vendors i8_sobol (include/sobol/sobol.cpp:14191,
   * 62-bit indices) but never calls it; its i4_sobol stops at 2^30 points (maxcol = 30), the 32 columns used here follow
   * the same recurrence.  Runs beyond 2^32 points per seed are split over seeds (each a different digital shift). */
  SMMC_MODE_SOBOL_BOOTSTRAP = 3,
  SMMC_MODE_SOBOL_GAUSSIAN = 4
} smmc_mode;

/* smmc_config.flags */
#define SMMC_FLAG_DEVICE_POINTERS 1u   /* `returns` and every smmc_result buffer are DEVICE pointers on cfg->device;
                                          the call is asynchronous on `stream` except for the scalar statistics,
                                          which are only filled when SMMC_FLAG_NO_SYNC is absent */
#define SMMC_FLAG_NO_SYNC 2u           /* with DEVICE_POINTERS: do not synchronise; scalars in smmc_result are left
                                          untouched and the caller reads `stats_dev` (see below) itself */

typedef struct smmc_config {
  uint32_t abi_version;        /* SMMC_ABI_VERSION */
  int32_t mode;                /* smmc_mode */
  int64_t n_simulations;       /* N: paths simulated by this call (reference: max_n_simulations) */
  int32_t n_periods;           /* P: compounding periods (months) per path */
  int32_t n_returns;           /* n: length of `returns` (bootstrap modes), else 0 */
  const float *returns;        /* historical monthly returns in PERCENT (reference units), bootstrap modes */
  uint64_t seed;               /* Philox key (native modes); added to the PCG seed input in COMPAT (0 = reference) */
  uint64_t stream_offset;      /* global id of path 0 of this call: path i uses RNG stream stream_offset + i.
                                  Multi-GPU sharding = disjoint [offset, offset+N) ranges, multiples of 256 */
  float initial_capital;       /* V0 */
  float gauss_mean;            /* percent per period (GAUSSIAN modes) */
  float gauss_std;             /* percent per period */
  float target;                /* below_count = #{ final < target }  (strict, fp32 compare:
                                  reference examples/benchmark_mc_gpu.cpp:30-41) */
  float hist_lo, hist_hi;      /* histogram range [lo, hi); bins of equal width */
  int32_t hist_bins;           /* 0 = no histogram; else 1..SMMC_MAX_HIST_BINS; output has hist_bins+2 counters:
                                  [0] = underflow (x < lo or NaN), [1..bins] = bins, [bins+1] = overflow (x >= hi) */
  int32_t device;              /* CUDA device ordinal */
  uint32_t flags;
  void *stream;                /* cudaStream_t (DEVICE_POINTERS only; NULL = legacy default stream) */
  /* ---- ABI 2: per-period withdrawals (the reference README's open "withdrawal strategies" item, README.md:130-133).
   * After the return of period m:  value <- max(value * (1 - pct_m/100) - withdraw_fixed, 0);  0 is absorbing.
   * All zero / NULL = no withdrawals: the default kernels, results bit-identical to ABI 1. */
  float withdraw_fixed;        /* fixed amount taken out every period (unit of initial_capital) */
  float withdraw_pct;          /* percent of the value taken out every period */
  const float *withdraw_pct_schedule; /* optional [n_periods] percent per period (overrides withdraw_pct);
                                         always a HOST pointer, copied by the call */
  /* ---- ABI 3: round count of the counter-based generator of BOOTSTRAP / GAUSSIAN (part of the stream contract).
   * 0 or 10 = Philox4x32-10 (default; the Random123 recommendation).  7 = Philox4x32-7, the smallest round count that
   * Salmon et al. (SC'11, table 2) report as Crush-resistant and for which Random123 publishes known-answer vectors:
   * an explicit opt-in that trades the safety margin for ~30 % less generator work.  Other values are rejected.
   * Ignored by BOOTSTRAP_COMPAT (reference stream) and the Sobol modes (their digital shift always uses 10 rounds). */
  int32_t philox_rounds;
  int32_t reserved0;           /* must be 0 */
} smmc_config;

/* Device-side reduced statistics, laid out as doubles so that ONE sum-allreduce (ncclFloat64) over
 * ranks merges every field exactly: the integer counters are < 2^53 and therefore exact in double.
 *   [0] count   [1] below_count   [2] S1 = sum (x - V0)   [3] S2 = sum (x - V0)^2
 *   [4] reserved (0)   [5] reserved (0)   [6 .. 6+hist_bins+2) histogram counters
 * smmc_stats_len(bins) gives the length; smmc_stats_finalize() turns a (possibly all-reduced) vector into
 * count / mean / M2 / below / hist. */
#define SMMC_STATS_HEADER 6

typedef struct smmc_result {
  /* caller-owned output buffers (host pointers unless SMMC_FLAG_DEVICE_POINTERS); NULL = not requested */
  float *finals;               /* [N]  final value of path i (reference `totals`, src/simulations.cu:151) */
  float *block_means;          /* [ceil(N/256)] per-256-path-block mean (reference `totals[blockIdx.x]`, :244) */
  float *block_vars;           /* [ceil(N/256)] per-block population variance (reference `variances`, :245) */
  uint64_t *hist;              /* [hist_bins+2] (host pointer always; filled unless NO_SYNC) */
  double *stats_dev;           /* DEVICE_POINTERS only: device buffer of smmc_stats_len(hist_bins) doubles */
  /* scalars filled by the call */
  uint64_t count;              /* N */
  uint64_t below_count;
  double mean;                 /* mean of the N finals */
  double m2;                   /* sum (x - mean)^2   (population variance = m2 / count) */
  float device_ms;             /* CUDA-event time of the kernel(s) of this call */
  int32_t gpu_launches;        /* kernels launched by this call */
} smmc_result;

/* ---- run ---------------------------------------------------------------------------------------------- */

/* One blocking simulation call on cfg->device.  Host buffers by default: the table is copied H2D, results D2H,
 * device workspace comes from a per-device cache owned by the library (freed by smmc_shutdown).
 * Replaces: mc_simulations_gpu_launcher / _reduceBlock_launcher / _multi_gpu_launcher_async
 * (reference src/simulations.cu:345-473, :576-655). */
smmc_status smmc_run(const smmc_config *cfg, smmc_result *out);

/* Multi-GPU, single process, one host thread (the reference's driver model, src/simulations.cu:576-655): shards
 * [0,N) into n_gpus contiguous 256-aligned path ranges by RNG offset, runs them concurrently on devices
 * 0..n_gpus-1 and merges the partial statistics (see below: on the devices).  Results are identical to a 1-GPU run of the
 * same config (histogram / counts bit-identical; mean/M2 to fp64 rounding). Host pointers only. */
smmc_status smmc_run_multi(const smmc_config *cfg, int n_gpus, smmc_result *out);
/* (statistics-only runs of smmc_run_multi merge on the DEVICES: one communicator per device from ncclCommInitAll, one
 * grouped ncclAllReduce, one 864-byte download from device 0; runs that return all finals stream them through the host
 * anyway and sum the per-chunk vectors there.  When NCCL cannot be loaded or ncclCommInitAll fails, smmc_run_multi and
 * smmc_quantiles(n_gpus > 1) download every device's vector and add them on the host: they never fail for want of NCCL.) */

/* ---- one rank per GPU: the collective behind the C ABI ---------------------------------------------------------
 * The reference drives all GPUs from one host thread and re-runs the SAME paths on every device
 * (src/simulations.cu:576-655, :621-622).  Here a job of one process (or thread) per GPU shards ONE global run by RNG
 * offset and merges the O(KB) statistics vector with a single ncclAllReduce(ncclFloat64, ncclSum) enqueued on the
 * same stream as the kernel: count / below / S1 / S2 / histogram are all sum-reducible and the integer fields are
 * exact in fp64, so the merged histogram and counts are bit-identical for every number of ranks.  No per-path data
 * ever crosses NVLink.  NCCL is bound at run time (dlopen of libnccl.so.2): when it is absent the rank entry points
 * (smmc_comm_*, smmc_run_rank, smmc_quantiles_rank) fail with SMMC_ERR_UNSUPPORTED -- ranks of different processes have
 * no host-side stand-in. */
typedef struct smmc_comm smmc_comm;     /* opaque: an ncclComm_t plus the rank's device */
#define SMMC_COMM_ID_BYTES 128          /* = NCCL_UNIQUE_ID_BYTES */
/* rank 0: ncclGetUniqueId into id[SMMC_COMM_ID_BYTES]; distribute the bytes to the other ranks by any means. */
smmc_status smmc_comm_unique_id(void *id);
/* every rank: ncclCommInitRank on `device` (collective call). */
smmc_status smmc_comm_init_rank(const void *id, int32_t n_ranks, int32_t rank, int32_t device, smmc_comm **comm);
/* wraps a communicator the application already owns (ncclComm_t as void*); not destroyed by smmc_comm_destroy. */
smmc_status smmc_comm_adopt(void *nccl_comm, int32_t device, smmc_comm **comm);
smmc_status smmc_comm_destroy(smmc_comm *comm);
smmc_status smmc_comm_info(const smmc_comm *comm, int32_t *n_ranks, int32_t *rank, int32_t *device, int32_t *nccl_version);
/* how many ncclAllReduce calls smmc_run_rank has issued through `comm` so far (one per call, on every path) */
smmc_status smmc_comm_allreduce_count(const smmc_comm *comm, uint64_t *count);
/* This rank's part of the GLOBAL run `cfg` (cfg->n_simulations, cfg->stream_offset describe the whole run; cfg->device
 * is ignored in favour of the communicator's device): rank r simulates the r-th of n_ranks contiguous ranges of whole
 * 256-path blocks (the partition of smmc_run_multi; smmc_rank_range reports it), then
 *   kernel -> ncclAllReduce(statistics vector) on the same stream -> (unless NO_SYNC) download + finalize,
 * so count / mean / m2 / below_count / hist in `out` describe the GLOBAL run on every rank.  finals / block_means /
 * block_vars, when requested, cover this rank's own range only (they never travel).  Flags as for smmc_run: host
 * pointers (blocking) by default; DEVICE_POINTERS [+ NO_SYNC] runs on cfg->stream, fully asynchronously with NO_SYNC
 * (read the merged vector from out->stats_dev).  Collective: every rank must call it.  Host-pointer runs that return block
 * partials of >= 16 Mi paths per rank are launched in two parts (the first part's download hides behind the second
 * kernel); the first part's statistics are added into the second part's on the device, so this path too issues exactly ONE
 * all-reduce per call (smmc_comm_allreduce_count lets a caller check). */
smmc_status smmc_run_rank(const smmc_config *cfg, smmc_result *out, smmc_comm *comm);
/* first path (relative to cfg->stream_offset) and number of paths of rank `rank` of `n_ranks` for a run of n paths */
smmc_status smmc_rank_range(int64_t n_simulations, int32_t n_ranks, int32_t rank, int64_t *first, int64_t *count);
/* Exact order statistics of the GLOBAL run, one rank per GPU: every rank replays its own range; the digit histograms
 * (exact integers) are merged by ncclAllReduce(ncclUint64, ncclSum) per pass.  values[t] on every rank. Host pointers. */
smmc_status smmc_quantiles_rank(const smmc_config *cfg, smmc_comm *comm, int32_t n_targets, const uint64_t *ranks,
                                float *values, float *device_ms);

/* CUDA-event durations (ms) of the simulation kernel of the most recent launches on `device` (oldest first, at
 * most 256 are kept), whichever entry point or flags enqueued them.  Blocks until those kernels have finished.
 * This is how an asynchronous caller (SMMC_FLAG_NO_SYNC) times the kernel on the stream it was launched on. */
smmc_status smmc_kernel_times(int32_t device, float *ms_out, int32_t max_n, int32_t *n_out);
/* WHICH kernel each of those entries was (same ring, same order): a full simulation / replay of the paths, or the pass
 * over the candidate keys that replaces the third replay of a select.  Lets a caller (and the tests) see what a call
 * launched without inferring it from durations.  No counterpart in the reference (it has one kernel per call). */
#define SMMC_KERNEL_SIMULATION 1          /* smmc::mc_kernel<...> */
#define SMMC_KERNEL_SELECT_CANDIDATES 2   /* smmc::select_candidates_kernel */
smmc_status smmc_kernel_kinds(int32_t device, int32_t *kinds_out, int32_t max_n, int32_t *n_out);

int32_t smmc_stats_len(int32_t hist_bins);
/* stats: HOST vector of smmc_stats_len(bins) doubles (e.g. after an allreduce). hist may be NULL. */
smmc_status smmc_stats_finalize(const double *stats, int32_t hist_bins, float initial_capital,
                                uint64_t *count, uint64_t *below_count, double *mean, double *m2, uint64_t *hist);

/* ---- the remaining reference GPU entry points ------------------------------------------------------------ */

/* mean of n host floats (reference reduce_mean_gpu, src/simulations.cu:269-341), single-pass fp64-accumulating
 * device reduction; 64-bit indexing. */
smmc_status smmc_reduce_mean(const float *vec, int64_t n, int32_t device, float *mean_out);
/* The same reduction over n floats already RESIDENT on `device` (e.g. the `finals` buffer of a DEVICE_POINTERS run):
 * launched on `stream` (NULL = legacy default stream), blocks until done.  vec_dev: any 4-byte-aligned device pointer.
 * kernel_ms (nullable): CUDA-event duration of the reduction kernel on that stream -- the kernel is HBM-bound at
 * 4 bytes per element. */
smmc_status smmc_reduce_mean_device(const float *vec_dev, int64_t n, int32_t device, void *stream, float *mean_out,
                                    float *kernel_ms);
/* The mean of n device-resident floats for hosts in which this library must not allocate device memory.  NO COUNTERPART IN
 * THE REFERENCE (its reduce_mean_gpu cudaMallocs on every call, src/simulations.cu:283-289).
 * Callers it is for: a service whose own arena or caching allocator already owns the whole GPU, so that a cudaMalloc made by
 * a plugin fails; a host with work in flight on other streams, which a cudaMalloc/cudaFree would stall device-wide; a
 * deployment with a memory budget fixed at load time.  Every other compute entry point here builds a per-device context
 * and grows cudaMalloc'ed buffers on demand; this one calls no cudaMalloc / cudaFree / allocating routine at all.
 * Its scratch (one fp64 partial per 32768 floats + ticket + sum) is a fixed 262 400-byte __device__ array inside the
 * library: every GPU that loads the library's module holds it until the process exits.
 *   - n must be <= SMMC_NOALLOC_MAX_N; larger n returns SMMC_ERR_INVALID_ARG (use smmc_reduce_mean_device or reduce in pieces).
 *   - vec_dev: any 4-byte-aligned pointer into `device` memory; mean_out: HOST float.  Launched on `stream` (NULL = legacy
 *     default stream); blocks until the result is on the host.  Same arithmetic as smmc_reduce_mean_device (fp64
 *     accumulation, mean = (float)(sum / n)); the order of the additions differs, so the two agree to fp64 rounding.
 *   - All calls on a device share the one array and it keeps the previous call's contents; each call clears the words that
 *     must start at zero.  Calls are serialised inside the process; GPU work of two calls that overlapped would race.
 *   - The FIRST call on each device resolves the array's device address (cudaGetSymbolAddress, may load the module): make
 *     that call before starting any stream capture. */
#define SMMC_NOALLOC_MAX_N 1073741824LL /* 2^30 floats = 4 GiB */
smmc_status smmc_reduce_mean_device_noalloc(const float *vec_dev, int64_t n, int32_t device, void *stream, float *mean_out);
/* out = a + b on the device (reference vector_add_gpu, src/gpu.cu:17-47). Host pointers. */
smmc_status smmc_vector_add(float *out, const float *a, const float *b, int64_t n, int32_t device);
/* The same over DEVICE-resident vectors (any 4-byte alignment; 16-byte aligned pointers take the vector path), on `stream`
 * (NULL = legacy default stream); blocks until done.  kernel_ms (nullable): CUDA-event duration of the kernel, which is
 * HBM-bound at 12 bytes per element. */
smmc_status smmc_vector_add_device(float *out_dev, const float *a_dev, const float *b_dev, int64_t n, int32_t device,
                                   void *stream, float *kernel_ms);

/* ---- exact order statistics of the finals without writing them to HBM ------------------- */

/* The reference sorts all N finals on the host and reads quartiles off the sorted vector (update_quartiles,
 * reference examples/visualize_returns_gpu.cpp:83-110: v[0], v[N/4], v[N/2], v[N/4 + N/2], v[N-1]).  Here the
 * value of rank r (0-based, in ascending order) is found by a radix select over the finals' order-preserving
 * 32-bit keys in three passes (11 + 11 + 10 key bits).  Every pass REPLAYS the simulation -- the RNG is
 * counter-based, so a replay reproduces every path bit for bit -- and only builds digit histograms in shared
 * memory; no per-path value is ever stored.  The result is exactly sorted(finals)[r].
 *
 * Pass protocol (this is what a one-process-per-GPU job runs, summing the histograms over ranks in between):
 *   smmc_select_begin(&st, n, ranks);
 *   while (st.bits_done < 32) { smmc_select_pass(cfg_of_my_shard, &st, hist); <all-reduce hist>; smmc_select_advance(&st, hist); }
 *   smmc_select_values(&st, values);
 * smmc_quantiles() runs the same protocol on n_gpus devices of this process. */
#define SMMC_SELECT_MAX_TARGETS 8
#define SMMC_SELECT_BINS 2048

typedef struct smmc_select_state {
  int32_t n_targets;                          /* 1..SMMC_SELECT_MAX_TARGETS */
  int32_t bits_done;                          /* key bits resolved so far: 0 -> 11 -> 22 -> 32 */
  uint64_t rank[SMMC_SELECT_MAX_TARGETS];     /* rank of target t among the keys that share its prefix */
  uint32_t prefix[SMMC_SELECT_MAX_TARGETS];   /* the top bits_done bits of target t's key */
} smmc_select_state;

smmc_status smmc_select_begin(smmc_select_state *st, int32_t n_targets, const uint64_t *ranks);
/* One replay of cfg's paths.  hist: SMMC_SELECT_MAX_TARGETS * SMMC_SELECT_BINS counters, row t = digit histogram
 * of the keys under target t's prefix (first pass, bits_done == 0: only row 0 is produced and serves every target).
 * Host pointer by default; with SMMC_FLAG_DEVICE_POINTERS `hist` (and cfg->returns) are device pointers, the call
 * is asynchronous on cfg->stream and rows >= the number produced are zeroed.  Counters are exact and
 * SUM-reducible across shards. */
smmc_status smmc_select_pass(const smmc_config *cfg, const smmc_select_state *st, uint64_t *hist);
/* smmc_run with the FIRST select pass riding along: the same launch that produces the statistics / histogram / partials /
 * finals of `cfg` also builds the first 11-bit digit histogram (st->bits_done must be 0), so statistics + order statistics
 * cost three replays instead of four.  hist as in smmc_select_pass (host); host pointers only. */
smmc_status smmc_run_select(const smmc_config *cfg, smmc_result *out, const smmc_select_state *st, uint64_t *hist);
/* Consume the (globally summed) HOST histogram of the pass just run: fixes the next 11/11/10 key bits per target. */
smmc_status smmc_select_advance(smmc_select_state *st, const uint64_t *hist);
/* After the last pass (bits_done == 32): the order statistics themselves. */
smmc_status smmc_select_values(const smmc_select_state *st, float *values);
/* values[t] = sorted(finals)[ranks[t]] of cfg's N paths, sharded over devices 0..n_gpus-1 like smmc_run_multi.
 * Host pointers.  device_ms (nullable): summed kernel time of the three passes (max over devices per pass). */
smmc_status smmc_quantiles(const smmc_config *cfg, int32_t n_gpus, int32_t n_targets, const uint64_t *ranks,
                           float *values, float *device_ms);

/* ---- trajectory mode ------------------------------------------------------------------- */

/* Full value paths of n_paths paths starting at cfg->stream_offset, from the SAME RNG streams and arithmetic as
 * the statistics kernel: values[p*(P+1) + 0] = V0, values[p*(P+1) + m + 1] = value after month m, so that
 * values[p*(P+1) + P] equals finals[p] of smmc_run bit for bit (reference: mc_simulations_keepdata,
 * src/simulations.cpp:139-202, `mc_data[id] = many_updates(...)`, P+1 values per path).
 * returns_pct (nullable) [n_paths * P]: the monthly return of each step in PERCENT, the reference's unit
 * (bootstrap modes: the table entry drawn; Gaussian modes: 100 (g - 1)).  Host pointers; cfg->n_simulations is
 * ignored. */
smmc_status smmc_trajectories(const smmc_config *cfg, int32_t n_paths, float *values, float *returns_pct);

/* ---- data-format edges: host code, no device needed ---------------------------------------- */

/* Reads the column named "returns" of a CSV file with a header line (reference read_historical_returns,
 * src/simulations.cpp:83-93: io::CSVReader<1>, read_header(ignore_extra_column, "returns")).  Two-call pattern:
 * out == NULL or cap too small still reports the row count in *n_out (status SMMC_ERR_INVALID_ARG if cap was
 * too small and out != NULL). */
smmc_status smmc_read_returns_csv(const char *path, float *out, int32_t cap, int32_t *n_out);
/* Writes one trajectory in the reference's output format (write_data_file, src/helpers.cpp:23-39, read back by
 * python/plot_returns.py:40):  "Returns,,r0,r1,...,\nValues,v0,v1,...,"  with ostream default float formatting. */
smmc_status smmc_write_data_file(const char *path, const float *returns, int32_t n_returns, const float *values,
                                 int32_t n_values);

/* ---- launch planning (host only; used by the tests and for capacity questions) --------------------------------- */

typedef struct smmc_plan {
  int32_t grid;            /* CTAs (one per SM) */
  int32_t threads;         /* threads per CTA: 32 x warps, warps chosen to fill the last wave of 256-path blocks */
  int32_t table_copies;    /* bank-private copies of the return table per entry (32 = conflict-free gather) */
  int32_t ticket;          /* 1: 256-path blocks are handed to the warps by an atomic ticket and their moments merged in block
                            * order (launches of 1..48 waves); 0: static assignment.  Results do not depend on it beyond the
                            * fp64 rounding of mean / M2 (integer outputs and every per-path / per-block output are identical) */
  uint64_t smem_bytes;     /* dynamic shared memory per CTA */
} smmc_plan;
/* The launch smmc_run would use for cfg on a device with n_sms SMs and smem_optin bytes of opt-in shared memory
 * (B200: 148, 232448).  select_targets > 0 plans a select pass.  SMMC_ERR_UNSUPPORTED if the table cannot fit. */
smmc_status smmc_plan_launch(const smmc_config *cfg, int32_t select_targets, int32_t n_sms, uint64_t smem_optin,
                             smmc_plan *plan);

/* ---- introspection used by the parity tests (device code paths, tiny outputs) ---------------------------------- */

/* Raw per-month draws of `n_paths` paths starting at cfg->stream_offset, computed ON THE DEVICE with the same
 * device functions the simulation kernel uses.  out[p * n_periods + m]:
 *   bootstrap modes: table index drawn for month m;  GAUSSIAN: radius field (even m) / hashed angle word (odd m)
 *   of the Box-Muller pair behind months m, m+1;  SOBOL_GAUSSIAN: the (shifted) Sobol word of month m. */
smmc_status smmc_debug_draws(const smmc_config *cfg, int32_t n_paths, uint32_t *out);
/* Philox4x32-10 block function on the device: out[4*i..] = philox(ctr[4*i..], key[0..1]) for i < n. */
smmc_status smmc_debug_philox(const uint32_t *ctr, const uint32_t *key, int32_t n, int32_t device, uint32_t *out);
/* The same with `rounds` in {7, 10} (smmc_config.philox_rounds): the device code behind both stream contracts. */
smmc_status smmc_debug_philox_rounds(const uint32_t *ctr, const uint32_t *key, int32_t rounds, int32_t n, int32_t device,
                                     uint32_t *out);
/* Test hook: selects among the compiled variants of a kernel.  Every variant of a mode consumes the same INTEGER draws
 * (Philox words, table indices, Sobol words) and the integer / table modes return identical results under every key; the
 * tests use the hook to prove that.  The two Gaussian modes are bit-reproducible only for a fixed build AND a fixed table
 * mask: "gauss_tbl" / "sobol_tbl" move sin/cos between the MUFU unit and the shared-memory table, which changes finals by
 * about 1e-5 relative, and the default GAUSSIAN mask depends on philox_rounds (5 for 10 rounds, 7 for 7).  The Gaussian
 * finals also changed between ABI 2 and ABI 3 (table pre-scaled by sqrt(-k)).  The hook is process-wide: changing a mask
 * between smmc_run, smmc_trajectories and smmc_quantiles of one configuration breaks their bit-identity, so set it once,
 * before the first call, or not at all.  Keys: "ilp" (paths per lane in flight: 2|4|8), "warps" (warps per CTA, 1..32), "reduce"
 * (reduce_mean kernel: 1 = streaming loads, 2 = TMA ring), "gauss_tbl" (GAUSSIAN: 4-bit mask of the Box-Muller pairs of a
 * call that take sin/cos from the shared-memory table; 0 = all MUFU), "sobol_tbl" (SOBOL_GAUSSIAN: 0 = MUFU, 1 = table), "sobol_unroll",
 * "finals_chunks" (host-pointer finals pipeline: 1 = one monolithic launch), "finals_threads" (its host copy-out threads),
 * "select_candidates" (0 = the third select pass always replays), "ticket" (256-path blocks handed to warps 0 = statically,
 * 1 = by an atomic ticket with the block moments merged in block order; default: ticket for launches of 1..48 waves).
 * "host_merge" (1 = smmc_run_multi / smmc_quantiles add their shards on the host, as they do when NCCL cannot be loaded),
 * "verbose" (1 = smmc_comm_init_rank prints one line per rank on stderr; the library prints nothing otherwise).
 * value < 0 (or 0 where 0 is not a choice) restores the default; key NULL resets everything.  Process-wide; nothing in the library reads the environment. */
smmc_status smmc_debug_set_variant(const char *key, int32_t value);

/* ---- housekeeping ---------------------------------------------------------------------------------------- */

const char *smmc_last_error(void);      /* thread-local message of the last failing call */
const char *smmc_version(void);
int32_t smmc_device_count(void);        /* number of usable (cc 10.x) CUDA devices; 0 if none */
void smmc_shutdown(void);               /* free cached device workspaces */

#ifdef __cplusplus
}
#endif
#endif /* SMMC_H_ */
