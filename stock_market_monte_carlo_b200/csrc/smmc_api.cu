// smmc_api.cu -- host side of the C ABI declared in include/smmc.h.
//
// Replaces the reference's launchers (src/simulations.cu:345-473 single GPU, :479-655 multi GPU plans): device
// workspace is cached per device instead of being cudaMalloc'ed/freed on every call, the table upload and the
// result download are the only host<->device copies, and multi-GPU runs shard the path-id range (RNG offset)
// instead of re-running identical paths on every device (reference :621-622).
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#ifdef __linux__
#include <sys/mman.h>
#endif
#include <atomic>
#include <chrono>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/smmc.h"
#include "smmc_kernels.cuh"
#include "smmc_nccl.h"

// opaque in include/smmc.h
struct smmc_comm {
  smmc::NcclComm nccl = nullptr;
  int n_ranks = 1, rank = 0, device = 0;
  bool owned = false;
  unsigned long long allreduces = 0;  // ncclAllReduce calls issued through this communicator (smmc_comm_allreduce_count)
};

namespace smmc {
const uint32_t *sobol_table();  // smmc_sobol_data.cpp
int sobol_table_dims();
}  // namespace smmc

namespace {

thread_local char g_err[512] = "";

smmc_status fail(smmc_status st, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return st;
}

}  // namespace

namespace smmc {
// error reporting for the other translation units of the library (smmc_io.cpp)
smmc_status set_error(smmc_status st, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return st;
}
}  // namespace smmc

namespace {

// Every entry point may switch the current device; the caller's choice is restored on every return path.
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() {
    if (cudaGetDevice(&prev) != cudaSuccess) {
      cudaGetLastError();
      prev = -1;
    }
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// kernel-variant selection of the test hook smmc_debug_set_variant (never the environment)
smmc::Variant g_variant;
std::mutex g_variant_mu;
smmc::Variant current_variant() {
  std::lock_guard<std::mutex> lk(g_variant_mu);
  return g_variant;
}

#define CU(call)                                                                                       \
  do {                                                                                                 \
    cudaError_t e__ = (call);                                                                          \
    if (e__ != cudaSuccess)                                                                            \
      return fail(SMMC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

// grow-only device / pinned buffers
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

constexpr int kEventRing = 256;

// everything that determines the paths of a select replay (validity key of the candidate buffer)
struct CandFingerprint {
  int32_t mode, periods, n_returns, rounds, tbl_mask, pad;
  int64_t n;
  uint64_t seed, offset, table, schedule;
  float v0, gm, gs, wfixed, wpct, pad2;
};

struct Context {
  int device = -1;
  bool ready = false;
  int n_sms = 0;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;
  // ring of event pairs around the simulation kernel of the last kEventRing launches (smmc_kernel_times)
  cudaEvent_t ev0[kEventRing] = {}, ev1[kEventRing] = {};
  uint64_t ev_head = 0;  // launches recorded so far
  int ev_kind[kEventRing] = {};  // SMMC_KERNEL_* of each ring entry (smmc_kernel_kinds)
  cudaEvent_t red0 = nullptr, red1 = nullptr;  // around the reduce_mean kernel (smmc_reduce_mean_device)
  DevBuf returns, sorted, tbl, work, stats, finals, bmeans, bvars, scratch, sobol_full, sobol_shift, selhist, sincos, block_moments;
  bool sincos_ready = false;
  uint32_t sincos_kbits = 0;  // bits of the GaussConsts::k the table was scaled for
  bool sobol_full_ready = false;
  // Small host-built tables (withdrawal schedule, per-launch Sobol tables) are cached by content in a ring of device
  // versions: an unchanged table costs no upload at all; a changed one goes into the next slot, ordered behind the last
  // kernel that read that slot by a stream-side event wait -- the host never blocks (SMMC_FLAG_NO_SYNC stays asynchronous).
  static constexpr int kVersions = 4;
  struct Versioned {
    DevBuf dev[kVersions];
    void *pinned[kVersions] = {};
    size_t pinned_cap[kVersions] = {};
    cudaEvent_t copied[kVersions] = {}, last_read[kVersions] = {};
    bool read_pending[kVersions] = {};
    std::vector<unsigned char> content[kVersions];
    uint64_t tag[kVersions] = {};
    int next = 0;
  } flow_v, sobol_v;
  // cross-stream ordering of the cached workspace: every launch records `last_use` on its stream; a launch on a different
  // stream waits for it (stream-side) before touching the shared buffers
  cudaEvent_t last_use = nullptr;
  cudaStream_t last_stream = nullptr;
  bool has_last_use = false;
  // candidate keys of the last second select replay (select_pass_enqueue)
  DevBuf cand_keys, cand_cursor;
  bool cand_valid = false;
  CandFingerprint cand_fp;
  int cand_groups = 0;
  uint32_t cand_prefix[SMMC_SELECT_MAX_TARGETS] = {};
  unsigned long long cand_capacity = 0;
  // pipelined all-finals path (run_finals_pipelined): second stream for the downloads, pinned staging ring, chunk events
  cudaStream_t copy_stream = nullptr;
  static constexpr int kStageSlots = 3;   // pinned staging slots (host side of the D2H copies)
  static constexpr int kChunkDepth = 8;   // kernel chunks enqueued ahead = device-side ring of chunk buffers
  void *stage[kStageSlots] = {};
  size_t stage_cap = 0;
  DevBuf chunk_stats;                     // [kChunkDepth][stats_len] doubles
  double *chunk_stats_host = nullptr;     // pinned, same shape
  size_t chunk_stats_host_cap = 0;
  cudaEvent_t kdone[kChunkDepth] = {}, cdone[kChunkDepth] = {}, sdone[kChunkDepth] = {};
  double *stats_host = nullptr;  // pinned
  size_t stats_host_cap = 0;
  std::mutex mu;
};

constexpr int kMaxDevices = 64;
Context g_ctx[kMaxDevices];
std::mutex g_ctx_mu;

int g_device_count = -2;  // -2 = not probed

int probe_devices() {
  std::lock_guard<std::mutex> lk(g_ctx_mu);
  if (g_device_count != -2) return g_device_count;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  int usable = 0;
  for (int d = 0; d < n && d < kMaxDevices; d++) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) usable = d + 1;
  }
  g_device_count = usable;
  return usable;
}

smmc_status get_context(int device, Context **out) {
  const int n = probe_devices();
  if (n <= 0) return fail(SMMC_ERR_NO_DEVICE, "no CUDA device of compute capability 10.x (sm_100a) is visible; this library has no CPU fallback");
  if (device < 0 || device >= n) return fail(SMMC_ERR_INVALID_ARG, "device %d out of range [0,%d)", device, n);
  Context &c = g_ctx[device];
  std::lock_guard<std::mutex> lk(g_ctx_mu);
  if (!c.ready) {
    CU(cudaSetDevice(device));
    int major = 0, sms = 0, optin = 0;
    CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) return fail(SMMC_ERR_NO_DEVICE, "device %d is not compute capability 10.x", device);
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    CU(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    for (int i = 0; i < kEventRing; i++) {
      CU(cudaEventCreate(&c.ev0[i]));
      CU(cudaEventCreate(&c.ev1[i]));
    }
    CU(cudaEventCreate(&c.red0));
    CU(cudaEventCreate(&c.red1));
    CU(cudaEventCreateWithFlags(&c.last_use, cudaEventDisableTiming));
    c.has_last_use = false;
    CU(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < Context::kChunkDepth; i++) {
      CU(cudaEventCreateWithFlags(&c.kdone[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&c.cdone[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&c.sdone[i], cudaEventDisableTiming));
    }
    c.device = device;
    c.n_sms = sms;
    c.smem_optin = (size_t)optin;
    c.ready = true;
  }
  *out = &c;
  return SMMC_OK;
}

bool is_bootstrap(int mode) {
  return mode == SMMC_MODE_BOOTSTRAP_COMPAT || mode == SMMC_MODE_BOOTSTRAP || mode == SMMC_MODE_SOBOL_BOOTSTRAP;
}
bool has_withdrawals(const smmc_config *cfg) {
  return cfg->withdraw_fixed != 0.0f || cfg->withdraw_pct != 0.0f || cfg->withdraw_pct_schedule != nullptr;
}
bool is_sobol(int mode) { return mode == SMMC_MODE_SOBOL_BOOTSTRAP || mode == SMMC_MODE_SOBOL_GAUSSIAN; }

smmc_status validate(const smmc_config *cfg, const smmc_result *out, bool read_table = true) {
  if (!cfg || !out) return fail(SMMC_ERR_INVALID_ARG, "null config/result");
  if (cfg->abi_version != SMMC_ABI_VERSION)
    return fail(SMMC_ERR_INVALID_ARG, "abi_version %u != %u", cfg->abi_version, SMMC_ABI_VERSION);
  if (cfg->mode < SMMC_MODE_BOOTSTRAP_COMPAT || cfg->mode > SMMC_MODE_SOBOL_GAUSSIAN)
    return fail(SMMC_ERR_INVALID_ARG, "unknown mode %d", cfg->mode);
  if (cfg->n_simulations < 0) return fail(SMMC_ERR_INVALID_ARG, "n_simulations < 0");
  if (cfg->n_simulations > (1ll << 38))  // per-CTA 32-bit histogram / digit counters: 2^38 / #SMs < 2^32
    return fail(SMMC_ERR_INVALID_ARG, "n_simulations per call must be <= 2^38; split the run by stream_offset");
  if (cfg->n_periods < 0) return fail(SMMC_ERR_INVALID_ARG, "n_periods < 0");
  if (cfg->stream_offset % SMMC_BLOCK) return fail(SMMC_ERR_INVALID_ARG, "stream_offset must be a multiple of %d", SMMC_BLOCK);
  if (is_bootstrap(cfg->mode)) {
    if (!cfg->returns || cfg->n_returns < 1) return fail(SMMC_ERR_INVALID_ARG, "bootstrap modes need a non-empty returns table");
    if (cfg->n_returns >= (1 << 21)) return fail(SMMC_ERR_INVALID_ARG, "n_returns must be < 2^21");
  }
  if (cfg->mode == SMMC_MODE_BOOTSTRAP_COMPAT && cfg->stream_offset + (uint64_t)cfg->n_simulations > 0x7fffffffull)
    return fail(SMMC_ERR_INVALID_ARG, "BOOTSTRAP_COMPAT uses the reference's 32-bit path ids: stream_offset + N must be < 2^31");
  if (is_sobol(cfg->mode)) {
    if (cfg->n_periods > 1111) return fail(SMMC_ERR_INVALID_ARG, "Sobol modes support at most 1111 periods (dimensions)");
    if (cfg->stream_offset + (uint64_t)cfg->n_simulations + 256ull > 0x100000000ull)
      return fail(SMMC_ERR_INVALID_ARG, "Sobol modes support point indices < 2^32");
  }
  if (cfg->hist_bins < 0 || cfg->hist_bins > SMMC_MAX_HIST_BINS) return fail(SMMC_ERR_INVALID_ARG, "hist_bins out of range");
  if (cfg->hist_bins > 0 && !(cfg->hist_hi > cfg->hist_lo)) return fail(SMMC_ERR_INVALID_ARG, "hist_hi must be > hist_lo");
  if (cfg->hist_bins > 0 && (!std::isfinite(cfg->hist_lo) || !std::isfinite(cfg->hist_hi)))
    return fail(SMMC_ERR_INVALID_ARG, "hist_lo / hist_hi must be finite");
  if (!std::isfinite(cfg->withdraw_fixed) || !std::isfinite(cfg->withdraw_pct))
    return fail(SMMC_ERR_INVALID_ARG, "withdraw_fixed / withdraw_pct must be finite");
  if (!std::isfinite(cfg->initial_capital)) return fail(SMMC_ERR_INVALID_ARG, "initial_capital must be finite");
  if ((cfg->mode == SMMC_MODE_GAUSSIAN || cfg->mode == SMMC_MODE_SOBOL_GAUSSIAN) &&
      (!std::isfinite(cfg->gauss_mean) || !std::isfinite(cfg->gauss_std)))
    return fail(SMMC_ERR_INVALID_ARG, "gauss_mean / gauss_std must be finite");
  if (cfg->philox_rounds != 0 && cfg->philox_rounds != 7 && cfg->philox_rounds != 10)
    return fail(SMMC_ERR_INVALID_ARG, "philox_rounds must be 0 (= 10), 7 or 10");
  if (read_table && is_bootstrap(cfg->mode) && !(cfg->flags & SMMC_FLAG_DEVICE_POINTERS)) {
    // a NaN / inf in the table would poison every path; reject it here (device-pointer tables are the caller's business:
    // the sorted-table kernel still assigns every rank exactly once)
    for (int i = 0; i < cfg->n_returns; i++)
      if (!std::isfinite(cfg->returns[i])) return fail(SMMC_ERR_INVALID_ARG, "returns[%d] is not finite", i);
  }
  if ((cfg->flags & SMMC_FLAG_NO_SYNC) && !(cfg->flags & SMMC_FLAG_DEVICE_POINTERS))
    return fail(SMMC_ERR_INVALID_ARG, "SMMC_FLAG_NO_SYNC requires SMMC_FLAG_DEVICE_POINTERS");
  if ((cfg->flags & SMMC_FLAG_NO_SYNC) && !out->stats_dev)
    return fail(SMMC_ERR_INVALID_ARG, "SMMC_FLAG_NO_SYNC requires stats_dev");
  return SMMC_OK;
}

smmc::PhiloxKeys make_keys(uint64_t seed) {
  smmc::PhiloxKeys k;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; r++) {
    k.k0[r] = k0;
    k.k1[r] = k1;
    k0 += smmc::kPhiloxW0;
    k1 += smmc::kPhiloxW1;
  }
  return k;
}

// does this configuration take sin/cos from the shared-memory table?  A function of (mode, variant hook) only -- never of
// what else has to fit into shared memory -- so that every launch of a configuration (statistics, select replays, shards,
// chunks) computes bit-identical paths.
// Which Box-Muller pairs take sin/cos from the shared-memory table: GAUSSIAN a 4-bit mask over the pairs of a Philox call,
// SOBOL_GAUSSIAN an 8-bit mask over a lane's path slots.  Only the masks a build has kernels for are handed out (anything
// else falls back to the default), so the statistics kernel, the trajectory kernel and the candidate fingerprint agree.
int gauss_table_mask(int mode, const smmc::Variant &v, int philox_rounds) {
  if (mode == SMMC_MODE_GAUSSIAN) {
    const int def = smmc::gauss_tbl_default(philox_rounds == 7 ? 7 : 10);
    if (v.gauss_tbl < 0) return def;
#ifdef SMMC_EXPERIMENT
    return v.gauss_tbl;
#else
    return (v.gauss_tbl == 0 || v.gauss_tbl == def) ? v.gauss_tbl : def;
#endif
  }
  if (mode == SMMC_MODE_SOBOL_GAUSSIAN) {
    if (v.sobol_tbl < 0) return smmc::kSobolTblDefault;
    if (v.sobol_tbl == 0) return 0;
    if (v.sobol_tbl == 1 || v.sobol_tbl == 0xFF) return 0xFF;
#ifdef SMMC_EXPERIMENT
    return v.sobol_tbl;
#else
    return smmc::kSobolTblDefault;
#endif
  }
  return 0;
}
bool uses_sincos_table(int mode, const smmc::Variant &v, int philox_rounds) { return gauss_table_mask(mode, v, philox_rounds) != 0; }

// the RNG-stream / arithmetic constants of a config (shared by the introspection and trajectory kernels; enqueue()
// sets the same fields for the simulation kernel)
void fill_stream_params(const smmc_config *cfg, smmc::KParams *p) {
  p->path0 = cfg->stream_offset;
  p->n_periods = cfg->n_periods;
  p->n_returns = cfg->n_returns;
  p->compat_seed = (uint32_t)cfg->seed;
  p->keys = make_keys(cfg->seed);
  p->philox_rounds = cfg->philox_rounds == 7 ? 7 : 10;
  p->gauss_tbl_mask = gauss_table_mask(cfg->mode, current_variant(), cfg->philox_rounds);
  const float mu = cfg->gauss_mean * 0.01f, sigma = cfg->gauss_std * 0.01f;
  p->gc.a = 1.0f + mu;
  p->gc.k = (float)(-2.0 * 0.69314718055994530942 * (double)sigma * (double)sigma);
  p->sigma_sqrt2 = (float)(1.41421356237309504880 * (double)sigma);
  p->one = 0x3F800000u;
  p->cell_mask = smmc::kSinCosCellMask;
  p->v0 = cfg->initial_capital;
  p->idx_scale = (float)cfg->n_returns * 2.3283064365386963e-10f;  // exact: n < 2^24
  p->bc.c21 = (float)(4 * cfg->n_returns);                          // exact: n < 2^21 (validated)
  p->bc.k21 = (float)(8388608 - 4 * cfg->n_returns);
  p->bc.c22 = (float)(2 * cfg->n_returns);
  p->bc.k22 = (float)(8388608 - 2 * cfg->n_returns);
}

// host Philox4x32-10 (only for the per-month Sobol digital shift; same counter layout as the device stream)
void philox_host(const uint32_t ctr[4], uint64_t seed, uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; r++) {
    const uint64_t p0 = (uint64_t)smmc::kPhiloxM0 * c0, p1 = (uint64_t)smmc::kPhiloxM1 * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
    k0 += smmc::kPhiloxW0;
    k1 += smmc::kPhiloxW1;
  }
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}

// digital shift of Sobol dimension m: 0 for seed 0 (plain sequence), else word (m & 3) of the Philox call
// (q = 0, path = m >> 2, tag "SOBL") keyed by the seed
uint32_t sobol_shift_host(int m, uint64_t seed) {
  if (seed == 0) return 0;
  const uint32_t ctr[4] = {0u, (uint32_t)(m >> 2), smmc::kTagSobolShift, 0u};
  uint32_t w[4];
  philox_host(ctr, seed, w);
  return w[m & 3];
}

// ---- content-cached, versioned uploads (Context::Versioned) ---------------------------------------------------------
// Returns the slot that holds (tag, host[0..bytes)) on the device, uploading on `st` if no slot does.  host == nullptr
// only looks up by tag (-1 = not cached).
int versioned_find(const Context::Versioned &v, uint64_t tag, const void *host, size_t bytes) {
  for (int i = 0; i < Context::kVersions; i++) {
    if (v.content[i].empty() || v.tag[i] != tag) continue;
    if (!host) return i;
    if (v.content[i].size() == bytes && memcmp(v.content[i].data(), host, bytes) == 0) return i;
  }
  return -1;
}

smmc_status versioned_upload(Context::Versioned &v, uint64_t tag, const void *host, size_t bytes, cudaStream_t st, int *slot_out) {
  int slot = versioned_find(v, tag, host, bytes);
  if (slot < 0) {
    slot = v.next;
    v.next = (v.next + 1) % Context::kVersions;
    if (!v.copied[slot]) {
      CU(cudaEventCreateWithFlags(&v.copied[slot], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&v.last_read[slot], cudaEventDisableTiming));
    } else {
      // the staging copy of this slot was enqueued kVersions uploads ago: normally long complete (host-side wait only if not)
      CU(cudaEventSynchronize(v.copied[slot]));
    }
    if (v.read_pending[slot]) CU(cudaStreamWaitEvent(st, v.last_read[slot], 0));  // kernels still reading the old version
    if (v.pinned_cap[slot] < bytes) {
      if (v.pinned[slot]) cudaFreeHost(v.pinned[slot]);
      v.pinned[slot] = nullptr;
      v.pinned_cap[slot] = 0;
      CU(cudaMallocHost(&v.pinned[slot], bytes));
      v.pinned_cap[slot] = bytes;
    }
    CU(v.dev[slot].reserve(bytes));
    memcpy(v.pinned[slot], host, bytes);
    CU(cudaMemcpyAsync(v.dev[slot].p, v.pinned[slot], bytes, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(v.copied[slot], st));
    v.content[slot].assign((const unsigned char *)host, (const unsigned char *)host + bytes);
    v.tag[slot] = tag;
    v.read_pending[slot] = false;
  }
  *slot_out = slot;
  return SMMC_OK;
}

// call after the kernel that reads slot `slot` has been launched on `st`
smmc_status versioned_mark_read(Context::Versioned &v, int slot, cudaStream_t st) {
  if (slot < 0) return SMMC_OK;
  CU(cudaEventRecord(v.last_read[slot], st));
  v.read_pending[slot] = true;
  return SMMC_OK;
}

void versioned_release(Context::Versioned &v) {
  for (int i = 0; i < Context::kVersions; i++) {
    v.dev[i].release();
    if (v.pinned[i]) cudaFreeHost(v.pinned[i]);
    v.pinned[i] = nullptr;
    v.pinned_cap[i] = 0;
    if (v.copied[i]) cudaEventDestroy(v.copied[i]);
    if (v.last_read[i]) cudaEventDestroy(v.last_read[i]);
    v.copied[i] = v.last_read[i] = nullptr;
    v.read_pending[i] = false;
    v.content[i].clear();
    v.tag[i] = 0;
  }
  v.next = 0;
}

// The per-launch Sobol tables for (n_periods, seed): [P_pad][8] low direction numbers, then [98][P_pad] (KParams::sobol_hi).  Cached by (n_periods, seed); nothing synchronises.
smmc_status prepare_sobol(Context *ctx, int n_periods, uint64_t seed, cudaStream_t st, smmc::KParams *p, int *slot_out) {
  const int ppad = (n_periods + 31) & ~31;
  const uint32_t *V = smmc::sobol_table();
  *slot_out = -1;
  if (n_periods > smmc::sobol_table_dims()) return fail(SMMC_ERR_INVALID_ARG, "Sobol modes support at most %d periods", smmc::sobol_table_dims());
  p->sobol_ppad = ppad;
  p->sobol_lo_bytes = (uint32_t)ppad * 32u;
  if (ppad == 0) return SMMC_OK;
  const uint64_t tag = (seed * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)n_periods << 1) ^ 1ull;
  // The tag only narrows the search (it can collide for some (seed, n_periods) pairs): a cached version is reused only when
  // the exact key stored behind its tables -- n_periods and both halves of the seed -- is the one asked for.
  const uint32_t key[3] = {(uint32_t)n_periods, (uint32_t)seed, (uint32_t)(seed >> 32)};
  int slot = -1;
  for (int i = 0; i < Context::kVersions && slot < 0; i++) {
    const std::vector<unsigned char> &c = ctx->sobol_v.content[i];
    if (c.size() >= sizeof key && ctx->sobol_v.tag[i] == tag && memcmp(c.data() + c.size() - sizeof key, key, sizeof key) == 0) slot = i;
  }
  if (slot < 0) {
    std::vector<uint32_t> buf((size_t)ppad * 8 + (size_t)smmc::kSobolHiRows * ppad + 3, 0u);
    uint32_t *lo = buf.data(), *hi = buf.data() + (size_t)ppad * 8;
    for (int m = 0; m < n_periods; m++) {
      const uint32_t *Vm = V + (size_t)m * 32;
      for (int k = 0; k < 8; k++) lo[(size_t)m * 8 + k] = Vm[k];
      const uint32_t sh = sobol_shift_host(m, seed);
      hi[(size_t)0 * ppad + m] = sh;
      hi[(size_t)1 * ppad + m] = sh ^ Vm[7];
      for (int g = 0; g < 6; g++)
        for (int c = 0; c < 16; c++) {
          uint32_t x = 0;
          for (int k = 0; k < 4; k++)
            if ((c >> k) & 1) x ^= Vm[8 + 4 * g + k];
          hi[(size_t)(2 + 16 * g + c) * ppad + m] = x;
        }
    }
    // the exact (n_periods, seed) rides behind the tables: compared above on lookup, and by versioned_upload (whole content)
    memcpy(&buf[buf.size() - 3], key, sizeof key);
    smmc_status us = versioned_upload(ctx->sobol_v, tag, buf.data(), buf.size() * 4, st, &slot);
    if (us != SMMC_OK) return us;
  }
  const uint32_t *dev = (const uint32_t *)ctx->sobol_v.dev[slot].p;
  p->sobol_lo = dev;
  p->sobol_hi = dev + (size_t)ppad * 8;
  *slot_out = slot;
  return SMMC_OK;
}

// Full [dims][32] direction-number table + per-month digital shift on the device (plain Gray-code evaluation:
// debug_draws, trajectories).  Uses ctx->stream; synchronises it.
smmc_status prepare_sobol_full(Context *ctx, int n_periods, uint64_t seed, const uint32_t **full_dev,
                               const uint32_t **shift_dev) {
  const int dims = smmc::sobol_table_dims();
  if (n_periods > dims) return fail(SMMC_ERR_INVALID_ARG, "Sobol modes support at most %d periods", dims);
  if (!ctx->sobol_full_ready) {
    CU(ctx->sobol_full.reserve((size_t)dims * 32 * 4));
    CU(cudaMemcpyAsync(ctx->sobol_full.p, smmc::sobol_table(), (size_t)dims * 32 * 4, cudaMemcpyHostToDevice, ctx->stream));
    ctx->sobol_full_ready = true;
  }
  std::vector<uint32_t> shift((size_t)(n_periods > 0 ? n_periods : 1), 0u);
  for (int m = 0; m < n_periods; m++) shift[m] = sobol_shift_host(m, seed);
  CU(ctx->sobol_shift.reserve(shift.size() * 4));
  CU(cudaMemcpyAsync(ctx->sobol_shift.p, shift.data(), shift.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *full_dev = (const uint32_t *)ctx->sobol_full.p;
  *shift_dev = (const uint32_t *)ctx->sobol_shift.p;
  return SMMC_OK;
}

// Per-period keep-fractions q[m] = 1 - pct_m/100 on the device (withdrawal variant of the kernels); null when the config
// asks for no withdrawals.  Cached by content; nothing synchronises.
smmc_status prepare_flow(Context *ctx, const smmc_config *cfg, cudaStream_t st, smmc::KParams *p, int *slot_out) {
  p->flow_q = nullptr;
  p->flow_w = 0.0f;
  *slot_out = -1;
  if (!has_withdrawals(cfg)) return SMMC_OK;
  const int P = cfg->n_periods;
  std::vector<float> q((size_t)(P > 0 ? P : 1), 1.0f);
  for (int m = 0; m < P; m++) {
    const float pct = cfg->withdraw_pct_schedule ? cfg->withdraw_pct_schedule[m] : cfg->withdraw_pct;
    if (!std::isfinite(pct)) return fail(SMMC_ERR_INVALID_ARG, "withdraw_pct_schedule[%d] is not finite", m);
    q[(size_t)m] = 1.0f - pct * 0.01f;
  }
  int slot = -1;
  smmc_status us = versioned_upload(ctx->flow_v, 1, q.data(), q.size() * 4, st, &slot);
  if (us != SMMC_OK) return us;
  p->flow_q = (const float *)ctx->flow_v.dev[slot].p;
  p->flow_w = cfg->withdraw_fixed;
  *slot_out = slot;
  return SMMC_OK;
}

// one radix-select pass over the finals' ordered keys (smmc_select_pass / smmc_quantiles)
struct SelectPass {
  int targets = 0;       // histograms kept by the kernel (1 in the first pass)
  int digit_shift = 0;
  uint32_t digit_mask = 0;
  int prefix_shift = 32;
  uint32_t prefix[SMMC_SELECT_MAX_TARGETS] = {};
  unsigned long long *hist_dev = nullptr;  // caller's device buffer (DEVICE_POINTERS) or null = library workspace
  // candidate collection riding along with this replay (second pass): see KParams
  int cand_groups = 0;
  uint32_t cand_prefix[SMMC_SELECT_MAX_TARGETS] = {};
  uint32_t *cand_keys = nullptr;
  unsigned long long *cand_cursor = nullptr;
  unsigned long long cand_capacity = 0;
};

// digit width of the pass that starts with `bits_done` key bits known: 11, 11, 10
int select_width(int bits_done) { return bits_done >= 22 ? 10 : 11; }

smmc_status make_select_pass(const smmc_select_state *st, SelectPass *sp) {
  if (!st || st->n_targets < 1 || st->n_targets > SMMC_SELECT_MAX_TARGETS)
    return fail(SMMC_ERR_INVALID_ARG, "select: n_targets must be in [1,%d]", SMMC_SELECT_MAX_TARGETS);
  if (st->bits_done != 0 && st->bits_done != 11 && st->bits_done != 22)
    return fail(SMMC_ERR_INVALID_ARG, "select: no pass left (bits_done = %d)", st->bits_done);
  const int w = select_width(st->bits_done);
  sp->targets = st->bits_done == 0 ? 1 : st->n_targets;
  sp->digit_shift = 32 - st->bits_done - w;
  sp->digit_mask = (1u << w) - 1u;
  sp->prefix_shift = 32 - st->bits_done;
  for (int t = 0; t < SMMC_SELECT_MAX_TARGETS; t++) sp->prefix[t] = t < st->n_targets && st->bits_done ? st->prefix[t] : 0u;
  return SMMC_OK;
}

// everything a launched shard needs to be collected later
struct Pending {
  Context *ctx = nullptr;
  cudaStream_t stream = nullptr;
  int64_t n = 0;
  int64_t nb = 0;
  int stats_len = 0;
  bool device_ptrs = false;
  float *finals_dev = nullptr, *bmeans_dev = nullptr, *bvars_dev = nullptr;
  double *stats_dev = nullptr;
  unsigned long long *sel_dev = nullptr;
  int sel_words = 0;
  int launches = 0;
  int ev_slot = 0;
};

// Cross-stream ordering of the per-device cached workspace (ctx->work / tbl / returns / sorted / stats / scratch ...):
// before a call touches those buffers on `st`, it waits (stream-side, the host does not block) for the last call that used
// them on a different stream; after its last enqueue it records the event again.
smmc_status order_after_last_use(Context *ctx, cudaStream_t st) {
  if (ctx->has_last_use && ctx->last_stream != st) CU(cudaStreamWaitEvent(st, ctx->last_use, 0));
  return SMMC_OK;
}
smmc_status mark_last_use(Context *ctx, cudaStream_t st) {
  CU(cudaEventRecord(ctx->last_use, st));
  ctx->last_stream = st;
  ctx->has_last_use = true;
  return SMMC_OK;
}

// the (cos, sin) table of the Gaussian modes, built once per device
const float *sincos_table_of(const Context *ctx, int mode) {
  const char *base = (const char *)ctx->sincos.p;
  return (const float *)(mode == SMMC_MODE_GAUSSIAN ? base + smmc::kSinCosBytes : base);
}

// The table holds sqrt(-k) * (cos, sin): it belongs to a value of gauss_std and is rebuilt (one 3 us kernel on `st`, ordered behind
// the last use of the workspace like every other write to it) when that changes.  *rebuilt: a kernel was launched.
smmc_status prepare_sincos(Context *ctx, float k, cudaStream_t st, bool *rebuilt = nullptr) {
  if (rebuilt) *rebuilt = false;
  uint32_t kbits;
  memcpy(&kbits, &k, sizeof kbits);
  if (ctx->sincos_ready && ctx->sincos_kbits == kbits) return SMMC_OK;
  CU(ctx->sincos.reserve(2 * (size_t)smmc::kSinCosBytes));  // [swizzled copy (SOBOL_GAUSSIAN)][linear copy (GAUSSIAN)]
  smmc::launch_sincos_table((float *)ctx->sincos.p, std::sqrt(-(double)k), st);
  CU(cudaGetLastError());
  ctx->sincos_ready = true;
  ctx->sincos_kbits = kbits;
  if (rebuilt) *rebuilt = true;
  return SMMC_OK;
}

// where the launch's outputs go (device pointers); null = library workspace / not requested
struct DeviceOutputs {
  float *finals = nullptr, *bmeans = nullptr, *bvars = nullptr;
  double *stats = nullptr;
  bool want_finals = false, want_bmeans = false, want_bvars = false;  // host-pointer runs: allocate in the context cache
};

// how a launch deviates from "one call = one launch with the caller's buffers"
struct EnqueueOpts {
  const DeviceOutputs *dout = nullptr;  // explicit device outputs (chunked / pipelined runs)
  bool reuse_table = false;             // the scaled table of the previous launch on this context is still valid: skip upload + prep
};

// Enqueue one shard [cfg->stream_offset, +N) on cfg->device.  Caller holds ctx->mu.
smmc_status enqueue(const smmc_config *cfg, const smmc_result *out, Context *ctx, Pending *pd,
                    const SelectPass *sel = nullptr, const EnqueueOpts &opts = EnqueueOpts()) {
  CU(cudaSetDevice(ctx->device));
  const bool devp = (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) != 0;
  const DeviceOutputs *dout = opts.dout;
  cudaStream_t st = devp ? (cudaStream_t)cfg->stream : ctx->stream;
  const int64_t N = cfg->n_simulations;
  const int64_t nb = (N + SMMC_BLOCK - 1) / SMMC_BLOCK;
  const int bins = cfg->hist_bins;
  const int slen = smmc_stats_len(bins);
  const smmc::Variant variant = current_variant();
  const bool sincos = uses_sincos_table(cfg->mode, variant, cfg->philox_rounds);

  smmc::LaunchInfo li;
  if (!smmc::plan_launch(cfg->mode, cfg->n_returns, cfg->n_periods, bins, sel ? sel->targets : 0, has_withdrawals(cfg), sincos, N,
                         ctx->n_sms, ctx->smem_optin, variant.warps, variant.ticket, &li))
    return fail(SMMC_ERR_UNSUPPORTED, "returns table of %d entries (+ histograms / schedules) does not fit in %zu bytes of shared memory",
                cfg->n_returns, ctx->smem_optin);

  {
    smmc_status os = order_after_last_use(ctx, st);
    if (os != SMMC_OK) return os;
  }

  smmc::KParams p;
  memset(&p, 0, sizeof p);
  fill_stream_params(cfg, &p);
  p.n_paths = N;
  p.target = cfg->target;
  p.hist_lo = cfg->hist_lo;
  p.hist_inv_w = bins > 0 ? (float)((double)bins / ((double)cfg->hist_hi - (double)cfg->hist_lo)) : 0.0f;
  p.hist_bins = bins;
  p.rep4 = 4u;
  pd->launches = 0;

  // ---- table: upload (host pointers) and expand to the scaled, bank-private replicated layout
  if (is_bootstrap(cfg->mode)) {
    const int n = cfg->n_returns;
    const uint32_t tbl_bytes = (uint32_t)((((size_t)(n + 1) * li.rep * 4) + 15) & ~(size_t)15);
    if (!opts.reuse_table) {
      const float *src = cfg->returns;
      if (!devp) {
        CU(ctx->returns.reserve((size_t)n * 4));
        CU(cudaMemcpyAsync(ctx->returns.p, cfg->returns, (size_t)n * 4, cudaMemcpyHostToDevice, st));
        src = (const float *)ctx->returns.p;
      }
      if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP) {  // quasi-random draws sample the SORTED table (quantile function)
        CU(ctx->sorted.reserve((size_t)n * 4));
        smmc::launch_sort_table(src, n, (float *)ctx->sorted.p, st);
        CU(cudaGetLastError());
        pd->launches++;
        src = (const float *)ctx->sorted.p;
      }
      CU(ctx->tbl.reserve(tbl_bytes));
      smmc::launch_prep_table(src, n, li.rep, (float *)ctx->tbl.p, tbl_bytes / 4, st);
      CU(cudaGetLastError());
      pd->launches++;
    }
    p.tbl_rep = (const float *)ctx->tbl.p;
    p.tbl_bytes = tbl_bytes;
    p.rep4 = (uint32_t)li.rep * 4u;
    p.bc.row_shift = 2;
    while ((1u << p.bc.row_shift) < (uint32_t)li.rep * 4u) p.bc.row_shift++;
  } else if (sincos) {
    bool fresh = false;
    smmc_status ss = prepare_sincos(ctx, p.gc.k, st, &fresh);
    if (ss != SMMC_OK) return ss;
    if (fresh) pd->launches++;
    p.tbl_rep = sincos_table_of(ctx, cfg->mode);
    p.tbl_bytes = smmc::kSinCosBytes;
  }

  // ---- workspace: [cta_partials grid*4 doubles][hist_g (bins+2) u64][ticket u32 (+pad)]
  const size_t part_bytes = (size_t)li.grid * 4 * sizeof(double);
  const size_t hist_bytes = (size_t)(bins + 2) * sizeof(unsigned long long);
  CU(ctx->work.reserve(part_bytes + hist_bytes + 16));
  char *w = (char *)ctx->work.p;
  p.cta_partials = (double *)w;
  p.hist_g = (unsigned long long *)(w + part_bytes);
  p.ticket = (unsigned int *)(w + part_bytes + hist_bytes);
  CU(cudaMemsetAsync(w + part_bytes, 0, hist_bytes + 16, st));
  if (li.ticket && nb > 0) {  // blocks handed out by ticket, their moments merged in block order (smmc_kernels.cu, plan_ticket)
    CU(ctx->block_moments.reserve((size_t)nb * sizeof(float2)));
    p.block_moments = (float2 *)ctx->block_moments.p;
    p.block_ticket = p.ticket + 1;  // inside the 16 bytes zeroed above
  }

  // ---- outputs
  if (dout) {
    p.finals = dout->finals;
    p.bmeans = dout->bmeans;
    p.bvars = dout->bvars;
    p.stats = dout->stats;
  } else if (devp) {
    p.finals = out->finals;
    p.bmeans = out->block_means;
    p.bvars = out->block_vars;
    p.stats = out->stats_dev;
  } else {
    if (out->finals && N > 0) {
      CU(ctx->finals.reserve((size_t)N * 4));
      p.finals = (float *)ctx->finals.p;
    }
    if (out->block_means && nb > 0) {
      CU(ctx->bmeans.reserve((size_t)nb * 4));
      p.bmeans = (float *)ctx->bmeans.p;
    }
    if (out->block_vars && nb > 0) {
      CU(ctx->bvars.reserve((size_t)nb * 4));
      p.bvars = (float *)ctx->bvars.p;
    }
  }
  if (!p.stats) {
    CU(ctx->stats.reserve((size_t)slen * sizeof(double)));
    p.stats = (double *)ctx->stats.p;
  }

  int sobol_slot = -1, flow_slot = -1;
  if (is_sobol(cfg->mode)) {
    smmc_status sst = prepare_sobol(ctx, cfg->n_periods, cfg->seed, st, &p, &sobol_slot);
    if (sst != SMMC_OK) return sst;
  }
  {
    smmc_status fst = prepare_flow(ctx, cfg, st, &p, &flow_slot);
    if (fst != SMMC_OK) return fst;
  }

  pd->sel_dev = nullptr;
  pd->sel_words = 0;
  if (sel && sel->targets > 0) {
    const size_t sel_bytes = (size_t)sel->targets * SMMC_SELECT_BINS * sizeof(unsigned long long);
    unsigned long long *hd = sel->hist_dev;
    if (!hd) {
      CU(ctx->selhist.reserve(sel_bytes));
      hd = (unsigned long long *)ctx->selhist.p;
    }
    CU(cudaMemsetAsync(hd, 0, sel_bytes, st));
    p.sel_targets = sel->targets;
    p.sel_digit_shift = sel->digit_shift;
    p.sel_digit_mask = sel->digit_mask;
    p.sel_prefix_shift = sel->prefix_shift;
    for (int t = 0; t < SMMC_SELECT_MAX_TARGETS; t++) p.sel_prefix[t] = sel->prefix[t];
    p.sel_hist_g = hd;
    p.cand_groups = sel->cand_groups;
    for (int t = 0; t < SMMC_SELECT_MAX_TARGETS; t++) p.cand_prefix[t] = sel->cand_prefix[t];
    p.cand_keys = sel->cand_keys;
    p.cand_cursor = sel->cand_cursor;
    p.cand_capacity = sel->cand_capacity;
    if (sel->cand_groups > 0) CU(cudaMemsetAsync(sel->cand_cursor, 0, sizeof(unsigned long long), st));
    pd->sel_dev = hd;
    pd->sel_words = sel->targets * SMMC_SELECT_BINS;
  }

  const int slot = (int)(ctx->ev_head % kEventRing);
  CU(cudaEventRecord(ctx->ev0[slot], st));
  CU(smmc::launch_mc(cfg->mode, p, li, variant, st));
  CU(cudaEventRecord(ctx->ev1[slot], st));
  ctx->ev_kind[slot] = SMMC_KERNEL_SIMULATION;
  ctx->ev_head++;
  pd->ev_slot = slot;
  pd->launches++;
  if (sobol_slot >= 0) {
    smmc_status ms = versioned_mark_read(ctx->sobol_v, sobol_slot, st);
    if (ms != SMMC_OK) return ms;
  }
  if (flow_slot >= 0) {
    smmc_status ms = versioned_mark_read(ctx->flow_v, flow_slot, st);
    if (ms != SMMC_OK) return ms;
  }
  {
    smmc_status ms = mark_last_use(ctx, st);
    if (ms != SMMC_OK) return ms;
  }

  pd->ctx = ctx;
  pd->stream = st;
  pd->n = N;
  pd->nb = nb;
  pd->stats_len = slen;
  pd->device_ptrs = devp;
  pd->finals_dev = p.finals;
  pd->bmeans_dev = p.bmeans;
  pd->bvars_dev = p.bvars;
  pd->stats_dev = p.stats;
  return SMMC_OK;
}

// Download results of a shard into `out` (host pointers) and its statistics vector into stats_host.
// First touch of a large caller-owned OUTPUT buffer, in parallel, while the kernel that will fill it is still running:
// a fresh 400 MB vector (10^8 finals) otherwise takes its 100 000 page faults one by one inside the D2H copy.  The
// buffer is about to be overwritten, so writing a zero per page is harmless.
void prefault_output(void *ptr, size_t bytes) {
  constexpr size_t kPage = 4096, kMin = 32u << 20;
  if (!ptr || bytes < kMin) return;
  unsigned hw = std::thread::hardware_concurrency();
  const unsigned nt = hw >= 16 ? 8u : (hw >= 4 ? hw / 2 : 1u);
  char *base = static_cast<char *>(ptr);
  const size_t first = (kPage - (reinterpret_cast<uintptr_t>(base) & (kPage - 1))) & (kPage - 1);
  const size_t pages = bytes > first ? (bytes - first + kPage - 1) / kPage : 0;
  base[0] = 0;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; t++)
    th.emplace_back([=] {
      for (size_t pg = pages * t / nt; pg < pages * (t + 1) / nt; pg++) {
        const size_t off = first + pg * kPage;
        if (off < bytes) *reinterpret_cast<volatile char *>(base + off) = 0;
      }
    });
  for (auto &x : th) x.join();
}

// Issue the downloads of a shard's results into `out` (host pointers) and of its statistics vector (pinned).
smmc_status collect_issue(const Pending &pd, const smmc_result *out) {
  Context *ctx = pd.ctx;
  CU(cudaSetDevice(ctx->device));
  if (ctx->stats_host_cap < (size_t)pd.stats_len) {
    if (ctx->stats_host) cudaFreeHost(ctx->stats_host);
    ctx->stats_host = nullptr;
    CU(cudaMallocHost(&ctx->stats_host, (size_t)pd.stats_len * sizeof(double)));
    ctx->stats_host_cap = (size_t)pd.stats_len;
  }
  CU(cudaMemcpyAsync(ctx->stats_host, pd.stats_dev, (size_t)pd.stats_len * sizeof(double), cudaMemcpyDeviceToHost, pd.stream));
  if (!pd.device_ptrs) {
    if (out->finals && pd.n > 0) prefault_output(out->finals, (size_t)pd.n * 4);  // overlaps the running kernel
    if (out->finals && pd.n > 0)
      CU(cudaMemcpyAsync(out->finals, pd.finals_dev, (size_t)pd.n * 4, cudaMemcpyDeviceToHost, pd.stream));
    if (out->block_means && pd.nb > 0)
      CU(cudaMemcpyAsync(out->block_means, pd.bmeans_dev, (size_t)pd.nb * 4, cudaMemcpyDeviceToHost, pd.stream));
    if (out->block_vars && pd.nb > 0)
      CU(cudaMemcpyAsync(out->block_vars, pd.bvars_dev, (size_t)pd.nb * 4, cudaMemcpyDeviceToHost, pd.stream));
  }
  return SMMC_OK;
}

smmc_status collect_wait(const Pending &pd, double *stats_host, float *device_ms) {
  Context *ctx = pd.ctx;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(pd.stream));
  memcpy(stats_host, ctx->stats_host, (size_t)pd.stats_len * sizeof(double));
  float ms = 0.0f;
  CU(cudaEventElapsedTime(&ms, ctx->ev0[pd.ev_slot], ctx->ev1[pd.ev_slot]));
  *device_ms = ms;
  return SMMC_OK;
}

smmc_status collect(const Pending &pd, const smmc_result *out, double *stats_host, float *device_ms) {
  smmc_status st = collect_issue(pd, out);
  if (st != SMMC_OK) return st;
  return collect_wait(pd, stats_host, device_ms);
}

// ---- shards: one contiguous, 256-aligned path range of a host-pointer run on one device ---------------------------------
struct Shard {
  Context *ctx = nullptr;
  smmc_config cfg;              // stream_offset / n_simulations / device of this shard
  float *finals = nullptr;      // host destinations (nullable)
  float *bmeans = nullptr, *bvars = nullptr;
};

// Splits [0, N) of cfg over the given devices in whole 256-path blocks (the partition smmc_run_multi documents).
// Whole 256-path blocks [b0, b1) of shard `rank` of `n`: floor(nb / n) blocks each, the remainder one apiece to the lowest
// ranks -- balanced to one block (the ONE partition rule: smmc_run_multi, smmc_run_rank, smmc_rank_range, distributed.py).
void shard_blocks(int64_t N, int n, int rank, int64_t *b0, int64_t *b1) {
  const int64_t nb = (N + SMMC_BLOCK - 1) / SMMC_BLOCK;
  const int64_t base = nb / n, extra = nb % n;
  *b0 = base * rank + (rank < extra ? rank : extra);
  *b1 = *b0 + base + (rank < extra ? 1 : 0);
}

void make_shards(const smmc_config *cfg, const smmc_result *out, const std::vector<Context *> &ctxs, std::vector<Shard> *shards) {
  const int n_gpus = (int)ctxs.size();
  const int64_t N = cfg->n_simulations;
  shards->assign((size_t)n_gpus, Shard());
  for (int g = 0; g < n_gpus; g++) {
    int64_t b0 = 0, b1 = 0;
    shard_blocks(N, n_gpus, g, &b0, &b1);
    int64_t p0 = b0 * SMMC_BLOCK, p1 = b1 * SMMC_BLOCK;
    if (p1 > N) p1 = N;
    if (p0 >= p1) p0 = p1 = 0;  // an empty shard keeps a block-aligned offset
    Shard &sh = (*shards)[(size_t)g];
    sh.ctx = ctxs[(size_t)g];
    sh.cfg = *cfg;
    sh.cfg.device = ctxs[(size_t)g]->device;
    sh.cfg.stream_offset = cfg->stream_offset + (uint64_t)p0;
    sh.cfg.n_simulations = p1 - p0;
    sh.finals = out->finals ? out->finals + p0 : nullptr;
    sh.bmeans = out->block_means ? out->block_means + b0 : nullptr;
    sh.bvars = out->block_vars ? out->block_vars + b0 : nullptr;
  }
}

// ---- the all-finals path, pipelined ------------------------------------------------------------------------------------
// The reference's primary entry point (mc_simulations_gpu, src/simulations.cu:576-655) returns every final value: 4 bytes
// per path over PCIe into pageable caller memory, which costs more than the simulation (10^8 x 360: 14 ms of kernel).
// Each shard is cut into chunks of whole launch waves; the path id is the RNG offset, so a chunk is just a launch with
// another stream_offset and every value is bit-identical to the monolithic launch.  Per device:
//   compute stream: kernel(c) -> stats(c) D2H (864 B, pinned)           [kChunkDepth chunks ahead, device-side ring]
//   copy stream:    wait kernel(c) -> finals(c) D2H into a pinned staging slot             [kStageSlots slots]
//   host workers:   wait copy(c) -> memcpy slot -> caller's buffer (first touch of its pages included), slot freed
// so the kernel of chunk c+1.. overlaps the download of chunk c and the host copy of chunk c-1.
struct Chunk {
  int shard = 0;
  int64_t p0 = 0, n = 0;  // path range relative to the shard
  int index = 0;          // chunk number within the shard
};

int64_t pipeline_chunk_paths(int64_t n_shard, int n_sms, int forced_chunks) {
  const int64_t wave = (int64_t)n_sms * smmc::kWarps * SMMC_BLOCK;  // one full wave of 256-path blocks
  const int want = forced_chunks > 1 ? forced_chunks : 16;
  int64_t c = (n_shard + want - 1) / want;
  c = (c + wave - 1) / wave * wave;                        // whole waves: only the last chunk has a ragged tail
  const int64_t cmax = ((int64_t)16 << 20) / wave * wave;  // <= 64 MB of finals per staging slot
  if (c > cmax && forced_chunks <= 1) c = cmax > 0 ? cmax : wave;
  if (c < wave) c = wave;
  return c;
}

constexpr int64_t kPipelineMinPaths = (int64_t)4 << 20;  // below 16 MB of finals the monolithic path is as fast

// A large, freshly allocated destination buffer is first touched by the download threads: ask for transparent huge pages
// on its 2 MB-aligned interior (a hint: no effect where THP is off), 512 times fewer page faults than 4 KB pages.
void hint_huge_pages(void *ptr, size_t bytes) {
#ifdef __linux__
  constexpr uintptr_t kHuge = (uintptr_t)2 << 20;
  if (!ptr || bytes < 4 * kHuge) return;
  const uintptr_t a = (reinterpret_cast<uintptr_t>(ptr) + kHuge - 1) & ~(kHuge - 1);
  const uintptr_t b = (reinterpret_cast<uintptr_t>(ptr) + bytes) & ~(kHuge - 1);
  if (b > a) madvise(reinterpret_cast<void *>(a), b - a, MADV_HUGEPAGE);
#else
  (void)ptr;
  (void)bytes;
#endif
}

// Runs `shards` (host pointers, finals requested) through the chunk pipeline.  total: summed statistics vector (shard, chunk
// order: deterministic).  Caller holds every shard's ctx->mu.
smmc_status run_finals_pipelined(std::vector<Shard> &shards, const smmc_config *cfg, std::vector<double> *total, float *device_ms,
                                 int *launches) {
  const int slen = smmc_stats_len(cfg->hist_bins);
  const smmc::Variant variant = current_variant();
  const int S = (int)shards.size();
  constexpr int kDepth = Context::kChunkDepth, kSlots = Context::kStageSlots;

  // ---- chunk lists, interleaved over the shards so that every device makes progress
  std::vector<std::vector<Chunk>> per((size_t)S);
  std::vector<int64_t> csize((size_t)S, 0);
  size_t most = 0;
  for (int s = 0; s < S; s++) {
    const int64_t n = shards[(size_t)s].cfg.n_simulations;
    csize[(size_t)s] = pipeline_chunk_paths(n, shards[(size_t)s].ctx->n_sms, variant.finals_chunks);
    for (int64_t p0 = 0; p0 < n; p0 += csize[(size_t)s]) {
      Chunk c;
      c.shard = s;
      c.p0 = p0;
      c.n = std::min(csize[(size_t)s], n - p0);
      c.index = (int)per[(size_t)s].size();
      per[(size_t)s].push_back(c);
    }
    most = std::max(most, per[(size_t)s].size());
  }
  std::vector<Chunk> order;
  for (size_t i = 0; i < most; i++)
    for (int s = 0; s < S; s++)
      if (i < per[(size_t)s].size()) order.push_back(per[(size_t)s][i]);
  const int G = (int)order.size();

  // ---- per-device resources (grow-only, cached in the context)
  std::vector<std::vector<double *>> stats_host((size_t)S);
  for (int s = 0; s < S; s++) {
    Shard &sh = shards[(size_t)s];
    Context *ctx = sh.ctx;
    const size_t nch = per[(size_t)s].size();
    if (nch == 0) continue;
    CU(cudaSetDevice(ctx->device));
    const size_t cbytes = (size_t)csize[(size_t)s] * 4;
    CU(ctx->finals.reserve(cbytes * kDepth));
    if (ctx->stage_cap < cbytes) {
      for (int i = 0; i < kSlots; i++) {
        if (ctx->stage[i]) cudaFreeHost(ctx->stage[i]);
        ctx->stage[i] = nullptr;
      }
      ctx->stage_cap = 0;
      for (int i = 0; i < kSlots; i++) CU(cudaMallocHost(&ctx->stage[i], cbytes));
      ctx->stage_cap = cbytes;
    }
    CU(ctx->chunk_stats.reserve((size_t)kDepth * slen * sizeof(double)));
    const size_t need = nch * (size_t)slen;
    if (ctx->chunk_stats_host_cap < need) {
      if (ctx->chunk_stats_host) cudaFreeHost(ctx->chunk_stats_host);
      ctx->chunk_stats_host = nullptr;
      ctx->chunk_stats_host_cap = 0;
      CU(cudaMallocHost(&ctx->chunk_stats_host, need * sizeof(double)));
      ctx->chunk_stats_host_cap = need;
    }
    const int64_t nb = (sh.cfg.n_simulations + SMMC_BLOCK - 1) / SMMC_BLOCK;
    if (sh.bmeans) CU(ctx->bmeans.reserve((size_t)nb * 4));
    if (sh.bvars) CU(ctx->bvars.reserve((size_t)nb * 4));
  }

  // ---- host workers: staging slot -> caller's buffer
  unsigned hw = std::thread::hardware_concurrency();
  int T = hw >= 16 ? 8 : (hw >= 4 ? (int)hw / 2 : 1);
  if (variant.finals_threads > 0) T = variant.finals_threads;
  for (const Shard &sh : shards) hint_huge_pages(sh.finals, (size_t)sh.cfg.n_simulations * 4);
  std::vector<std::atomic<int>> issued((size_t)G), out_done((size_t)G);
  for (int g = 0; g < G; g++) {
    issued[(size_t)g].store(0);
    out_done[(size_t)g].store(0);
  }
  std::vector<cudaEvent_t> copied((size_t)G, nullptr);
  std::atomic<int> abort_flag(0);
  auto worker = [&](int t) {
    for (int g = 0; g < G; g++) {
      while (!issued[(size_t)g].load(std::memory_order_acquire)) {
        if (abort_flag.load(std::memory_order_relaxed)) return;
        std::this_thread::yield();
      }
      if (cudaEventSynchronize(copied[(size_t)g]) != cudaSuccess) {
        abort_flag.store(1);
        return;
      }
      const Chunk &c = order[(size_t)g];
      const Shard &sh = shards[(size_t)c.shard];
      const size_t bytes = (size_t)c.n * 4, lo = bytes * (size_t)t / (size_t)T & ~(size_t)63, hi = (t == T - 1) ? bytes : (bytes * (size_t)(t + 1) / (size_t)T & ~(size_t)63);
      if (hi > lo)
        memcpy(reinterpret_cast<char *>(sh.finals + c.p0) + lo, static_cast<const char *>(sh.ctx->stage[c.index % kSlots]) + lo, hi - lo);
      out_done[(size_t)g].fetch_add(1, std::memory_order_release);
    }
  };
  std::vector<std::thread> threads;
  for (int t = 0; t < T; t++) threads.emplace_back(worker, t);
  auto stop_workers = [&]() {
    abort_flag.store(1);
    for (auto &th : threads) th.join();
  };

  // ---- issue loop
  std::vector<int> next_kernel((size_t)S, 0);
  std::vector<int> gidx_of((size_t)S * most + 1, -1);  // (shard, chunk index) -> position in `order`
  for (int g = 0; g < G; g++) gidx_of[(size_t)order[(size_t)g].shard * most + (size_t)order[(size_t)g].index] = g;
  smmc_status rc = SMMC_OK;
  *launches = 0;
  auto enqueue_kernel = [&](int s, int i) -> smmc_status {
    Shard &sh = shards[(size_t)s];
    Context *ctx = sh.ctx;
    const Chunk &c = per[(size_t)s][(size_t)i];
    const int slot = i % kDepth;
    CU(cudaSetDevice(ctx->device));
    if (i == 0) CU(cudaEventRecord(ctx->red0, ctx->stream));
    if (i >= kDepth) CU(cudaStreamWaitEvent(ctx->stream, ctx->cdone[slot], 0));  // ring slot: its previous chunk has been downloaded
    smmc_config cc = sh.cfg;
    cc.stream_offset = sh.cfg.stream_offset + (uint64_t)c.p0;
    cc.n_simulations = c.n;
    DeviceOutputs d;
    d.finals = (float *)ctx->finals.p + (size_t)slot * (size_t)csize[(size_t)s];
    d.bmeans = sh.bmeans ? (float *)ctx->bmeans.p + c.p0 / SMMC_BLOCK : nullptr;
    d.bvars = sh.bvars ? (float *)ctx->bvars.p + c.p0 / SMMC_BLOCK : nullptr;
    d.stats = (double *)ctx->chunk_stats.p + (size_t)slot * slen;
    EnqueueOpts o;
    o.dout = &d;
    o.reuse_table = i > 0;
    smmc_result dummy;
    memset(&dummy, 0, sizeof dummy);
    Pending pd;
    smmc_status st = enqueue(&cc, &dummy, ctx, &pd, nullptr, o);
    if (st != SMMC_OK) return st;
    *launches += pd.launches;
    CU(cudaEventRecord(ctx->kdone[slot], ctx->stream));
    CU(cudaMemcpyAsync(ctx->chunk_stats_host + (size_t)i * slen, d.stats, (size_t)slen * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (i + 1 == (int)per[(size_t)s].size()) CU(cudaEventRecord(ctx->red1, ctx->stream));
    return SMMC_OK;
  };
  for (int g = 0; g < G && rc == SMMC_OK; g++) {
    const Chunk &c = order[(size_t)g];
    const int s = c.shard, i = c.index;
    Context *ctx = shards[(size_t)s].ctx;
    const int nch = (int)per[(size_t)s].size();
    while (rc == SMMC_OK && next_kernel[(size_t)s] < nch && next_kernel[(size_t)s] < i + kDepth) {
      rc = enqueue_kernel(s, next_kernel[(size_t)s]);
      next_kernel[(size_t)s]++;
    }
    if (rc != SMMC_OK) break;
    if (i >= kSlots) {  // the staging slot is free once every worker has copied chunk i - kSlots out of it
      const int gp = gidx_of[(size_t)s * most + (size_t)(i - kSlots)];
      while (out_done[(size_t)gp].load(std::memory_order_acquire) < T) {
        if (abort_flag.load()) break;
        std::this_thread::yield();
      }
    }
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream, ctx->kdone[i % kDepth], 0);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(ctx->stage[i % kSlots], (float *)ctx->finals.p + (size_t)(i % kDepth) * (size_t)csize[(size_t)s], (size_t)c.n * 4,
                          cudaMemcpyDeviceToHost, ctx->copy_stream);
    if (e == cudaSuccess) e = cudaEventRecord(ctx->cdone[i % kDepth], ctx->copy_stream);
    if (e != cudaSuccess || abort_flag.load()) {
      rc = fail(SMMC_ERR_CUDA, "finals pipeline: %s", cudaGetErrorString(e));
      break;
    }
    copied[(size_t)g] = ctx->cdone[i % kDepth];
    issued[(size_t)g].store(1, std::memory_order_release);
  }
  if (rc != SMMC_OK) {
    stop_workers();
    for (int s = 0; s < S; s++) {
      cudaSetDevice(shards[(size_t)s].ctx->device);
      cudaStreamSynchronize(shards[(size_t)s].ctx->stream);
      cudaStreamSynchronize(shards[(size_t)s].ctx->copy_stream);
    }
    return rc;
  }
  for (auto &th : threads) th.join();
  if (abort_flag.load()) return fail(SMMC_ERR_CUDA, "finals pipeline: a download failed");

  // ---- block partials, statistics, timing
  total->assign((size_t)slen, 0.0);
  *device_ms = 0.0f;
  for (int s = 0; s < S; s++) {
    Shard &sh = shards[(size_t)s];
    Context *ctx = sh.ctx;
    const int nch = (int)per[(size_t)s].size();
    if (nch == 0) continue;
    CU(cudaSetDevice(ctx->device));
    const int64_t nb = (sh.cfg.n_simulations + SMMC_BLOCK - 1) / SMMC_BLOCK;
    if (sh.bmeans) CU(cudaMemcpyAsync(sh.bmeans, ctx->bmeans.p, (size_t)nb * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (sh.bvars) CU(cudaMemcpyAsync(sh.bvars, ctx->bvars.p, (size_t)nb * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaStreamSynchronize(ctx->copy_stream));
    for (int i = 0; i < nch; i++)
      for (int k = 0; k < slen; k++) (*total)[(size_t)k] += ctx->chunk_stats_host[(size_t)i * slen + k];
    float ms = 0.0f;
    CU(cudaEventElapsedTime(&ms, ctx->red0, ctx->red1));
    if (ms > *device_ms) *device_ms = ms;
  }
  return SMMC_OK;
}

// ---- block partials into host memory: the download of the first part hides behind the second part's kernel -------------
// reduceBlock-style runs (block_means / block_vars, no finals) return 8 bytes per 256 paths: 3.1 MB at 10^8 paths, into the
// caller's pageable -- usually freshly allocated -- arrays: 0.55 ms behind a 17.4 ms kernel when it is one copy after one
// launch (measured, tools/exp_e2e.py).  Each shard is therefore launched in two parts (7/8 of its launch waves, then the
// rest; the path id is the RNG offset, so the values are those of one launch); the first part's partials are downloaded on
// the copy stream while the second part computes -- a pageable cudaMemcpyAsync blocks the host, which has nothing else to
// do -- and only the last eighth is exposed.  Statistics: the two allreduce-ready vectors are added on the host.
constexpr int64_t kSplitMinPaths = (int64_t)16 << 20;

// api / nccl (smmc_run_rank): the first part's vector is added into the second part's on the device (108 doubles, one tiny
// kernel behind the second launch) and that ONE vector is all-reduced over the ranks on the kernel's stream before it is
// downloaded -- one collective per call on this path as on every other; the early download of the first part's partials is kept.
smmc_status run_partials_split(std::vector<Shard> &shards, const smmc_config *cfg, std::vector<double> *total, float *device_ms,
                               int *launches, const smmc::NcclApi *api = nullptr, smmc::NcclComm nccl = nullptr,
                               unsigned long long *allreduces = nullptr) {
  const int slen = smmc_stats_len(cfg->hist_bins);
  const size_t S = shards.size();
  std::vector<int64_t> n_first(S, 0);
  *launches = 0;
  // ---- enqueue both parts of every shard
  for (size_t s = 0; s < S; s++) {
    Shard &sh = shards[s];
    Context *ctx = sh.ctx;
    const int64_t n = sh.cfg.n_simulations;
    if (n <= 0 && !api) continue;  // (with a communicator an empty rank still takes part in the two collectives)
    CU(cudaSetDevice(ctx->device));
    const int64_t wave = (int64_t)ctx->n_sms * smmc::kWarps * SMMC_BLOCK;
    const int64_t waves = n / wave;
    n_first[s] = waves >= 8 ? (waves - (waves + 7) / 8) * wave : 0;  // whole waves; 0 = too small to split
    const int64_t nb = (n + SMMC_BLOCK - 1) / SMMC_BLOCK;
    if (sh.bmeans) CU(ctx->bmeans.reserve((size_t)nb * 4));
    if (sh.bvars) CU(ctx->bvars.reserve((size_t)nb * 4));
    CU(ctx->chunk_stats.reserve((size_t)2 * slen * sizeof(double)));
    if (ctx->chunk_stats_host_cap < (size_t)2 * slen) {
      if (ctx->chunk_stats_host) cudaFreeHost(ctx->chunk_stats_host);
      ctx->chunk_stats_host = nullptr;
      ctx->chunk_stats_host_cap = 0;
      CU(cudaMallocHost(&ctx->chunk_stats_host, (size_t)2 * slen * sizeof(double)));
      ctx->chunk_stats_host_cap = (size_t)2 * slen;
    }
    CU(cudaEventRecord(ctx->red0, ctx->stream));
    for (int part = 0; part < 2; part++) {
      const int64_t p0 = part == 0 ? 0 : n_first[s], cnt = part == 0 ? n_first[s] : n - n_first[s];
      double *stats_dev = (double *)ctx->chunk_stats.p + (size_t)part * slen;
      if (cnt <= 0) {  // unsplit shard: part 0 is empty (an empty rank: both)
        CU(cudaMemsetAsync(stats_dev, 0, (size_t)slen * sizeof(double), ctx->stream));
      } else {
        smmc_config cc = sh.cfg;
        cc.stream_offset = sh.cfg.stream_offset + (uint64_t)p0;
        cc.n_simulations = cnt;
        DeviceOutputs d;
        d.bmeans = sh.bmeans ? (float *)ctx->bmeans.p + p0 / SMMC_BLOCK : nullptr;
        d.bvars = sh.bvars ? (float *)ctx->bvars.p + p0 / SMMC_BLOCK : nullptr;
        d.stats = stats_dev;
        EnqueueOpts o;
        o.dout = &d;
        o.reuse_table = part == 1 && n_first[s] > 0;
        smmc_result dummy;
        memset(&dummy, 0, sizeof dummy);
        Pending pd;
        smmc_status st = enqueue(&cc, &dummy, ctx, &pd, nullptr, o);
        if (st != SMMC_OK) return st;
        *launches += pd.launches;
      }
      if (part == 0) CU(cudaEventRecord(ctx->kdone[0], ctx->stream));
      if (api) {
        if (part == 0) {  // folded into part 1 below: nothing of part 0 is reduced or downloaded by itself
          memset(ctx->chunk_stats_host, 0, (size_t)slen * sizeof(double));
          continue;
        }
        CU(smmc::launch_add_f64(stats_dev, (const double *)ctx->chunk_stats.p, slen, ctx->stream));
        const int r = api->AllReduce(stats_dev, stats_dev, (size_t)slen, smmc::kNcclFloat64, smmc::kNcclSum, nccl, ctx->stream);
        if (r != 0) return fail(SMMC_ERR_CUDA, "ncclAllReduce failed: %s", api->GetErrorString(r));
        if (allreduces) ++*allreduces;
      }
      CU(cudaMemcpyAsync(ctx->chunk_stats_host + (size_t)part * slen, stats_dev, (size_t)slen * sizeof(double), cudaMemcpyDeviceToHost,
                         ctx->stream));
    }
    CU(cudaEventRecord(ctx->red1, ctx->stream));
  }
  // ---- first parts: downloaded while the second parts compute
  for (size_t s = 0; s < S; s++) {
    Shard &sh = shards[s];
    Context *ctx = sh.ctx;
    const size_t nbA = (size_t)(n_first[s] / SMMC_BLOCK);
    if (nbA == 0) continue;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamWaitEvent(ctx->copy_stream, ctx->kdone[0], 0));
    if (sh.bmeans) CU(cudaMemcpyAsync(sh.bmeans, ctx->bmeans.p, nbA * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (sh.bvars) CU(cudaMemcpyAsync(sh.bvars, ctx->bvars.p, nbA * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
  }
  // ---- second parts, statistics, timing
  total->assign((size_t)slen, 0.0);
  *device_ms = 0.0f;
  for (size_t s = 0; s < S; s++) {
    Shard &sh = shards[s];
    Context *ctx = sh.ctx;
    const int64_t n = sh.cfg.n_simulations;
    if (n <= 0 && !api) continue;
    CU(cudaSetDevice(ctx->device));
    const size_t nb = (size_t)((n + SMMC_BLOCK - 1) / SMMC_BLOCK), nbA = (size_t)(n_first[s] / SMMC_BLOCK);
    if (sh.bmeans && nb > nbA)
      CU(cudaMemcpyAsync(sh.bmeans + nbA, (float *)ctx->bmeans.p + nbA, (nb - nbA) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (sh.bvars && nb > nbA)
      CU(cudaMemcpyAsync(sh.bvars + nbA, (float *)ctx->bvars.p + nbA, (nb - nbA) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaStreamSynchronize(ctx->copy_stream));
    for (int part = 0; part < 2; part++)
      for (int k = 0; k < slen; k++) (*total)[(size_t)k] += ctx->chunk_stats_host[(size_t)part * slen + k];
    float ms = 0.0f;
    CU(cudaEventElapsedTime(&ms, ctx->red0, ctx->red1));
    if (ms > *device_ms) *device_ms = ms;
  }
  return SMMC_OK;
}

smmc_status multi_comms(int n_gpus, const std::vector<smmc::NcclComm> **out);
smmc_status need_nccl(const smmc::NcclApi **api);
const std::vector<smmc::NcclComm> *multi_comms_or_null(int n_gpus);

// Host-pointer run of cfg over the devices of `ctxs` (1 = smmc_run, more = smmc_run_multi); caller holds every ctx->mu.
smmc_status run_host(const smmc_config *cfg, const std::vector<Context *> &ctxs, smmc_result *out) {
  std::vector<Shard> shards;
  make_shards(cfg, out, ctxs, &shards);
  const int slen = smmc_stats_len(cfg->hist_bins);
  std::vector<double> total((size_t)slen, 0.0);
  float ms_max = 0.0f;
  int launches = 0;
  const smmc::Variant variant = current_variant();
  int64_t biggest = 0;
  for (const Shard &sh : shards) biggest = std::max(biggest, (int64_t)sh.cfg.n_simulations);
  if (out->finals && variant.finals_chunks != 1 && (biggest >= kPipelineMinPaths || variant.finals_chunks > 1)) {
    smmc_status st = run_finals_pipelined(shards, cfg, &total, &ms_max, &launches);
    if (st != SMMC_OK) return st;
  } else if (!out->finals && (out->block_means || out->block_vars) && variant.finals_chunks != 1 && biggest >= kSplitMinPaths) {
    smmc_status st = run_partials_split(shards, cfg, &total, &ms_max, &launches);
    if (st != SMMC_OK) return st;
  } else {
    // every device's launch is enqueued before anything waits, and every download is issued before any is awaited
    std::vector<Pending> pend(shards.size());
    std::vector<smmc_result> outs(shards.size(), *out);
    smmc_status rc = SMMC_OK;
    size_t launched = 0;
    for (size_t g = 0; g < shards.size() && rc == SMMC_OK; g++) {
      outs[g].finals = shards[g].finals;
      outs[g].block_means = shards[g].bmeans;
      outs[g].block_vars = shards[g].bvars;
      rc = enqueue(&shards[g].cfg, &outs[g], shards[g].ctx, &pend[g]);
      if (rc == SMMC_OK) launched++;
    }
    // several devices: ONE grouped ncclAllReduce of the statistics vectors, in place, each on its kernel's stream -- the
    // merge happens on the devices (NVLink), and only device 0's 864 bytes are read back
    // Without NCCL (libnccl.so.2 not loadable, or no communicator over these devices) the per-shard vectors are downloaded
    // and added on the host instead, as the pipelined-finals and split-partials branches above do: a single-process
    // multi-GPU call never fails for want of NCCL.  (smmc_run_rank / smmc_quantiles_rank have no such stand-in.)
    bool merged_on_device = false;
    if (shards.size() > 1 && rc == SMMC_OK) {
      const std::vector<smmc::NcclComm> *comms = multi_comms_or_null((int)shards.size());
      const smmc::NcclApi *api = comms ? smmc::nccl_api() : nullptr;
      if (api) {
        merged_on_device = true;
        int r = api->GroupStart();
        for (size_t g = 0; g < launched && r == 0; g++)
          r = api->AllReduce(pend[g].stats_dev, pend[g].stats_dev, (size_t)slen, smmc::kNcclFloat64, smmc::kNcclSum, (*comms)[g], pend[g].stream);
        const int r2 = api->GroupEnd();
        if (r == 0) r = r2;
        if (r != 0) rc = fail(SMMC_ERR_CUDA, "ncclAllReduce failed: %s", api->GetErrorString(r));
      }
    }
    for (size_t g = 0; g < launched; g++) {
      smmc_status s2 = collect_issue(pend[g], &outs[g]);
      if (s2 != SMMC_OK && rc == SMMC_OK) rc = s2;
    }
    std::vector<double> part((size_t)slen);
    for (size_t g = 0; g < launched; g++) {
      float ms = 0.0f;
      smmc_status s2 = collect_wait(pend[g], part.data(), &ms);
      if (s2 != SMMC_OK && rc == SMMC_OK) rc = s2;
      if (!merged_on_device || g == 0)
        for (int i = 0; i < slen; i++) total[(size_t)i] += part[(size_t)i];
      if (ms > ms_max) ms_max = ms;
      launches += pend[g].launches;
    }
    if (rc != SMMC_OK) return rc;
  }
  out->device_ms = ms_max;
  out->gpu_launches = launches;
  return smmc_stats_finalize(total.data(), cfg->hist_bins, cfg->initial_capital, &out->count, &out->below_count, &out->mean,
                             &out->m2, out->hist);
}

// ---- NCCL (bound at run time, smmc_nccl.h) -------------------------------------------------------------------------------
#define NCCL_CALL(api, call)                                                                                            \
  do {                                                                                                                  \
    int r__ = (call);                                                                                                   \
    if (r__ != 0) return fail(SMMC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, (api)->GetErrorString(r__), __FILE__, __LINE__); \
  } while (0)

smmc_status need_nccl(const smmc::NcclApi **api) {
  *api = smmc::nccl_api();
  if (!*api)
    return fail(SMMC_ERR_UNSUPPORTED, "NCCL is not available (%s); the multi-GPU collective has no host-side stand-in", smmc::nccl_load_error());
  return SMMC_OK;
}

// first block-aligned path and path count of shard `rank` of `n` (the partition of make_shards)
void rank_range(int64_t N, int n, int rank, int64_t *first, int64_t *count) {
  int64_t b0 = 0, b1 = 0;
  shard_blocks(N, n, rank, &b0, &b1);
  int64_t p0 = b0 * SMMC_BLOCK, p1 = b1 * SMMC_BLOCK;
  if (p1 > N) p1 = N;
  if (p0 >= p1) p0 = p1 = 0;  // an empty shard keeps a block-aligned offset
  *first = p0;
  *count = p1 - p0;
}

// communicators of the single-process multi-GPU entry points (smmc_run_multi), one set per device count
std::mutex g_multi_mu;
std::vector<smmc::NcclComm> g_multi_comms[kMaxDevices + 1];

smmc_status multi_comms(int n_gpus, const std::vector<smmc::NcclComm> **out) {
  const smmc::NcclApi *api = nullptr;
  smmc_status st = need_nccl(&api);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(g_multi_mu);
  std::vector<smmc::NcclComm> &v = g_multi_comms[n_gpus];
  if (v.empty()) {
    std::vector<int> devs((size_t)n_gpus);
    for (int g = 0; g < n_gpus; g++) devs[(size_t)g] = g;
    std::vector<smmc::NcclComm> c((size_t)n_gpus, nullptr);
    NCCL_CALL(api, api->CommInitAll(c.data(), n_gpus, devs.data()));
    v = c;
  }
  *out = &v;
  return SMMC_OK;
}

// smmc_shutdown: the communicators die with the contexts whose streams they were used on
void release_multi_comms() {
  std::lock_guard<std::mutex> lk(g_multi_mu);
  const smmc::NcclApi *api = smmc::nccl_api();
  for (std::vector<smmc::NcclComm> &v : g_multi_comms) {
    if (api)
      for (smmc::NcclComm c : v)
        if (c) api->CommDestroy(c);
    v.clear();
  }
}

// The communicators of smmc_run_multi / smmc_quantiles, or null when the shards have to be merged on the host: NCCL cannot be
// loaded, ncclCommInitAll fails on these devices, or the test hook "host_merge" asks for that path.
const std::vector<smmc::NcclComm> *multi_comms_or_null(int n_gpus) {
  if (current_variant().host_merge == 1 || !smmc::nccl_api()) return nullptr;
  const std::vector<smmc::NcclComm> *comms = nullptr;
  if (multi_comms(n_gpus, &comms) != SMMC_OK) {
    g_err[0] = 0;  // not an error of the call: the host merge takes over
    return nullptr;
  }
  return comms;
}

// ---- select passes: replay, or -- for the last 10 key bits -- a pass over the candidate keys of the previous replay --------
// After the first pass fixes 11 key bits per target, the second replay also appends every key under one of the targets'
// 11-bit prefixes to a device buffer (mc_kernel epilogue, one atomicAdd per 256-path block).  The third pass then reads
// that buffer (4 B per candidate, HBM-bound, sub-millisecond at 10^8 paths) instead of replaying the simulation: exact
// order statistics in two replays (one when the first rides along with the statistics launch).  The buffer is sized for
// every path of the shard, so it cannot overflow; when that does not fit (budget below) the third replay runs as before.
// Validity is checked, never assumed: the fingerprint covers every field that determines the paths.
uint64_t fnv1a(const void *data, size_t bytes, uint64_t h = 1469598103934665603ull) {
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t i = 0; i < bytes; i++) h = (h ^ p[i]) * 1099511628211ull;
  return h;
}

CandFingerprint fingerprint_of(const smmc_config *cfg) {
  CandFingerprint f;
  memset(&f, 0, sizeof f);
  f.mode = cfg->mode;
  f.n = cfg->n_simulations;
  f.periods = cfg->n_periods;
  f.n_returns = cfg->n_returns;
  f.seed = cfg->seed;
  f.offset = cfg->stream_offset;
  f.v0 = cfg->initial_capital;
  f.gm = cfg->gauss_mean;
  f.gs = cfg->gauss_std;
  f.wfixed = cfg->withdraw_fixed;
  f.wpct = cfg->withdraw_pct;
  f.rounds = cfg->philox_rounds == 7 ? 7 : 10;
  const bool devp = (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) != 0;
  if (is_bootstrap(cfg->mode))
    f.table = devp ? (uint64_t)reinterpret_cast<uintptr_t>(cfg->returns) ^ 0x5555555555555555ull
                   : fnv1a(cfg->returns, (size_t)cfg->n_returns * 4);
  if (cfg->withdraw_pct_schedule) f.schedule = fnv1a(cfg->withdraw_pct_schedule, (size_t)cfg->n_periods * 4);
  const smmc::Variant v = current_variant();
  f.tbl_mask = gauss_table_mask(cfg->mode, v, cfg->philox_rounds);
  return f;
}

// Enqueues the select pass described by `sst` for the local shard `cfg` on ctx (caller holds ctx->mu).  hist_dev: the
// caller's device histogram (DEVICE_POINTERS) or null = ctx->selhist.  *hist_out / *words_out: where the digit histograms
// are on the device.  The kernel's events go into the context's ring (smmc_kernel_times) either way.
smmc_status select_pass_enqueue(Context *ctx, const smmc_config *cfg, const smmc_select_state *sst, unsigned long long *hist_dev,
                                cudaStream_t *stream_out, unsigned long long **hist_out, int *words_out, int *ev_slot_out,
                                int *launches_out) {
  SelectPass sp;
  smmc_status st = make_select_pass(sst, &sp);
  if (st != SMMC_OK) return st;
  sp.hist_dev = hist_dev;
  const bool devp = (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) != 0;
  cudaStream_t stream = devp ? (cudaStream_t)cfg->stream : ctx->stream;
  const CandFingerprint fp = fingerprint_of(cfg);
  CU(cudaSetDevice(ctx->device));

  // ---- last pass over the candidates of the previous replay?
  if (sst->bits_done == 22 && ctx->cand_valid && memcmp(&fp, &ctx->cand_fp, sizeof fp) == 0) {
    bool covered = true;
    for (int t = 0; t < sst->n_targets && covered; t++) {
      bool hit = false;
      for (int g = 0; g < ctx->cand_groups; g++) hit = hit || ((sst->prefix[t] >> 11) == ctx->cand_prefix[g]);
      covered = hit;
    }
    if (covered) {
      st = order_after_last_use(ctx, stream);
      if (st != SMMC_OK) return st;
      const size_t sel_bytes = (size_t)sp.targets * SMMC_SELECT_BINS * sizeof(unsigned long long);
      unsigned long long *hd = hist_dev;
      if (!hd) {
        CU(ctx->selhist.reserve(sel_bytes));
        hd = (unsigned long long *)ctx->selhist.p;
      }
      CU(cudaMemsetAsync(hd, 0, sel_bytes, stream));
      const int slot = (int)(ctx->ev_head % kEventRing);
      CU(cudaEventRecord(ctx->ev0[slot], stream));
      CU(smmc::launch_select_candidates((const uint32_t *)ctx->cand_keys.p, (const unsigned long long *)ctx->cand_cursor.p,
                                        ctx->cand_capacity, sp.targets, sp.digit_shift, sp.digit_mask, sp.prefix_shift, sp.prefix, hd,
                                        ctx->n_sms, stream));
      CU(cudaEventRecord(ctx->ev1[slot], stream));
      ctx->ev_kind[slot] = SMMC_KERNEL_SELECT_CANDIDATES;
      ctx->ev_head++;
      st = mark_last_use(ctx, stream);
      if (st != SMMC_OK) return st;
      *stream_out = stream;
      *hist_out = hd;
      *words_out = sp.targets * SMMC_SELECT_BINS;
      *ev_slot_out = slot;
      *launches_out = 1;
      return SMMC_OK;
    }
  }

  // ---- a replay; the second one (11 bits known) collects the candidates for the third
  ctx->cand_valid = false;
  bool collecting = false;
  if (sst->bits_done == 11 && cfg->n_simulations > 0 && current_variant().select_candidates != 0) {
    size_t free_b = 0, total_b = 0;
    const size_t need = (size_t)cfg->n_simulations * 4;
    const bool have = ctx->cand_keys.cap >= need;
    if (have || (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && need <= free_b / 2 && need <= ((size_t)8 << 30))) {
      if (ctx->cand_keys.reserve(need) == cudaSuccess && ctx->cand_cursor.reserve(16) == cudaSuccess) {
        int groups = 0;
        for (int t = 0; t < sst->n_targets; t++) {
          bool seen = false;
          for (int g = 0; g < groups; g++) seen = seen || (sp.cand_prefix[g] == sst->prefix[t]);
          if (!seen) sp.cand_prefix[groups++] = sst->prefix[t];
        }
        sp.cand_groups = groups;
        sp.cand_keys = (uint32_t *)ctx->cand_keys.p;
        sp.cand_cursor = (unsigned long long *)ctx->cand_cursor.p;
        sp.cand_capacity = (unsigned long long)cfg->n_simulations;
        collecting = true;
      } else {
        cudaGetLastError();  // out of memory: no candidates, the third pass replays
      }
    }
  }
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  Pending pd;
  st = enqueue(cfg, &dummy, ctx, &pd, &sp);
  if (st != SMMC_OK) return st;
  if (collecting) {
    ctx->cand_valid = true;
    ctx->cand_fp = fp;
    ctx->cand_groups = sp.cand_groups;
    for (int g = 0; g < SMMC_SELECT_MAX_TARGETS; g++) ctx->cand_prefix[g] = sp.cand_prefix[g];
    ctx->cand_capacity = sp.cand_capacity;
  }
  *stream_out = pd.stream;
  *hist_out = pd.sel_dev;
  *words_out = pd.sel_words;
  *ev_slot_out = pd.ev_slot;
  *launches_out = pd.launches;
  return SMMC_OK;
}

}  // namespace

extern "C" {

int32_t smmc_stats_len(int32_t hist_bins) { return SMMC_STATS_HEADER + (hist_bins > 0 ? hist_bins : 0) + 2; }

smmc_status smmc_stats_finalize(const double *stats, int32_t hist_bins, float initial_capital, uint64_t *count,
                                uint64_t *below_count, double *mean, double *m2, uint64_t *hist) {
  if (!stats) return fail(SMMC_ERR_INVALID_ARG, "null stats");
  const double n = stats[0];
  if (count) *count = (uint64_t)n;
  if (below_count) *below_count = (uint64_t)stats[1];
  if (n > 0.0) {
    const double s1 = stats[2], s2 = stats[3];
    if (mean) *mean = (double)initial_capital + s1 / n;
    if (m2) {
      const double v = s2 - s1 * s1 / n;
      *m2 = v > 0.0 ? v : 0.0;
    }
  } else {
    if (mean) *mean = 0.0;
    if (m2) *m2 = 0.0;
  }
  if (hist && hist_bins > 0)
    for (int i = 0; i < hist_bins + 2; i++) hist[i] = (uint64_t)stats[SMMC_STATS_HEADER + i];
  return SMMC_OK;
}

smmc_status smmc_run(const smmc_config *cfg, smmc_result *out) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_status st = validate(cfg, out);
  if (st != SMMC_OK) return st;
  Context *ctx = nullptr;
  st = get_context(cfg->device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!(cfg->flags & SMMC_FLAG_DEVICE_POINTERS)) return run_host(cfg, std::vector<Context *>(1, ctx), out);
  Pending pd;
  st = enqueue(cfg, out, ctx, &pd);
  if (st != SMMC_OK) return st;
  out->gpu_launches = pd.launches;
  if (cfg->flags & SMMC_FLAG_NO_SYNC) return SMMC_OK;
  std::vector<double> stats((size_t)pd.stats_len);
  float ms = 0.0f;
  st = collect(pd, out, stats.data(), &ms);
  if (st != SMMC_OK) return st;
  out->device_ms = ms;
  return smmc_stats_finalize(stats.data(), cfg->hist_bins, cfg->initial_capital, &out->count, &out->below_count,
                             &out->mean, &out->m2, out->hist);
}

smmc_status smmc_run_multi(const smmc_config *cfg, int n_gpus, smmc_result *out) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_status st = validate(cfg, out);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) return fail(SMMC_ERR_INVALID_ARG, "smmc_run_multi takes host pointers");
  const int ndev = probe_devices();
  if (ndev <= 0) return fail(SMMC_ERR_NO_DEVICE, "no CUDA device of compute capability 10.x (sm_100a) is visible; this library has no CPU fallback");
  if (n_gpus < 1 || n_gpus > ndev) return fail(SMMC_ERR_INVALID_ARG, "n_gpus %d out of range [1,%d]", n_gpus, ndev);
  if (n_gpus == 1) return smmc_run(cfg, out);
  std::vector<Context *> ctxs((size_t)n_gpus, nullptr);
  for (int g = 0; g < n_gpus; g++) {
    st = get_context(g, &ctxs[(size_t)g]);
    if (st != SMMC_OK) return st;
  }
  for (int g = 0; g < n_gpus; g++) ctxs[(size_t)g]->mu.lock();
  st = run_host(cfg, ctxs, out);
  for (int g = 0; g < n_gpus; g++) ctxs[(size_t)g]->mu.unlock();
  return st;
}

// ---- one rank per GPU ------------------------------------------------------------------------------------------------------

smmc_status smmc_comm_unique_id(void *id) {
  if (!id) return fail(SMMC_ERR_INVALID_ARG, "comm_unique_id: null buffer");
  const smmc::NcclApi *api = nullptr;
  smmc_status st = need_nccl(&api);
  if (st != SMMC_OK) return st;
  smmc::NcclUniqueId uid;
  NCCL_CALL(api, api->GetUniqueId(&uid));
  static_assert(sizeof uid == SMMC_COMM_ID_BYTES, "NCCL_UNIQUE_ID_BYTES");
  memcpy(id, &uid, sizeof uid);
  return SMMC_OK;
}

smmc_status smmc_comm_init_rank(const void *id, int32_t n_ranks, int32_t rank, int32_t device, smmc_comm **comm) {
  DeviceGuard device_guard;
  if (!id || !comm || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(SMMC_ERR_INVALID_ARG, "comm_init_rank: bad arguments");
  const smmc::NcclApi *api = nullptr;
  smmc_status st = need_nccl(&api);
  if (st != SMMC_OK) return st;
  Context *ctx = nullptr;
  st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  CU(cudaSetDevice(device));
  smmc::NcclUniqueId uid;
  memcpy(&uid, id, sizeof uid);
  smmc::NcclComm c = nullptr;
  NCCL_CALL(api, api->CommInitRank(&c, n_ranks, uid, rank));
  smmc_comm *out = new smmc_comm;
  out->nccl = c;
  out->n_ranks = n_ranks;
  out->rank = rank;
  out->device = device;
  out->owned = true;
  *comm = out;
  int ver = 0;
  api->GetVersion(&ver);
  // a library stays silent unless asked: smmc_debug_set_variant("verbose", 1) (bench.py does) gives one line per rank on
  // stderr (stdout stays free for the caller's own output, e.g. bench.py's JSON line)
  if (current_variant().verbose == 1)
    fprintf(stderr, "smmc: NCCL %d communicator ready: rank %d nranks %d cudaDev %d\n", ver, rank, n_ranks, device);
  return SMMC_OK;
}

smmc_status smmc_comm_adopt(void *nccl_comm, int32_t device, smmc_comm **comm) {
  DeviceGuard device_guard;  // get_context may switch the device; the caller's choice is restored on every return path
  if (!nccl_comm || !comm) return fail(SMMC_ERR_INVALID_ARG, "comm_adopt: null argument");
  const smmc::NcclApi *api = nullptr;
  smmc_status st = need_nccl(&api);
  if (st != SMMC_OK) return st;
  Context *ctx = nullptr;
  st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  int n = 0, r = 0;
  NCCL_CALL(api, api->CommCount((smmc::NcclComm)nccl_comm, &n));
  NCCL_CALL(api, api->CommUserRank((smmc::NcclComm)nccl_comm, &r));
  smmc_comm *out = new smmc_comm;
  out->nccl = (smmc::NcclComm)nccl_comm;
  out->n_ranks = n;
  out->rank = r;
  out->device = device;
  out->owned = false;
  *comm = out;
  return SMMC_OK;
}

smmc_status smmc_comm_allreduce_count(const smmc_comm *comm, uint64_t *count) {
  if (!comm || !count) return fail(SMMC_ERR_INVALID_ARG, "comm_allreduce_count: null argument");
  *count = comm->allreduces;
  return SMMC_OK;
}

smmc_status smmc_comm_destroy(smmc_comm *comm) {
  if (!comm) return SMMC_OK;
  if (comm->owned && comm->nccl) {
    const smmc::NcclApi *api = smmc::nccl_api();
    if (api) api->CommDestroy(comm->nccl);
  }
  delete comm;
  return SMMC_OK;
}

smmc_status smmc_comm_info(const smmc_comm *comm, int32_t *n_ranks, int32_t *rank, int32_t *device, int32_t *nccl_version) {
  if (!comm) return fail(SMMC_ERR_INVALID_ARG, "comm_info: null communicator");
  if (n_ranks) *n_ranks = comm->n_ranks;
  if (rank) *rank = comm->rank;
  if (device) *device = comm->device;
  if (nccl_version) {
    int v = 0;
    const smmc::NcclApi *api = smmc::nccl_api();
    if (api) api->GetVersion(&v);
    *nccl_version = v;
  }
  return SMMC_OK;
}

smmc_status smmc_rank_range(int64_t n_simulations, int32_t n_ranks, int32_t rank, int64_t *first, int64_t *count) {
  if (n_simulations < 0 || n_ranks < 1 || rank < 0 || rank >= n_ranks || !first || !count)
    return fail(SMMC_ERR_INVALID_ARG, "rank_range: bad arguments");
  rank_range(n_simulations, n_ranks, rank, first, count);
  return SMMC_OK;
}

smmc_status smmc_run_rank(const smmc_config *cfg, smmc_result *out, smmc_comm *comm) {
  DeviceGuard device_guard;
  smmc_status st = validate(cfg, out);
  if (st != SMMC_OK) return st;
  if (!comm) return fail(SMMC_ERR_INVALID_ARG, "run_rank: null communicator");
  const smmc::NcclApi *api = nullptr;
  st = need_nccl(&api);
  if (st != SMMC_OK) return st;
  Context *ctx = nullptr;
  st = get_context(comm->device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  smmc_config c2 = *cfg;
  int64_t first = 0, count = 0;
  rank_range(cfg->n_simulations, comm->n_ranks, comm->rank, &first, &count);
  c2.device = comm->device;
  c2.stream_offset = cfg->stream_offset + (uint64_t)first;
  c2.n_simulations = count;
  st = validate(&c2, out, false);  // 32-bit id limits of COMPAT / Sobol apply to the shard as well
  if (st != SMMC_OK) return st;
  const bool host_ptrs = (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) == 0;
  if (host_ptrs && !out->finals && (out->block_means || out->block_vars) && current_variant().finals_chunks != 1 &&
      cfg->n_simulations / comm->n_ranks >= kSplitMinPaths) {
    // block partials into host memory: two launches (ONE collective, after the second) so that most of the download hides
    // behind the second kernel; the criterion is a function of the job, not of the shard, so every rank takes the same branch
    std::vector<Shard> one(1);
    one[0].ctx = ctx;
    one[0].cfg = c2;
    one[0].bmeans = out->block_means;
    one[0].bvars = out->block_vars;
    std::vector<double> total;
    float ms = 0.0f;
    int launches = 0;
    st = run_partials_split(one, cfg, &total, &ms, &launches, api, comm->nccl, &comm->allreduces);
    if (st != SMMC_OK) return st;
    out->device_ms = ms;
    out->gpu_launches = launches;
    return smmc_stats_finalize(total.data(), cfg->hist_bins, cfg->initial_capital, &out->count, &out->below_count, &out->mean,
                               &out->m2, out->hist);
  }
  Pending pd;
  st = enqueue(&c2, out, ctx, &pd);
  if (st != SMMC_OK) return st;
  out->gpu_launches = pd.launches;
  // the ONE collective of the path: O(KB), on the kernel's stream (stream order = data dependency)
  NCCL_CALL(api, api->AllReduce(pd.stats_dev, pd.stats_dev, (size_t)pd.stats_len, smmc::kNcclFloat64, smmc::kNcclSum, comm->nccl, pd.stream));
  comm->allreduces++;
  st = mark_last_use(ctx, pd.stream);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_NO_SYNC) return SMMC_OK;
  std::vector<double> stats((size_t)pd.stats_len);
  float ms = 0.0f;
  st = collect(pd, out, stats.data(), &ms);
  if (st != SMMC_OK) return st;
  out->device_ms = ms;
  return smmc_stats_finalize(stats.data(), cfg->hist_bins, cfg->initial_capital, &out->count, &out->below_count, &out->mean,
                             &out->m2, out->hist);
}

namespace {
// sum of n device floats on `st`; blocks until done; kernel_ms (nullable) = CUDA-event time of the reduction kernel
smmc_status reduce_sum_device(Context *ctx, const float *v_dev, int64_t n, cudaStream_t st, double *sum, float *kernel_ms) {
  const size_t part = (size_t)ctx->n_sms * 4 * sizeof(double);
  // ctx->work also holds the CTA partials, histogram and ticket of a simulation launch: a SMMC_FLAG_NO_SYNC run that is still in
  // flight on another stream must finish with them before this reduction overwrites them (and later users wait for this one)
  smmc_status os = order_after_last_use(ctx, st);
  if (os != SMMC_OK) return os;
  CU(ctx->work.reserve(part + 32));
  char *w = (char *)ctx->work.p;
  CU(cudaMemsetAsync(w + part, 0, 32, st));
  cudaEvent_t e0 = ctx->red0, e1 = ctx->red1;
  if (kernel_ms) CU(cudaEventRecord(e0, st));
  CU(smmc::launch_reduce_sum(v_dev, n, (double *)w, (unsigned int *)(w + part + 16), (double *)(w + part), ctx->n_sms,
                             current_variant().reduce, st));
  if (kernel_ms) CU(cudaEventRecord(e1, st));
  os = mark_last_use(ctx, st);
  if (os != SMMC_OK) return os;
  CU(cudaMemcpyAsync(sum, w + part, sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (kernel_ms) CU(cudaEventElapsedTime(kernel_ms, e0, e1));
  return SMMC_OK;
}
}  // namespace

smmc_status smmc_reduce_mean(const float *vec, int64_t n, int32_t device, float *mean_out) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (!vec || !mean_out || n <= 0) return fail(SMMC_ERR_INVALID_ARG, "reduce_mean: need n > 0 and non-null pointers");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  CU(ctx->scratch.reserve((size_t)n * 4));
  CU(cudaMemcpyAsync(ctx->scratch.p, vec, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  double sum = 0.0;
  st = reduce_sum_device(ctx, (const float *)ctx->scratch.p, n, ctx->stream, &sum, nullptr);
  if (st != SMMC_OK) return st;
  *mean_out = (float)(sum / (double)n);
  return SMMC_OK;
}

smmc_status smmc_reduce_mean_device(const float *vec_dev, int64_t n, int32_t device, void *stream, float *mean_out,
                                    float *kernel_ms) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (!vec_dev || !mean_out || n <= 0)
    return fail(SMMC_ERR_INVALID_ARG, "reduce_mean_device: need n > 0 and non-null pointers");
  if (reinterpret_cast<uintptr_t>(vec_dev) & 3u) return fail(SMMC_ERR_INVALID_ARG, "reduce_mean_device: pointer must be 4-byte aligned");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  double sum = 0.0;
  st = reduce_sum_device(ctx, vec_dev, n, (cudaStream_t)stream, &sum, kernel_ms);
  if (st != SMMC_OK) return st;
  *mean_out = (float)(sum / (double)n);
  return SMMC_OK;
}

smmc_status smmc_vector_add(float *out, const float *a, const float *b, int64_t n, int32_t device) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (n < 0 || (n > 0 && (!out || !a || !b))) return fail(SMMC_ERR_INVALID_ARG, "vector_add: bad arguments");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  if (n == 0) return SMMC_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  CU(ctx->scratch.reserve((size_t)n * 12));
  float *da = (float *)ctx->scratch.p, *db = da + n, *dout = db + n;
  CU(cudaMemcpyAsync(da, a, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaMemcpyAsync(db, b, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(smmc::launch_vector_add(dout, da, db, n, ctx->n_sms, ctx->stream));
  CU(cudaMemcpyAsync(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return SMMC_OK;
}

smmc_status smmc_vector_add_device(float *out_dev, const float *a_dev, const float *b_dev, int64_t n, int32_t device,
                                   void *stream, float *kernel_ms) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (n < 0 || (n > 0 && (!out_dev || !a_dev || !b_dev))) return fail(SMMC_ERR_INVALID_ARG, "vector_add_device: bad arguments");
  if ((reinterpret_cast<uintptr_t>(out_dev) | reinterpret_cast<uintptr_t>(a_dev) | reinterpret_cast<uintptr_t>(b_dev)) & 3u)
    return fail(SMMC_ERR_INVALID_ARG, "vector_add_device: pointers must be 4-byte aligned");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  cudaStream_t s = (cudaStream_t)stream;
  if (kernel_ms) CU(cudaEventRecord(ctx->red0, s));
  CU(smmc::launch_vector_add(out_dev, a_dev, b_dev, n, ctx->n_sms, s));
  if (kernel_ms) CU(cudaEventRecord(ctx->red1, s));
  CU(cudaStreamSynchronize(s));
  if (kernel_ms) CU(cudaEventElapsedTime(kernel_ms, ctx->red0, ctx->red1));
  return SMMC_OK;
}

// ---- exact order statistics by replay (radix select) ------------------------------------------------------------

smmc_status smmc_select_begin(smmc_select_state *st, int32_t n_targets, const uint64_t *ranks) {
  if (!st || !ranks || n_targets < 1 || n_targets > SMMC_SELECT_MAX_TARGETS)
    return fail(SMMC_ERR_INVALID_ARG, "select_begin: n_targets must be in [1,%d]", SMMC_SELECT_MAX_TARGETS);
  memset(st, 0, sizeof *st);
  st->n_targets = n_targets;
  for (int t = 0; t < n_targets; t++) st->rank[t] = ranks[t];
  return SMMC_OK;
}

smmc_status smmc_select_pass(const smmc_config *cfg, const smmc_select_state *sst, uint64_t *hist) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  if (!hist) return fail(SMMC_ERR_INVALID_ARG, "select_pass: null histogram");
  smmc_config c2;
  if (cfg) {
    c2 = *cfg;
    c2.hist_bins = 0;
    c2.flags &= ~SMMC_FLAG_NO_SYNC;  // validated below without the stats_dev requirement
  }
  smmc_status st = validate(cfg ? &c2 : nullptr, &dummy);
  if (st != SMMC_OK) return st;
  SelectPass probe;
  st = make_select_pass(sst, &probe);
  if (st != SMMC_OK) return st;
  Context *ctx = nullptr;
  st = get_context(cfg->device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  const bool devp = (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) != 0;
  const size_t all_bytes = (size_t)SMMC_SELECT_MAX_TARGETS * SMMC_SELECT_BINS * sizeof(uint64_t);
  cudaStream_t stream = nullptr;
  unsigned long long *hd = nullptr;
  int words = 0, slot = 0, launches = 0;
  if (devp) {
    CU(cudaSetDevice(cfg->device));
    CU(cudaMemsetAsync(hist, 0, all_bytes, (cudaStream_t)cfg->stream));
    return select_pass_enqueue(ctx, &c2, sst, (unsigned long long *)hist, &stream, &hd, &words, &slot, &launches);  // asynchronous
  }
  st = select_pass_enqueue(ctx, &c2, sst, nullptr, &stream, &hd, &words, &slot, &launches);
  if (st != SMMC_OK) return st;
  memset(hist, 0, all_bytes);
  CU(cudaMemcpyAsync(hist, hd, (size_t)words * sizeof(uint64_t), cudaMemcpyDeviceToHost, stream));
  CU(cudaStreamSynchronize(stream));
  return SMMC_OK;
}

smmc_status smmc_run_select(const smmc_config *cfg, smmc_result *out, const smmc_select_state *sst, uint64_t *hist) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_status st = validate(cfg, out);
  if (st != SMMC_OK) return st;
  if (!hist) return fail(SMMC_ERR_INVALID_ARG, "run_select: null histogram");
  if (cfg->flags & (SMMC_FLAG_DEVICE_POINTERS | SMMC_FLAG_NO_SYNC))
    return fail(SMMC_ERR_INVALID_ARG, "run_select: host pointers only (use smmc_run + smmc_select_pass with device pointers)");
  SelectPass sp;
  st = make_select_pass(sst, &sp);
  if (st != SMMC_OK) return st;
  if (sst->bits_done != 0) return fail(SMMC_ERR_INVALID_ARG, "run_select: only the first pass can ride along (bits_done = %d)", sst->bits_done);
  Context *ctx = nullptr;
  st = get_context(cfg->device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  Pending pd;
  st = enqueue(cfg, out, ctx, &pd, &sp);
  if (st != SMMC_OK) return st;
  out->gpu_launches = pd.launches;
  std::vector<double> stats((size_t)pd.stats_len);
  float ms = 0.0f;
  memset(hist, 0, (size_t)SMMC_SELECT_MAX_TARGETS * SMMC_SELECT_BINS * sizeof(uint64_t));
  CU(cudaMemcpyAsync(hist, pd.sel_dev, (size_t)pd.sel_words * sizeof(uint64_t), cudaMemcpyDeviceToHost, pd.stream));
  st = collect(pd, out, stats.data(), &ms);
  if (st != SMMC_OK) return st;
  out->device_ms = ms;
  return smmc_stats_finalize(stats.data(), cfg->hist_bins, cfg->initial_capital, &out->count, &out->below_count,
                             &out->mean, &out->m2, out->hist);
}

smmc_status smmc_select_advance(smmc_select_state *st, const uint64_t *hist) {
  if (!st || !hist) return fail(SMMC_ERR_INVALID_ARG, "select_advance: null argument");
  SelectPass sp;
  smmc_status rc = make_select_pass(st, &sp);
  if (rc != SMMC_OK) return rc;
  const int w = select_width(st->bits_done);
  const int bins = 1 << w;
  for (int t = 0; t < st->n_targets; t++) {
    const uint64_t *h = hist + (size_t)(st->bits_done == 0 ? 0 : t) * SMMC_SELECT_BINS;
    uint64_t cum = 0;
    int b = 0;
    for (; b < bins; b++) {
      if (cum + h[b] > st->rank[t]) break;
      cum += h[b];
    }
    if (b == bins) return fail(SMMC_ERR_INVALID_ARG, "select_advance: rank of target %d is beyond the number of paths", t);
    st->rank[t] -= cum;
    st->prefix[t] = (st->prefix[t] << w) | (uint32_t)b;
  }
  st->bits_done += w;
  return SMMC_OK;
}

smmc_status smmc_select_values(const smmc_select_state *st, float *values) {
  if (!st || !values) return fail(SMMC_ERR_INVALID_ARG, "select_values: null argument");
  if (st->bits_done != 32) return fail(SMMC_ERR_INVALID_ARG, "select_values: %d of 32 key bits resolved", st->bits_done);
  for (int t = 0; t < st->n_targets; t++) {
    const uint32_t key = st->prefix[t];
    const uint32_t bits = (key & 0x80000000u) ? (key ^ 0x80000000u) : ~key;  // inverse of smmc::ordered_key
    memcpy(&values[t], &bits, 4);
  }
  return SMMC_OK;
}

namespace {
// The select loop shared by smmc_quantiles (devices of this process) and smmc_quantiles_rank (one rank of a job): per pass,
// every local shard enqueues its replay / candidate pass, the digit histograms are merged -- ncclAllReduce(ncclUint64) on
// the devices when more than one shard or rank takes part -- and ONE histogram is downloaded.  Caller holds the ctx locks.
smmc_status select_loop(std::vector<Shard> &shards, const std::vector<smmc::NcclComm> *comms, smmc_select_state *sel, float *ms_total) {
  const smmc::NcclApi *api = nullptr;
  if (comms) {
    smmc_status st = need_nccl(&api);
    if (st != SMMC_OK) return st;
  }
  const size_t words_all = (size_t)SMMC_SELECT_MAX_TARGETS * SMMC_SELECT_BINS;
  std::vector<uint64_t> total(words_all);
  *ms_total = 0.0f;
  const size_t S = shards.size();
  while (sel->bits_done < 32) {
    std::vector<cudaStream_t> streams(S, nullptr);
    std::vector<unsigned long long *> hd(S, nullptr);
    std::vector<int> words(S, 0), slots(S, 0);
    for (size_t g = 0; g < S; g++) {
      int launches = 0;
      smmc_config c2 = shards[g].cfg;
      c2.hist_bins = 0;
      smmc_status st = select_pass_enqueue(shards[g].ctx, &c2, sel, nullptr, &streams[g], &hd[g], &words[g], &slots[g], &launches);
      if (st != SMMC_OK) return st;
    }
    if (comms) {
      NCCL_CALL(api, api->GroupStart());
      for (size_t g = 0; g < S; g++) {
        int r = api->AllReduce(hd[g], hd[g], (size_t)words[g], smmc::kNcclUint64, smmc::kNcclSum, (*comms)[g], streams[g]);
        if (r != 0) {
          api->GroupEnd();
          return fail(SMMC_ERR_CUDA, "ncclAllReduce failed: %s", api->GetErrorString(r));
        }
      }
      NCCL_CALL(api, api->GroupEnd());
    }
    std::fill(total.begin(), total.end(), 0ull);
    float ms_pass = 0.0f;
    // merged on the devices: shard 0 holds the sum.  No communicator (one shard, or NCCL absent in smmc_quantiles): every
    // shard's histogram is downloaded and the words are added on the host.
    std::vector<uint64_t> parts(comms || S == 1 ? 0 : S * words_all, 0ull);
    for (size_t g = 0; g < S; g++) {
      Context *ctx = shards[g].ctx;
      CU(cudaSetDevice(ctx->device));
      if (g == 0 && parts.empty())
        CU(cudaMemcpyAsync(total.data(), hd[g], (size_t)words[g] * sizeof(uint64_t), cudaMemcpyDeviceToHost, streams[g]));
      else if (!parts.empty())
        CU(cudaMemcpyAsync(parts.data() + g * words_all, hd[g], (size_t)words[g] * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                           streams[g]));
    }
    for (size_t g = 0; g < S; g++) {
      Context *ctx = shards[g].ctx;
      CU(cudaSetDevice(ctx->device));
      CU(cudaStreamSynchronize(streams[g]));
      float ms = 0.0f;
      CU(cudaEventElapsedTime(&ms, ctx->ev0[slots[g]], ctx->ev1[slots[g]]));
      if (ms > ms_pass) ms_pass = ms;
      if (!parts.empty())
        for (size_t i = 0; i < (size_t)words[g]; i++) total[i] += parts[g * words_all + i];
    }
    *ms_total += ms_pass;
    smmc_status st = smmc_select_advance(sel, total.data());
    if (st != SMMC_OK) return st;
  }
  return SMMC_OK;
}
}  // namespace

smmc_status smmc_quantiles(const smmc_config *cfg, int32_t n_gpus, int32_t n_targets, const uint64_t *ranks,
                           float *values, float *device_ms) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  smmc_status st = validate(cfg, &dummy);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) return fail(SMMC_ERR_INVALID_ARG, "smmc_quantiles takes host pointers");
  if (!values) return fail(SMMC_ERR_INVALID_ARG, "quantiles: null output");
  smmc_select_state sel;
  st = smmc_select_begin(&sel, n_targets, ranks);
  if (st != SMMC_OK) return st;
  for (int t = 0; t < n_targets; t++)
    if (ranks[t] >= (uint64_t)cfg->n_simulations)
      return fail(SMMC_ERR_INVALID_ARG, "quantiles: rank %llu >= n_simulations", (unsigned long long)ranks[t]);
  const int ndev = probe_devices();
  if (ndev <= 0) return fail(SMMC_ERR_NO_DEVICE, "no CUDA device of compute capability 10.x (sm_100a) is visible; this library has no CPU fallback");
  if (n_gpus < 1 || n_gpus > ndev) return fail(SMMC_ERR_INVALID_ARG, "n_gpus %d out of range [1,%d]", n_gpus, ndev);
  std::vector<Context *> ctxs((size_t)n_gpus, nullptr);
  for (int g = 0; g < n_gpus; g++) {
    st = get_context(n_gpus == 1 ? cfg->device : g, &ctxs[(size_t)g]);
    if (st != SMMC_OK) return st;
  }
  // null = one device, or no NCCL: select_loop then adds the shards' digit histograms on the host
  const std::vector<smmc::NcclComm> *comms = n_gpus > 1 ? multi_comms_or_null(n_gpus) : nullptr;
  std::vector<Shard> shards;
  make_shards(cfg, &dummy, ctxs, &shards);
  for (int g = 0; g < n_gpus; g++) ctxs[(size_t)g]->mu.lock();
  float ms_total = 0.0f;
  st = select_loop(shards, comms, &sel, &ms_total);
  for (int g = 0; g < n_gpus; g++) ctxs[(size_t)g]->mu.unlock();
  if (st != SMMC_OK) return st;
  if (device_ms) *device_ms = ms_total;
  return smmc_select_values(&sel, values);
}

smmc_status smmc_quantiles_rank(const smmc_config *cfg, smmc_comm *comm, int32_t n_targets, const uint64_t *ranks, float *values,
                                float *device_ms) {
  DeviceGuard device_guard;
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  smmc_status st = validate(cfg, &dummy);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) return fail(SMMC_ERR_INVALID_ARG, "smmc_quantiles_rank takes host pointers");
  if (!values || !comm) return fail(SMMC_ERR_INVALID_ARG, "quantiles_rank: null argument");
  smmc_select_state sel;
  st = smmc_select_begin(&sel, n_targets, ranks);
  if (st != SMMC_OK) return st;
  for (int t = 0; t < n_targets; t++)
    if (ranks[t] >= (uint64_t)cfg->n_simulations)
      return fail(SMMC_ERR_INVALID_ARG, "quantiles: rank %llu >= n_simulations", (unsigned long long)ranks[t]);
  Context *ctx = nullptr;
  st = get_context(comm->device, &ctx);
  if (st != SMMC_OK) return st;
  std::vector<Shard> shards(1);
  int64_t first = 0, count = 0;
  rank_range(cfg->n_simulations, comm->n_ranks, comm->rank, &first, &count);
  shards[0].ctx = ctx;
  shards[0].cfg = *cfg;
  shards[0].cfg.device = comm->device;
  shards[0].cfg.stream_offset = cfg->stream_offset + (uint64_t)first;
  shards[0].cfg.n_simulations = count;
  const std::vector<smmc::NcclComm> comms(1, comm->nccl);
  std::lock_guard<std::mutex> lk(ctx->mu);
  float ms_total = 0.0f;
  st = select_loop(shards, &comms, &sel, &ms_total);
  if (st != SMMC_OK) return st;
  if (device_ms) *device_ms = ms_total;
  return smmc_select_values(&sel, values);
}

// ---- trajectory mode -----------------------------------------------------------------------------------------------

smmc_status smmc_trajectories(const smmc_config *cfg, int32_t n_paths, float *values, float *returns_pct) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  smmc_config c2;
  if (cfg) {
    c2 = *cfg;
    c2.n_simulations = n_paths;  // range checks (32-bit path ids, Sobol point indices) apply to the paths asked for
    c2.hist_bins = 0;
  }
  smmc_status st = validate(cfg ? &c2 : nullptr, &dummy);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) return fail(SMMC_ERR_INVALID_ARG, "trajectories takes host pointers");
  if (n_paths < 0 || (n_paths > 0 && !values)) return fail(SMMC_ERR_INVALID_ARG, "trajectories: bad arguments");
  Context *ctx = nullptr;
  st = get_context(cfg->device, &ctx);
  if (st != SMMC_OK) return st;
  if (n_paths == 0) return SMMC_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(cfg->device));
  cudaStream_t s = ctx->stream;
  st = order_after_last_use(ctx, s);  // the cached buffers written below may still be read by an asynchronous launch
  if (st != SMMC_OK) return st;
  const int P = cfg->n_periods, n = cfg->n_returns;
  smmc::KParams p;
  memset(&p, 0, sizeof p);
  fill_stream_params(cfg, &p);
  const float *tbl = nullptr, *raw = nullptr;
  if (is_bootstrap(cfg->mode)) {
    CU(ctx->returns.reserve((size_t)n * 4));
    CU(cudaMemcpyAsync(ctx->returns.p, cfg->returns, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    raw = (const float *)ctx->returns.p;
    if (cfg->mode == SMMC_MODE_SOBOL_BOOTSTRAP) {
      CU(ctx->sorted.reserve((size_t)n * 4));
      smmc::launch_sort_table(raw, n, (float *)ctx->sorted.p, s);
      CU(cudaGetLastError());
      raw = (const float *)ctx->sorted.p;
    }
    const uint32_t tbl_floats = (uint32_t)((n + 1 + 3) & ~3);
    CU(ctx->tbl.reserve((size_t)tbl_floats * 4));
    smmc::launch_prep_table(raw, n, 1, (float *)ctx->tbl.p, tbl_floats, s);
    CU(cudaGetLastError());
    tbl = (const float *)ctx->tbl.p;
  }
  const uint32_t *full_dev = nullptr, *shift_dev = nullptr;
  if (is_sobol(cfg->mode)) {
    st = prepare_sobol_full(ctx, P, cfg->seed, &full_dev, &shift_dev);
    if (st != SMMC_OK) return st;
  }
  const size_t vbytes = (size_t)n_paths * (size_t)(P + 1) * 4, rbytes = returns_pct ? (size_t)n_paths * (size_t)P * 4 : 0;
  CU(ctx->scratch.reserve(vbytes + rbytes + 16));
  float *vd = (float *)ctx->scratch.p, *rd = rbytes ? vd + (size_t)n_paths * (P + 1) : nullptr;
  int flow_slot = -1;
  st = prepare_flow(ctx, cfg, s, &p, &flow_slot);
  if (st != SMMC_OK) return st;
  if (p.gauss_tbl_mask) {  // the Gaussian modes' sin/cos table (read from global memory here)
    st = prepare_sincos(ctx, p.gc.k, s);
    if (st != SMMC_OK) return st;
    p.tbl_rep = sincos_table_of(ctx, cfg->mode);
  }
  CU(smmc::launch_trajectories(cfg->mode, p, n_paths, tbl, raw, full_dev, shift_dev, vd, rd, s));
  st = versioned_mark_read(ctx->flow_v, flow_slot, s);
  if (st != SMMC_OK) return st;
  CU(cudaMemcpyAsync(values, vd, vbytes, cudaMemcpyDeviceToHost, s));
  if (rbytes) CU(cudaMemcpyAsync(returns_pct, rd, rbytes, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  return SMMC_OK;
}

// ---- launch planning -------------------------------------------------------------------------------------------------

smmc_status smmc_plan_launch(const smmc_config *cfg, int32_t select_targets, int32_t n_sms, uint64_t smem_optin,
                             smmc_plan *plan) {
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  smmc_config c2;
  if (cfg) {
    c2 = *cfg;
    if (!c2.returns && c2.n_returns > 0) c2.returns = (const float *)(uintptr_t)16;  // planning never reads the table
  }
  smmc_status st = validate(cfg ? &c2 : nullptr, &dummy, false);
  if (st != SMMC_OK) return st;
  if (!plan || n_sms < 1 || select_targets < 0 || select_targets > SMMC_SELECT_MAX_TARGETS)
    return fail(SMMC_ERR_INVALID_ARG, "plan_launch: bad arguments");
  smmc::LaunchInfo li;
  const smmc::Variant variant = current_variant();
  if (!smmc::plan_launch(cfg->mode, cfg->n_returns, cfg->n_periods, cfg->hist_bins, select_targets, has_withdrawals(cfg),
                         uses_sincos_table(cfg->mode, variant, cfg->philox_rounds), cfg->n_simulations, n_sms, (size_t)smem_optin, variant.warps, variant.ticket, &li))
    return fail(SMMC_ERR_UNSUPPORTED, "returns table of %d entries does not fit in %llu bytes of shared memory", cfg->n_returns,
                (unsigned long long)smem_optin);
  plan->grid = li.grid;
  plan->threads = li.threads;
  plan->table_copies = li.rep;
  plan->ticket = li.ticket ? 1 : 0;
  plan->smem_bytes = li.smem_bytes;
  return SMMC_OK;
}

smmc_status smmc_debug_draws(const smmc_config *cfg, int32_t n_paths, uint32_t *outbuf) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  smmc_result dummy;
  memset(&dummy, 0, sizeof dummy);
  smmc_config c2;
  if (cfg) {
    c2 = *cfg;
    c2.n_simulations = n_paths;  // range checks (32-bit path ids, Sobol point indices) apply to the paths asked for
    c2.hist_bins = 0;
  }
  smmc_status st = validate(cfg ? &c2 : nullptr, &dummy);
  if (st != SMMC_OK) return st;
  if (cfg->flags & SMMC_FLAG_DEVICE_POINTERS) return fail(SMMC_ERR_INVALID_ARG, "debug_draws takes host pointers");
  if (n_paths < 0 || !outbuf) return fail(SMMC_ERR_INVALID_ARG, "debug_draws: bad arguments");
  Context *ctx = nullptr;
  st = get_context(cfg->device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(cfg->device));
  smmc::KParams p;
  memset(&p, 0, sizeof p);
  fill_stream_params(cfg, &p);
  const size_t bytes = (size_t)n_paths * (size_t)cfg->n_periods * 4;
  if (bytes == 0) return SMMC_OK;
  st = order_after_last_use(ctx, ctx->stream);  // the Sobol tables rebuilt below may still be read by an asynchronous launch
  if (st != SMMC_OK) return st;
  CU(ctx->scratch.reserve(bytes));
  const uint32_t *full_dev = nullptr, *shift_dev = nullptr;
  if (is_sobol(cfg->mode)) {
    st = prepare_sobol_full(ctx, cfg->n_periods, cfg->seed, &full_dev, &shift_dev);
    if (st != SMMC_OK) return st;
  }
  CU(smmc::launch_debug_draws(cfg->mode, p, n_paths, full_dev, shift_dev, (uint32_t *)ctx->scratch.p, ctx->stream));
  CU(cudaMemcpyAsync(outbuf, ctx->scratch.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return SMMC_OK;
}

smmc_status smmc_debug_philox(const uint32_t *ctr, const uint32_t *key, int32_t n, int32_t device, uint32_t *outbuf) {
  return smmc_debug_philox_rounds(ctr, key, 10, n, device, outbuf);
}

smmc_status smmc_debug_philox_rounds(const uint32_t *ctr, const uint32_t *key, int32_t rounds, int32_t n, int32_t device,
                                     uint32_t *outbuf) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (!ctr || !key || !outbuf || n < 0) return fail(SMMC_ERR_INVALID_ARG, "debug_philox: bad arguments");
  if (rounds != 7 && rounds != 10) return fail(SMMC_ERR_INVALID_ARG, "debug_philox: rounds must be 7 or 10");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  if (n == 0) return SMMC_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  CU(ctx->scratch.reserve((size_t)n * 32));
  uint32_t *dc = (uint32_t *)ctx->scratch.p, *dout = dc + (size_t)n * 4;
  CU(cudaMemcpyAsync(dc, ctr, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
  CU(smmc::launch_debug_philox(dc, make_keys((uint64_t)key[0] | ((uint64_t)key[1] << 32)), rounds, n, dout, ctx->stream));
  CU(cudaMemcpyAsync(outbuf, dout, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return SMMC_OK;
}

smmc_status smmc_debug_set_variant(const char *key, int32_t value) {
  std::lock_guard<std::mutex> lk(g_variant_mu);
  if (!key) {
    g_variant = smmc::Variant();
    return SMMC_OK;
  }
  const smmc::Variant def;
  if (!strcmp(key, "ilp")) g_variant.ilp = value > 0 ? value : def.ilp;
  else if (!strcmp(key, "warps")) g_variant.warps = (value >= 1 && value <= smmc::kWarps) ? value : def.warps;
  else if (!strcmp(key, "reduce")) g_variant.reduce = (value == 1 || value == 2) ? value : def.reduce;
  else if (!strcmp(key, "gauss_tbl")) g_variant.gauss_tbl = (value >= 0 && value <= 15) ? value : def.gauss_tbl;
  else if (!strcmp(key, "sobol_tbl")) g_variant.sobol_tbl = (value >= 0 && value <= 255) ? value : def.sobol_tbl;
  else if (!strcmp(key, "sobol_unroll")) g_variant.sobol_unroll = value > 0 ? value : def.sobol_unroll;
  else if (!strcmp(key, "finals_chunks")) g_variant.finals_chunks = value > 0 ? value : def.finals_chunks;
  else if (!strcmp(key, "select_candidates")) g_variant.select_candidates = (value == 0 || value == 1) ? value : def.select_candidates;
  else if (!strcmp(key, "ticket")) g_variant.ticket = (value == 0 || value == 1) ? value : def.ticket;
  else if (!strcmp(key, "verbose")) g_variant.verbose = value == 1 ? 1 : def.verbose;
  else if (!strcmp(key, "host_merge")) g_variant.host_merge = value == 1 ? 1 : def.host_merge;
  else if (!strcmp(key, "finals_threads")) g_variant.finals_threads = (value > 0 && value <= 64) ? value : def.finals_threads;
  else return fail(SMMC_ERR_INVALID_ARG, "debug_set_variant: unknown key '%s'", key);
  return SMMC_OK;
}

smmc_status smmc_kernel_kinds(int32_t device, int32_t *kinds_out, int32_t max_n, int32_t *n_out) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (!kinds_out || !n_out || max_n < 0) return fail(SMMC_ERR_INVALID_ARG, "kernel_kinds: bad arguments");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  uint64_t have = ctx->ev_head < (uint64_t)kEventRing ? ctx->ev_head : (uint64_t)kEventRing;
  if (have > (uint64_t)max_n) have = (uint64_t)max_n;
  for (uint64_t i = 0; i < have; i++) kinds_out[i] = ctx->ev_kind[(ctx->ev_head - have + i) % kEventRing];  // oldest first
  *n_out = (int32_t)have;
  return SMMC_OK;
}

smmc_status smmc_kernel_times(int32_t device, float *ms_out, int32_t max_n, int32_t *n_out) {
  DeviceGuard device_guard;  // the caller's current device is restored on every return path
  if (!ms_out || !n_out || max_n < 0) return fail(SMMC_ERR_INVALID_ARG, "kernel_times: bad arguments");
  Context *ctx = nullptr;
  smmc_status st = get_context(device, &ctx);
  if (st != SMMC_OK) return st;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(cudaSetDevice(device));
  uint64_t have = ctx->ev_head < (uint64_t)kEventRing ? ctx->ev_head : (uint64_t)kEventRing;
  if (have > (uint64_t)max_n) have = (uint64_t)max_n;
  for (uint64_t i = 0; i < have; i++) {  // oldest first
    const int slot = (int)((ctx->ev_head - have + i) % kEventRing);
    CU(cudaEventSynchronize(ctx->ev1[slot]));
    CU(cudaEventElapsedTime(&ms_out[i], ctx->ev0[slot], ctx->ev1[slot]));
  }
  *n_out = (int32_t)have;
  return SMMC_OK;
}

const char *smmc_last_error(void) { return g_err; }
const char *smmc_version(void) { return "smmc 0.2 (sm_100a, abi 3)"; }
int32_t smmc_device_count(void) { return probe_devices(); }

void smmc_shutdown(void) {
  DeviceGuard device_guard;
  std::lock_guard<std::mutex> lk(g_ctx_mu);
  for (int d = 0; d < kMaxDevices; d++) {
    Context &c = g_ctx[d];
    if (!c.ready) continue;
    cudaSetDevice(d);
    c.returns.release();
    c.sorted.release();
    c.sobol_full.release();
    c.sobol_shift.release();
    c.sobol_full_ready = false;
    c.sincos.release();
    c.sincos_ready = false;
    versioned_release(c.flow_v);
    versioned_release(c.sobol_v);
    if (c.last_use) cudaEventDestroy(c.last_use);
    c.last_use = nullptr;
    c.has_last_use = false;
    for (int i = 0; i < Context::kStageSlots; i++) {
      if (c.stage[i]) cudaFreeHost(c.stage[i]);
      c.stage[i] = nullptr;
    }
    c.stage_cap = 0;
    c.chunk_stats.release();
    c.cand_keys.release();
    c.cand_cursor.release();
    c.cand_valid = false;
    if (c.chunk_stats_host) cudaFreeHost(c.chunk_stats_host);
    c.chunk_stats_host = nullptr;
    c.chunk_stats_host_cap = 0;
    for (int i = 0; i < Context::kChunkDepth; i++) {
      cudaEventDestroy(c.kdone[i]);
      cudaEventDestroy(c.cdone[i]);
      cudaEventDestroy(c.sdone[i]);
    }
    if (c.copy_stream) cudaStreamDestroy(c.copy_stream);
    c.copy_stream = nullptr;
    c.tbl.release();
    c.work.release();
    c.block_moments.release();
    c.stats.release();
    c.finals.release();
    c.bmeans.release();
    c.bvars.release();
    c.scratch.release();
    c.selhist.release();
    if (c.stats_host) cudaFreeHost(c.stats_host);
    c.stats_host = nullptr;
    c.stats_host_cap = 0;
    for (int i = 0; i < kEventRing; i++) {
      cudaEventDestroy(c.ev0[i]);
      cudaEventDestroy(c.ev1[i]);
    }
    cudaEventDestroy(c.red0);
    cudaEventDestroy(c.red1);
    c.ev_head = 0;
    cudaStreamDestroy(c.stream);
    c.ready = false;
  }
  release_multi_comms();
}

}  // extern "C"
