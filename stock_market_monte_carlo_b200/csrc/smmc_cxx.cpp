// smmc_cxx.cpp -- the reference's C++ entry points as marshalling wrappers over the C ABI (include/smmc.h).
// Replaces the "CPP callable function" section of reference src/simulations.cu:657-697 and src/gpu.cpp.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <thread>
#include <vector>

#include "../../include/stock_market_monte_carlo/simulations.h"

namespace {

// reference error behaviour: message on stderr, then exit(code) (gpuErrchk, src/simulations.cu:23-30)
void die_on(smmc_status st, const char *what) {
  if (st == SMMC_OK) return;
  std::fprintf(stderr, "SMMCassert: %s: %s\n", what, smmc_last_error());
  std::exit((int)st);
}

smmc_config base_config(long N, int n_periods, float initial_capital, std::vector<float> &returns,
                        const smmc::Options &opt) {
  smmc_config c;
  std::memset(&c, 0, sizeof c);
  c.abi_version = SMMC_ABI_VERSION;
  c.mode = opt.mode;
  c.n_simulations = N;
  c.n_periods = n_periods;
  c.n_returns = (int)returns.size();  // the reference passes returns.size() as int (src/simulations.cu:670)
  c.returns = returns.empty() ? nullptr : returns.data();
  c.seed = opt.seed;
  c.stream_offset = opt.stream_offset;
  c.initial_capital = initial_capital;
  c.gauss_mean = opt.gauss_mean;
  c.gauss_std = opt.gauss_std;
  c.target = opt.target < 0.0f ? initial_capital : opt.target;
  c.hist_bins = opt.hist_bins;
  c.hist_lo = opt.hist_lo;
  c.hist_hi = opt.hist_hi;
  c.withdraw_fixed = opt.withdraw_fixed;
  c.withdraw_pct = opt.withdraw_pct;
  if (opt.withdraw_pct_schedule) {
    if ((long)opt.withdraw_pct_schedule->size() != (long)n_periods)
      throw std::invalid_argument("withdraw_pct_schedule needs one entry per period");
    c.withdraw_pct_schedule = opt.withdraw_pct_schedule->data();
  }
  return c;
}

// A std::vector<float> of n elements whose storage is NOT value-initialised: the reference's contract is "totals is
// replaced by a fresh N-element vector" (src/simulations.cu:643-644), and every element is about to be overwritten by
// the download, so the 400 MB zero-fill of resize() (10^8 paths: ~40 ms, single-threaded, page faults included) is pure
// overhead.  libstdc++'s vector is three pointers {begin, end, end_of_storage} -- the layout this drop-in is already
// bound to by the reference's mangled signatures -- so after reserve(n) the size is set by moving `end`.
// float is trivially constructible: no constructor is skipped.  Falls back to resize() if the layout check fails.
// Writing the vector's private pointers is outside the language's guarantees, so the shortcut is compiled only for the one
// library it was checked against in its plain configuration: libstdc++ without _GLIBCXX_DEBUG / _GLIBCXX_ASSERTIONS and
// without AddressSanitizer (whose container annotations keep [end, capacity) poisoned after reserve(), so the download
// into it would be reported as a container overflow).  Every other build (libc++, debug or sanitised libstdc++, or
// -DSMMC_PLAIN_RESIZE) takes the zero-filling resize(): ~40 ms slower at 10^8 paths, same vector.
#if defined(__has_feature)
#if __has_feature(address_sanitizer)
#define SMMC_ASAN 1
#endif
#endif
#if defined(__SANITIZE_ADDRESS__)
#define SMMC_ASAN 1
#endif
#if defined(__GLIBCXX__) && !defined(_GLIBCXX_DEBUG) && !defined(_GLIBCXX_ASSERTIONS) && !defined(SMMC_ASAN) && \
    !defined(SMMC_PLAIN_RESIZE)
#define SMMC_RESIZE_SHORTCUT 1
#endif
void resize_uninitialized(std::vector<float> &v, size_t n) {
#ifdef SMMC_RESIZE_SHORTCUT
  struct Rep {
    float *begin, *end, *cap;
  };
  static_assert(sizeof(std::vector<float>) == sizeof(Rep), "libstdc++ std::vector layout expected");
  v.clear();
  v.reserve(n);
  Rep rep;
  std::memcpy(&rep, static_cast<void *>(&v), sizeof rep);
  if (rep.begin == v.data() && rep.end == rep.begin && rep.cap >= rep.begin + n) {
    rep.end = rep.begin + n;
    std::memcpy(static_cast<void *>(&v), &rep, sizeof rep);
  } else {
    v.resize(n);
  }
#else
  v.clear();
  v.resize(n);
#endif
}

void export_stats(const smmc_result &r, smmc::Statistics *s) {
  if (!s) return;
  s->count = r.count;
  s->below_count = r.below_count;
  s->mean = r.mean;
  s->m2 = r.m2;
  s->device_ms = r.device_ms;
}

}  // namespace

void mc_simulations_gpu(std::atomic<long> &n_simulations, long N, int n_periods, float initial_capital,
                        std::vector<float> &returns, std::vector<float> *totals, int n_gpus,
                        const smmc::Options &opt, smmc::Statistics *stats) {
  smmc_config c = base_config(N, n_periods, initial_capital, returns, opt);
  smmc_result r;
  std::memset(&r, 0, sizeof r);
  std::vector<float> fresh;
  if (totals) {
    const size_t n = (size_t)(N > 0 ? N : 0);
    resize_uninitialized(fresh, n);  // pages are first touched by the library's download threads, in parallel
    r.finals = fresh.data();
  }
  if (stats && opt.hist_bins > 0) {
    stats->hist.assign((size_t)opt.hist_bins + 2, 0);
    r.hist = stats->hist.data();
  }
  die_on(smmc_run_multi(&c, n_gpus, &r), "mc_simulations_gpu");
  if (totals) totals->swap(fresh);  // the reference swaps in a fresh N-element vector (src/simulations.cu:643-644)
  export_stats(r, stats);
  n_simulations.store(N, std::memory_order_release);  // results are visible before the progress counter moves
}

void mc_simulations_gpu(std::atomic<long> &n_simulations, long N, int n_periods, float initial_capital,
                        std::vector<float> &returns, std::vector<float> &totals, int n_gpus) {
  mc_simulations_gpu(n_simulations, N, n_periods, initial_capital, returns, &totals, n_gpus, smmc::Options{}, nullptr);
}

void mc_simulations_gpu_reduceBlock(std::atomic<long> &n_simulations, long N, int n_periods, float initial_capital,
                                    std::vector<float> &returns, std::vector<float> *means,
                                    std::vector<float> *variances, int n_gpus, const smmc::Options &opt,
                                    smmc::Statistics *stats) {
  if (n_gpus != 1) throw std::invalid_argument(">1 GPU not yet supported for reduceBlock version!!");
  smmc_config c = base_config(N, n_periods, initial_capital, returns, opt);
  smmc_result r;
  std::memset(&r, 0, sizeof r);
  const size_t nb = N > 0 ? (size_t)((N + SMMC_BLOCK - 1) / SMMC_BLOCK) : 0;
  if (means) {
    means->assign(nb, 0.0f);  // reference: resize + fill 0 (src/simulations.cu:429-432)
    r.block_means = means->data();
  }
  if (variances) {
    variances->assign(nb, 0.0f);
    r.block_vars = variances->data();
  }
  if (stats && opt.hist_bins > 0) {
    stats->hist.assign((size_t)opt.hist_bins + 2, 0);
    r.hist = stats->hist.data();
  }
  die_on(smmc_run(&c, &r), "mc_simulations_gpu_reduceBlock");
  export_stats(r, stats);
  n_simulations.store(N, std::memory_order_release);
}

void mc_simulations_gpu_reduceBlock(std::atomic<long> &n_simulations, long N, int n_periods, float initial_capital,
                                    std::vector<float> &returns, std::vector<float> &means,
                                    std::vector<float> &variances, int n_gpus) {
  mc_simulations_gpu_reduceBlock(n_simulations, N, n_periods, initial_capital, returns, &means, &variances, n_gpus,
                                 smmc::Options{}, nullptr);
}

float reduce_mean_gpu(std::vector<float> &vec, long n) {
  float mean = 0.0f;
  die_on(smmc_reduce_mean(vec.data(), n, 0, &mean), "reduce_mean_gpu");
  return mean;
}

void vector_add(float *out, float *a, float *b, int n) {
  for (int i = 0; i < n; i++) out[i] = a[i] + b[i];
}

void vector_add_gpu(float *out, float *a, float *b, int n) {
  die_on(smmc_vector_add(out, a, b, n, 0), "vector_add_gpu");
}

// ---- SURVEY 8(f) rows ----------------------------------------------------------------------------------------------
namespace smmc {

void update_quartiles(std::vector<float> &quartiles, long N, int n_periods, float initial_capital,
                      std::vector<float> &returns, int n_gpus, const Options &opt) {
  smmc_config c = base_config(N, n_periods, initial_capital, returns, opt);
  c.hist_bins = 0;
  const long q1 = N / 4, q2 = N / 2;  // reference examples/visualize_returns_gpu.cpp:84-86
  const uint64_t ranks[5] = {0, (uint64_t)q1, (uint64_t)q2, (uint64_t)(q1 + q2), (uint64_t)(N - 1)};
  float v[5] = {0, 0, 0, 0, 0};
  die_on(smmc_quantiles(&c, n_gpus, 5, ranks, v, nullptr), "update_quartiles");
  quartiles.assign(v, v + 5);
}

void mc_simulations_keepdata(std::atomic<long> &n_simulations, long N, unsigned int n_periods, float initial_capital,
                             std::vector<float> &historical_returns, std::vector<std::vector<float>> &mc_data,
                             std::vector<float> &final_values, const Options &opt) {
  smmc_config c = base_config(N, (int)n_periods, initial_capital, historical_returns, opt);
  c.hist_bins = 0;
  const size_t row = (size_t)n_periods + 1;
  std::vector<float> flat((size_t)(N > 0 ? N : 0) * row);
  die_on(smmc_trajectories(&c, (int32_t)N, flat.data(), nullptr), "mc_simulations_keepdata");
  mc_data.resize((size_t)(N > 0 ? N : 0));
  final_values.resize((size_t)(N > 0 ? N : 0));
  for (long i = 0; i < N; i++) {
    mc_data[(size_t)i].assign(flat.begin() + (size_t)i * row, flat.begin() + ((size_t)i + 1) * row);
    final_values[(size_t)i] = flat[(size_t)i * row + n_periods];
  }
  n_simulations.store(N, std::memory_order_release);
}

std::vector<float> read_historical_returns(const std::string &csv_fpath) {
  int32_t n = 0;
  if (smmc_read_returns_csv(csv_fpath.c_str(), nullptr, 0, &n) != SMMC_OK) throw std::runtime_error(smmc_last_error());
  std::vector<float> v((size_t)n);
  if (smmc_read_returns_csv(csv_fpath.c_str(), v.data(), n, &n) != SMMC_OK) throw std::runtime_error(smmc_last_error());
  return v;
}

void write_data_file(const std::string &path, std::vector<float> &returns, std::vector<float> &values) {
  if (smmc_write_data_file(path.c_str(), returns.data(), (int32_t)returns.size(), values.data(), (int32_t)values.size()) != SMMC_OK)
    throw std::runtime_error(smmc_last_error());
}

}  // namespace smmc
