// extern "C" drivers around the UNMODIFIED reference CPU sources (src/simulations.cpp, src/helpers.cpp),
// compiled where they lie into oracle/_ref/libref_cpu.so.  TEST INFRASTRUCTURE: used to pin the oracle
// (tests/) and as the timed CPU baseline of bench.py; never linked into the product.
#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>
#include <omp.h>
#include <unistd.h>
#include <fcntl.h>

#include "stock_market_monte_carlo/simulations.h"

// defined in src/simulations.cpp:24 with an `unsigned int` third parameter (the header declares `long`)
std::vector<float> many_updates(float fund_value, std::vector<float> &returns, unsigned int n_periods);

namespace {
// the reference prints progress from inside its library functions; keep test logs readable
struct QuietStdout {
  int saved = -1;
  explicit QuietStdout(bool on) {
    if (!on) return;
    fflush(stdout);
    saved = dup(1);
    int devnull = open("/dev/null", O_WRONLY);
    dup2(devnull, 1);
    close(devnull);
  }
  ~QuietStdout() {
    if (saved < 0) return;
    fflush(stdout);
    dup2(saved, 1);
    close(saved);
  }
};
}  // namespace

extern "C" {

float ref_update_fund(float fund, float ret) { return update_fund(fund, ret); }

// many_updates (src/simulations.cpp:24-39): out[0..n_periods]
void ref_many_updates(float fund, const float *returns, unsigned n_periods, float *out) {
  std::vector<float> r(returns, returns + n_periods);
  std::vector<float> v = many_updates(fund, r, n_periods);
  for (unsigned i = 0; i <= n_periods; i++) out[i] = v[i];
}

// CPU v2 OpenMP bootstrap (src/simulations.cpp:204-266).  N must be >= 10000 (the reference divides by
// n_blocks/10, :255).  Returns the thread count the reference chose (hardware_concurrency()-1, :218-219).
int ref_mc_simulations(long N, unsigned n_periods, float initial_capital, const float *returns, int n_returns,
                       float *finals, int quiet) {
  QuietStdout q(quiet != 0);
  std::vector<float> hist(returns, returns + n_returns);
  std::vector<float> fin((size_t)N, initial_capital);
  std::atomic<long> n_sim{0};
  mc_simulations(n_sim, N, n_periods, initial_capital, hist, fin);
  for (long i = 0; i < N; i++) finals[i] = fin[i];
  int t = (int)std::thread::hardware_concurrency() - 1;
  return t < 1 ? 1 : t;
}

// Gaussian CPU path: sample_returns_gaussian + many_updates per path (src/simulations.cpp:41-55, :24-39, as
// composed by one_simulation_gaussian :57-66), under OpenMP with the reference's own thread-count rule.
int ref_gaussian_finals(long N, unsigned n_periods, float initial_capital, float mean_pct, float std_pct,
                        float *finals) {
  int t = (int)std::thread::hardware_concurrency() - 1;
  if (t < 1) t = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(t)
  for (long i = 0; i < N; i++) {
    std::vector<float> r = sample_returns_gaussian(n_periods, mean_pct, std_pct);
    std::vector<float> v = many_updates(initial_capital, r, n_periods);
    finals[i] = v[n_periods];
  }
  return t;
}

// read_historical_returns (src/simulations.cpp:83-93); returns the row count (<= max_n rows copied)
long ref_read_historical_returns(const char *path, float *out, long max_n) {
  std::vector<float> v = read_historical_returns(path);
  long n = (long)v.size();
  for (long i = 0; i < n && i < max_n; i++) out[i] = v[i];
  return n;
}
}

// write_data_file (src/helpers.cpp:23-39) writes ./outputs/<fname> relative to the working directory
#include "stock_market_monte_carlo/helpers.h"
extern "C" void ref_write_data_file(const char *fname, const float *returns, int n_returns, const float *values,
                                    int n_values) {
  QuietStdout q(true);
  std::vector<float> r(returns, returns + n_returns), v(values, values + n_values);
  write_data_file(fname, r, v);
}
