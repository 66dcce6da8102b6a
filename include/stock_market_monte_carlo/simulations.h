// B200-native drop-in for the GPU half of the reference's public header
// (reference include/stock_market_monte_carlo/simulations.h:69-86).
//
// The five functions below keep the reference's names, argument order, argument meaning and result layout, and
// are exported (Itanium-mangled) by libsmmc.so, so the reference's programs (examples/benchmark_mc_gpu.cpp:69,
// benchmark_mc_gpu_reduceBlock.cpp:97, benchmark_reduce_mean.cpp:38, example_gpu.cpp:29-32,
// visualize_returns_gpu.cpp:175) relink against it unchanged.  They are marshalling wrappers over the C ABI in
// include/smmc.h; everything new (seed, RNG mode, Gaussian / Sobol returns, fused statistics, histogram,
// P(target)) is reached through `smmc::Options` / `smmc::Statistics` overloads further down.
//
// Unlike the reference header this one includes what it uses (<atomic>, <vector>, <cstdint>).
#pragma once

#include <atomic>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../smmc.h"

// ---- reference signatures -----------------------------------------------------------------------------------

// All N final values, one per path: totals[i] is path i (reference src/simulations.cu:661-680).
//   returns        historical monthly returns in PERCENT, bootstrap-resampled (reference RNG stream, bit for bit)
//   totals         replaced by a fresh N-element vector
//   n_gpus         devices 0..n_gpus-1 each simulate a disjoint path range (the reference re-ran the same
//                  paths on every device, src/simulations.cu:621-622; here totals[] is independent of n_gpus)
//   n_simulations  stored = N after the results are visible (callers poll it from another thread)
// CUDA errors print to stderr and exit(code), like the reference's gpuErrchk (src/simulations.cu:23-30).
void mc_simulations_gpu(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                        float initial_capital, std::vector<float> &returns, std::vector<float> &totals, int n_gpus);

// Per-256-path-block mean and population variance (reference src/simulations.cu:682-697):
// means[b], variances[b] cover paths [256 b, 256 b + 255]; both vectors are resized to ceil(N/256).
// Throws std::invalid_argument unless n_gpus == 1, as the reference does (:690-694).
void mc_simulations_gpu_reduceBlock(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                                    float initial_capital, std::vector<float> &returns, std::vector<float> &means,
                                    std::vector<float> &variances, int n_gpus);

// Mean of the first n elements of vec, reduced on the GPU (reference src/simulations.cu:269-341).
float reduce_mean_gpu(std::vector<float> &vec, long n);

// reference include/stock_market_monte_carlo/gpu.h:1-2 (also declared at the end of its simulations.h)
void vector_add(float *out, float *a, float *b, int n);
void vector_add_gpu(float *out, float *a, float *b, int n);

// ---- extensions: seed / mode / fused statistics ----------------------------------------------------------------
namespace smmc {

struct Options {
  smmc_mode mode = SMMC_MODE_BOOTSTRAP_COMPAT;
  std::uint64_t seed = 0;
  std::uint64_t stream_offset = 0;
  float gauss_mean = 0.5f;            // percent per period (reference examples/monte_carlo_simulated.cpp:11)
  float gauss_std = 10.0f / 12.0f;    // percent per period (:12)
  float target = -1.0f;               // < 0: use initial_capital (reference examples/benchmark_mc_gpu.cpp:79)
  int hist_bins = 0;
  float hist_lo = 0.0f, hist_hi = 0.0f;
  // withdrawals after every period (reference README.md:130-133, unimplemented there): value <- max(value *
  // (1 - pct/100) - withdraw_fixed, 0); withdraw_pct_schedule (one percent per period) overrides withdraw_pct
  float withdraw_fixed = 0.0f;
  float withdraw_pct = 0.0f;
  const std::vector<float> *withdraw_pct_schedule = nullptr;
};

struct Statistics {
  std::uint64_t count = 0, below_count = 0;
  double mean = 0.0, m2 = 0.0;        // population variance = m2 / count
  std::vector<std::uint64_t> hist;    // hist_bins + 2 counters (under, bins..., over)
  float device_ms = 0.0f;
};

}  // namespace smmc

// Same contract as the reference-signature functions, plus options in and fused statistics out.  `totals`,
// `means`, `variances` may be nullptr: nothing per-path (or per-block) is then written to HBM or copied back.
void mc_simulations_gpu(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                        float initial_capital, std::vector<float> &returns, std::vector<float> *totals, int n_gpus,
                        const smmc::Options &opt, smmc::Statistics *stats);
void mc_simulations_gpu_reduceBlock(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                                    float initial_capital, std::vector<float> &returns, std::vector<float> *means,
                                    std::vector<float> *variances, int n_gpus, const smmc::Options &opt,
                                    smmc::Statistics *stats);

// ---- what the reference's callers do on the host around the GPU entry points ------------------
// (in namespace smmc: the reference defines CPU functions of the same names in the global namespace, and its
// programs keep linking those from their own src/simulations.cpp / src/helpers.cpp)
namespace smmc {

// quartiles <- {min, Q1, median, Q3, max} of the N finals, i.e. sorted[0], sorted[N/4], sorted[N/2],
// sorted[N/4 + N/2], sorted[N-1] exactly as update_quartiles computes them after a full host sort
// (reference examples/visualize_returns_gpu.cpp:83-110) -- by three replays of the counter-based simulation
// driving a radix select on the device; the finals are never materialised.
void update_quartiles(std::vector<float> &quartiles, long max_n_simulations, int n_periods, float initial_capital,
                      std::vector<float> &returns, int n_gpus = 1, const Options &opt = Options{});

// mc_data[i] <- the n_periods + 1 values of path i, final_values[i] <- the last one (reference
// mc_simulations_keepdata, src/simulations.cpp:139-202), from the same RNG streams as mc_simulations_gpu.
void mc_simulations_keepdata(std::atomic<long> &n_simulations, long max_n_simulations, unsigned int n_periods,
                             float initial_capital, std::vector<float> &historical_returns,
                             std::vector<std::vector<float>> &mc_data, std::vector<float> &final_values,
                             const Options &opt = Options{});

// reference read_historical_returns (src/simulations.cpp:83-93) / write_data_file (src/helpers.cpp:23-39; the
// reference prepends "./outputs/", here `path` is used as given).  Throw std::runtime_error on failure.
std::vector<float> read_historical_returns(const std::string &csv_fpath);
void write_data_file(const std::string &path, std::vector<float> &returns, std::vector<float> &values);

}  // namespace smmc
