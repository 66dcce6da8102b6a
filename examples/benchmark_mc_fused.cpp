// benchmark_mc_fused -- the reference's GPU benchmark program (examples/benchmark_mc_gpu.cpp) rewritten against the
// fused entry points: same command line, same CSV, same printed statistics, but nothing per-path leaves the GPU --
// mean / std / count-below-target come out of the simulation kernel, the quartiles out of three replays.
//
//   benchmark_mc_fused <n_gpus> <n_months> <n_simulations> [mode] [seed]
//     mode: compat (default: the reference's RNG stream, identical numbers to benchmark_mc_gpu) | bootstrap |
//           gaussian | sobol_bootstrap | sobol_gaussian
//   reads data/SP500_monthly_returns.csv (column "returns", percent) like the reference (examples/benchmark_mc_gpu.cpp:64)
//
// build: make -C examples      (links stock_market_monte_carlo_b200/lib/libsmmc.so)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "stock_market_monte_carlo/simulations.h"

int main(int argc, char *argv[]) {
  if (argc < 4) {
    std::printf("usage: benchmark_mc_fused <n_gpus> <n_months> <n_simulations> [mode] [seed], eg benchmark_mc_fused 1 360 100000\n");
    return 0;
  }
  const int n_gpus = std::atoi(argv[1]);
  const int n_periods = std::atoi(argv[2]);
  const long max_n_simulations = std::strtol(argv[3], nullptr, 10);
  const std::string mode = argc > 4 ? argv[4] : "compat";
  smmc::Options opt;
  opt.seed = argc > 5 ? std::strtoull(argv[5], nullptr, 10) : 0;
  if (mode == "compat") opt.mode = SMMC_MODE_BOOTSTRAP_COMPAT;
  else if (mode == "bootstrap") opt.mode = SMMC_MODE_BOOTSTRAP;
  else if (mode == "gaussian") opt.mode = SMMC_MODE_GAUSSIAN;
  else if (mode == "sobol_bootstrap") opt.mode = SMMC_MODE_SOBOL_BOOTSTRAP;
  else if (mode == "sobol_gaussian") opt.mode = SMMC_MODE_SOBOL_GAUSSIAN;
  else {
    std::fprintf(stderr, "unknown mode %s\n", mode.c_str());
    return 1;
  }
  std::printf("n_periods: %d | max_n_simulations: %ld | mode: %s | seed: %llu\n", n_periods, max_n_simulations, mode.c_str(),
              (unsigned long long)opt.seed);

  const float initial_capital = 1000;
  std::vector<float> historical_returns;
  if (opt.mode != SMMC_MODE_GAUSSIAN && opt.mode != SMMC_MODE_SOBOL_GAUSSIAN)
    historical_returns = smmc::read_historical_returns("data/SP500_monthly_returns.csv");
  std::atomic<long> n_simulations{0};
  smmc::Statistics stats;
  std::vector<float> quartiles;

  const auto begin = std::chrono::steady_clock::now();
  mc_simulations_gpu(n_simulations, max_n_simulations, n_periods, initial_capital, historical_returns, nullptr, n_gpus, opt, &stats);
  smmc::update_quartiles(quartiles, max_n_simulations, n_periods, initial_capital, historical_returns, n_gpus, opt);
  const auto end = std::chrono::steady_clock::now();
  const double secs = std::chrono::duration_cast<std::chrono::microseconds>(end - begin).count() * 1e-6;
  std::printf("All %ld simulation done in %.6f s!\n", n_simulations.load(), secs);

  // the statistics the reference program computes on the host from the N finals (examples/benchmark_mc_gpu.cpp:7-41)
  const double mean = stats.mean, sd = std::sqrt(stats.count ? stats.m2 / (double)stats.count : 0.0);
  std::printf("mean: %.2f | std: %.2f \n", mean, sd);
  std::printf("count_below %.1f: %llu (%4f%%)\n", initial_capital, (unsigned long long)stats.below_count,
              100.0 * (double)stats.below_count / (double)max_n_simulations);
  std::printf("min: %.2f | Q1: %.2f | median: %.2f | Q3: %.2f | max: %.2f\n", quartiles[0], quartiles[1], quartiles[2],
              quartiles[3], quartiles[4]);
  return 0;
}
