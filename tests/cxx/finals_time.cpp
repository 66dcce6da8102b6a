// Warm-process timing of the drop-in C++ entry point mc_simulations_gpu (all N finals, reference signature):
//   g++ -std=c++17 -O2 -I../../include finals_time.cpp -L<libdir> -lsmmc -Wl,-rpath,<libdir>     (tests/run_cxx_finals_time.py)
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "stock_market_monte_carlo/simulations.h"

int main(int argc, char **argv) {
  const long N = argc > 1 ? atol(argv[1]) : 100000000L;
  const int P = argc > 2 ? atoi(argv[2]) : 360;
  std::vector<float> returns(1127);
  for (int i = 0; i < 1127; i++) returns[i] = 0.6f + 4.0f * (float)((i * 7919) % 200 - 100) / 100.0f;
  std::atomic<long> n(0);
  for (int it = 0; it < 4; it++) {
    std::vector<float> totals;  // passed empty, as the reference's benchmark does (examples/benchmark_mc_gpu.cpp:65)
    const auto t0 = std::chrono::steady_clock::now();
    mc_simulations_gpu(n, N, P, 1000.0f, returns, totals, 1);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::printf("call %d: %.2f ms  (%ld finals, totals[0] = %.3f, totals[N-1] = %.3f)\n", it, ms, (long)totals.size(), totals[0],
                totals[N - 1]);
  }
  return 0;
}
