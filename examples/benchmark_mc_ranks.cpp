// benchmark_mc_ranks -- the one-rank-per-GPU binding of INTEGRATION.md section 4 as a program: every rank owns one GPU and one
// communicator, passes the SAME configuration of the whole run to smmc_run_rank, and gets the statistics of the whole run back
// (its path range -> kernel -> one ncclAllReduce of the statistics vector on the kernel's stream, all inside libsmmc.so).
// Ranks are host threads here so that the program needs no launcher; with MPI / torchrun each thread body is a process and the
// 128-byte id travels by MPI_Bcast / a store instead of a shared array.  Replaces the device loop of the reference's
// mc_simulations_multi_gpu_launcher_async (src/simulations.cu:576-655), which re-runs the same paths on every device.
//
//   benchmark_mc_ranks <n_ranks> <n_months> <n_simulations> [seed]        (plain C ABI: include/smmc.h only)
//
// build: make -C examples      (links stock_market_monte_carlo_b200/lib/libsmmc.so)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "smmc.h"

int main(int argc, char *argv[]) {
  if (argc < 4) {
    std::printf("usage: benchmark_mc_ranks <n_ranks> <n_months> <n_simulations> [seed], eg benchmark_mc_ranks 2 360 100000000\n");
    return 0;
  }
  const int n_ranks = std::atoi(argv[1]);
  const int n_periods = std::atoi(argv[2]);
  const long long n_simulations = std::strtoll(argv[3], nullptr, 10);
  const unsigned long long seed = argc > 4 ? std::strtoull(argv[4], nullptr, 10) : 1234ull;
  if (n_ranks < 1 || n_ranks > smmc_device_count()) {
    std::fprintf(stderr, "n_ranks %d: this box has %d sm_100a device(s)\n", n_ranks, smmc_device_count());
    return 2;
  }
  unsigned char id[SMMC_COMM_ID_BYTES];
  if (smmc_comm_unique_id(id) != SMMC_OK) {  // "rank 0"; SMMC_ERR_UNSUPPORTED when libnccl.so.2 cannot be loaded
    std::fprintf(stderr, "%s\n", smmc_last_error());
    return 3;
  }
  const int bins = 100;
  struct RankOut {
    smmc_status st = SMMC_OK;
    smmc_result r;
    std::vector<uint64_t> hist;
    float quart[5] = {0, 0, 0, 0, 0};
    char err[256] = "";
  };
  std::vector<RankOut> out((size_t)n_ranks);
  auto rank_main = [&](int rank) {
    RankOut &o = out[(size_t)rank];
    smmc_comm *comm = nullptr;
    o.st = smmc_comm_init_rank(id, n_ranks, rank, /*device=*/rank, &comm);  // collective: ncclCommInitRank
    if (o.st == SMMC_OK) {
      smmc_config c;
      std::memset(&c, 0, sizeof c);
      c.abi_version = SMMC_ABI_VERSION;
      c.mode = SMMC_MODE_GAUSSIAN;
      c.n_simulations = n_simulations;  // the WHOLE run: every rank passes the same configuration
      c.n_periods = n_periods;
      c.initial_capital = 1000.0f;
      c.gauss_mean = 0.5f;
      c.gauss_std = 10.0f / 12.0f;
      c.seed = seed;
      c.target = 1000.0f;
      c.hist_bins = bins;
      c.hist_lo = 0.0f;
      c.hist_hi = 50000.0f;
      o.hist.assign((size_t)bins + 2, (uint64_t)0);
      std::memset(&o.r, 0, sizeof o.r);
      o.r.hist = o.hist.data();
      o.st = smmc_run_rank(&c, &o.r, comm);
      if (o.st == SMMC_OK) {
        const uint64_t N = (uint64_t)n_simulations, ranks[5] = {0, N / 4, N / 2, N / 4 + N / 2, N - 1};
        o.st = smmc_quantiles_rank(&c, comm, 5, ranks, o.quart, nullptr);  // exact order statistics of the sharded run
      }
    }
    if (o.st != SMMC_OK) std::snprintf(o.err, sizeof o.err, "%s", smmc_last_error());  // (the message is per thread)
    if (comm) smmc_comm_destroy(comm);
  };
  std::vector<std::thread> threads;
  for (int r = 0; r < n_ranks; r++) threads.emplace_back(rank_main, r);
  for (auto &t : threads) t.join();
  for (int r = 0; r < n_ranks; r++)
    if (out[(size_t)r].st != SMMC_OK) {
      std::fprintf(stderr, "rank %d: %s\n", r, out[(size_t)r].err);
      return 4;
    }
  // every rank holds the statistics of the whole run: identical counters, histogram and quartiles
  const RankOut &a = out[0];
  for (int r = 1; r < n_ranks; r++) {
    const RankOut &b = out[(size_t)r];
    if (b.r.count != a.r.count || b.r.below_count != a.r.below_count || b.hist != a.hist || b.r.mean != a.r.mean ||
        std::memcmp(b.quart, a.quart, sizeof a.quart) != 0) {
      std::fprintf(stderr, "rank %d disagrees with rank 0\n", r);
      return 5;
    }
  }
  const double sd = std::sqrt(a.r.count ? a.r.m2 / (double)a.r.count : 0.0);
  std::printf("ranks: %d | n_periods: %d | n_simulations: %lld | seed: %llu\n", n_ranks, n_periods, (long long)a.r.count, seed);
  std::printf("mean: %.4f | std: %.4f | below target: %llu\n", a.r.mean, sd, (unsigned long long)a.r.below_count);
  std::printf("quartiles: %.4f %.4f %.4f %.4f %.4f\n", a.quart[0], a.quart[1], a.quart[2], a.quart[3], a.quart[4]);
  std::printf("kernel time of rank 0: %.3f ms (%.3e path-months/s over all ranks)\n", a.r.device_ms,
              (double)a.r.count * n_periods / (a.r.device_ms * 1e-3));
  return 0;
}
