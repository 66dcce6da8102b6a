// extern "C" drivers around the UNMODIFIED reference GPU library (src/simulations.cu compiled for sm_100a),
// built into oracle/_ref/libref_gpu.so.  TEST INFRASTRUCTURE: the "reference kernel on the same box" that
// pins BOOTSTRAP_COMPAT bit-for-bit and serves as the same-box GPU baseline.
// The reference kernels hold the table in a fixed __shared__ float[1127] (src/simulations.cu:123, :193):
// n_returns must be <= 1127.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

void mc_simulations_gpu(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                        float initial_capital, std::vector<float> &returns, std::vector<float> &totals, int n_gpus);
void mc_simulations_gpu_reduceBlock(std::atomic<long> &n_simulations, long max_n_simulations, int n_periods,
                                    float initial_capital, std::vector<float> &returns, std::vector<float> &means,
                                    std::vector<float> &variances, int n_gpus);

namespace {
// The reference library prints its own timing lines to stdout (src/simulations.cu:401, :463: "GPU kernel itself took X s!",
// host steady_clock in whole milliseconds).  They are captured into a scratch file instead of reaching the caller's
// stdout, and the kernel line is kept for ref_last_kernel_seconds().
double g_last_kernel_s = -1.0;
struct CaptureStdout {
  int saved = -1;
  FILE *tmp = nullptr;
  CaptureStdout() {
    fflush(stdout);
    saved = dup(1);
    tmp = tmpfile();
    if (tmp) {
      dup2(fileno(tmp), 1);
    } else {
      int devnull = open("/dev/null", O_WRONLY);
      dup2(devnull, 1);
      close(devnull);
    }
  }
  ~CaptureStdout() {
    fflush(stdout);
    dup2(saved, 1);
    close(saved);
    g_last_kernel_s = -1.0;
    if (tmp) {
      rewind(tmp);
      char line[512];
      while (fgets(line, sizeof line, tmp)) {
        double v = 0.0;
        const char *p = strstr(line, "GPU kernel itself took ");
        if (p && sscanf(p, "GPU kernel itself took %lf s", &v) == 1) g_last_kernel_s = v;
      }
      fclose(tmp);
    }
  }
};
typedef CaptureStdout QuietStdout;
}  // namespace

extern "C" {
// seconds of the last "GPU kernel itself took" line the reference printed (1 ms resolution), -1 if none
double ref_last_kernel_seconds(void) { return g_last_kernel_s; }

int ref_mc_simulations_gpu(long N, int n_periods, float initial_capital, const float *returns, int n_returns,
                           float *totals, int n_gpus) {
  if (n_returns > 1127) return 1;
  QuietStdout q;
  std::vector<float> r(returns, returns + n_returns), t;
  std::atomic<long> n_sim{0};
  mc_simulations_gpu(n_sim, N, n_periods, initial_capital, r, t, n_gpus);
  for (long i = 0; i < N && i < (long)t.size(); i++) totals[i] = t[i];
  return n_sim.load() == N ? 0 : 2;
}

int ref_mc_simulations_gpu_reduceBlock(long N, int n_periods, float initial_capital, const float *returns,
                                       int n_returns, float *means, float *variances) {
  if (n_returns > 1127) return 1;
  QuietStdout q;
  std::vector<float> r(returns, returns + n_returns), m, v;
  std::atomic<long> n_sim{0};
  mc_simulations_gpu_reduceBlock(n_sim, N, n_periods, initial_capital, r, m, v, 1);
  for (size_t i = 0; i < m.size(); i++) {
    means[i] = m[i];
    variances[i] = v[i];
  }
  return n_sim.load() == N ? 0 : 2;
}
}
