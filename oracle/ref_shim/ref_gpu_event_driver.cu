// The reference's two simulation kernels, UNMODIFIED, timed with CUDA events instead of the reference's own host
// steady_clock print (whole milliseconds, src/simulations.cu:456-463).  This file textually includes the reference source
// from where it lies under /root/reference (-I$(REF)) and launches
//   mc_simulations_gpu_kernel_reduceBlock<256>   (src/simulations.cu:185-247)
//   mc_simulations_gpu_kernel<256>               (src/simulations.cu:112-152)
// with the launch geometry of the reference's launchers (:363-366, :421-425: 256 threads, ceil(N / 256) blocks) between an
// event pair on the default stream.  Built into oracle/_ref/libref_gpu_events.so.  TEST / BASELINE INFRASTRUCTURE: nothing in
// the product links or calls this.
#include "src/simulations.cu"

extern "C" {
// which = 0: reduceBlock kernel (per-block mean / variance), 1: finals kernel (one float per path).
// Returns the mean CUDA-event duration in ms of `reps` launches after one warm-up launch, or a negative error code.
float ref_kernel_event_ms(int which, long N, int n_periods, float initial_capital, const float *returns, int n_returns, int reps) {
  if (n_returns > 1127 || N <= 0 || reps < 1) return -1.0f;
  const int block_size = THREADS_PER_BLOCK;
  const unsigned int n_blocks = (unsigned int)((N + block_size - 1) / block_size);
  float *returns_d = nullptr, *a_d = nullptr, *b_d = nullptr;
  if (cudaMalloc(&returns_d, n_returns * sizeof(float)) != cudaSuccess) return -2.0f;
  const size_t out_elems = which == 0 ? (size_t)n_blocks : (size_t)N;
  if (cudaMalloc(&a_d, out_elems * sizeof(float)) != cudaSuccess) return -2.0f;
  if (which == 0 && cudaMalloc(&b_d, out_elems * sizeof(float)) != cudaSuccess) return -2.0f;
  cudaMemcpy(returns_d, returns, n_returns * sizeof(float), cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double total = 0.0;
  for (int r = 0; r <= reps; r++) {
    if (which == 0) {
      cudaMemset(a_d, 0, out_elems * sizeof(float));
      cudaMemset(b_d, 0, out_elems * sizeof(float));
    }
    cudaEventRecord(e0);
    if (which == 0)
      mc_simulations_gpu_kernel_reduceBlock<block_size><<<n_blocks, block_size>>>(returns_d, n_returns, a_d, b_d, N, initial_capital, n_periods);
    else
      mc_simulations_gpu_kernel<block_size><<<n_blocks, block_size>>>(returns_d, n_returns, a_d, N, initial_capital, n_periods);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
      total = -3.0 * reps;
      break;
    }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (r > 0) total += ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(returns_d);
  cudaFree(a_d);
  if (b_d) cudaFree(b_d);
  return (float)(total / reps);
}
}
