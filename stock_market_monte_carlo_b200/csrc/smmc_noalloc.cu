// smmc_noalloc.cu -- smmc_reduce_mean_device_noalloc: the mean of a device-resident float vector for callers that cannot
// let this library allocate device memory.  No counterpart in the reference.
//
// Who needs it.  Every other compute entry point of libsmmc builds a per-device Context on first use and grows cudaMalloc'ed
// work buffers on demand (smmc_api.cu: DevBuf::reserve).  That is wrong for three kinds of hosts this library is linked into:
//   * a pricing / risk service that hands ALL of a GPU's memory to its own arena at start-up (a pool allocator, or a
//     framework's caching allocator with the whole device reserved): a cudaMalloc from inside a plugin then fails with
//     out-of-memory although the service has plenty of free memory of its own;
//   * a host that keeps other kernels in flight on other streams while it reduces a `finals` buffer: cudaMalloc and
//     cudaFree synchronise the whole device, so an allocating call stalls everything the host has queued;
//   * code running under a memory budget fixed at deployment (MIG slices, several tenants per GPU), where what a library
//     takes must be known when it is loaded and not depend on the sizes of later calls.
// For them the scratch of the reduction -- one fp64 partial per 32768 input floats, a ticket and the sum -- lives in a
// file-scope __device__ array of FIXED size (262 400 bytes).  Every GPU on which this library's module gets loaded holds
// that array until the process exits; nothing is allocated or freed by a call.
//
// Rules that follow from sharing ONE array per device:
//   * n <= SMMC_NOALLOC_MAX_N (2^30 floats, 4 GiB); longer vectors are refused with SMMC_ERR_INVALID_ARG.
//   * The array keeps what the last call left in it.  Every call clears the 256-byte header (ticket, sum) on its stream
//     before its kernel; the partials need no clearing because every slot that is read was written by the same launch.
//   * Calls whose GPU work overlaps (two host threads, two streams) on one device race on the array and return garbage.
//     Calls are serialised per process by a mutex and each one blocks until its result is on the host, so the entry point
//     by itself never overlaps itself; a second PROCESS has its own copy of the array.
//   * The first call on a device looks up the array's device address (cudaGetSymbolAddress, which may load the module).
//     Make that first call before any stream capture begins.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>

#include "../../include/smmc.h"

namespace smmc {
smmc_status set_error(smmc_status st, const char *fmt, ...);  // smmc_api.cu
}

namespace {

constexpr int kThreads = 512;
constexpr int64_t kChunkFloats = 32768;  // one partial per 128 KiB of input
constexpr int64_t kMaxPartials = 32768;
constexpr size_t kHeaderBytes = 256;     // [0,8) sum (double), [8,12) ticket (u32), rest unused
constexpr size_t kScratchBytes = kHeaderBytes + (size_t)kMaxPartials * sizeof(double);
static_assert(kChunkFloats * kMaxPartials == SMMC_NOALLOC_MAX_N, "smmc.h states the limit this file implements");
constexpr int kMaxDevices = 64;

// The scratch itself.  In HOST code `g_noalloc_scratch` is only a handle for cudaGetSymbolAddress: passing it (or its
// address) to a kernel would hand the kernel a host address.  scratch_of() below resolves it once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

struct Header {
  double sum;
  unsigned int ticket;
};

__device__ __forceinline__ double block_sum(double acc, double *ws) {
#pragma unroll
  for (int st = 16; st >= 1; st >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, st);
  __syncthreads();  // ws may still be read by the previous use
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kThreads / 32; w++) t += ws[w];
  return t;  // valid in thread 0
}

// Sum of n floats, fp64 accumulation.  The vector is cut at its first 16-byte boundary: `head` (< 4) leading floats go to
// chunk 0, the aligned rest is split into chunks of kChunkFloats that the CTAs take round-robin; chunk c leaves its sum in
// partials[c].  The CTA that finishes last adds the partials in index order -- the bits of the result depend on (pointer
// alignment, n) only, not on the grid or on which CTA ran when.  HBM-bound at 4 bytes per element.
__global__ void __launch_bounds__(kThreads) reduce_sum_noalloc_kernel(const float *__restrict__ v, int64_t n, int64_t n_chunks,
                                                                      unsigned char *scratch) {
  __shared__ double ws[kThreads / 32];
  __shared__ bool last;
  Header *hdr = reinterpret_cast<Header *>(scratch);
  double *partials = reinterpret_cast<double *>(scratch + kHeaderBytes);
  const int tid = threadIdx.x;
  int64_t head = (int64_t)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(v) & 15u)) & 15u) >> 2);
  if (head > n) head = n;
  const float *mid = v + head;
  const int64_t nm = n - head;
  for (int64_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const int64_t lo = c * kChunkFloats;
    const int64_t len = (nm - lo < kChunkFloats) ? (nm - lo) : kChunkFloats;  // >= 0
    const float4 *q4 = reinterpret_cast<const float4 *>(mid + lo);
    const int n4 = (int)(len >> 2);
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int i = tid;
    for (; i + 3 * kThreads < n4; i += 4 * kThreads) {
      const float4 q0 = __ldcs(q4 + i), q1 = __ldcs(q4 + i + kThreads), q2 = __ldcs(q4 + i + 2 * kThreads),
                   q3 = __ldcs(q4 + i + 3 * kThreads);
      a0 += ((double)q0.x + (double)q0.y) + ((double)q0.z + (double)q0.w);
      a1 += ((double)q1.x + (double)q1.y) + ((double)q1.z + (double)q1.w);
      a2 += ((double)q2.x + (double)q2.y) + ((double)q2.z + (double)q2.w);
      a3 += ((double)q3.x + (double)q3.y) + ((double)q3.z + (double)q3.w);
    }
    for (; i < n4; i += kThreads) {
      const float4 q = __ldcs(q4 + i);
      a0 += ((double)q.x + (double)q.y) + ((double)q.z + (double)q.w);
    }
    double acc = (a0 + a1) + (a2 + a3);
    if (tid < (int)(len & 3)) acc += (double)mid[lo + ((int64_t)n4 << 2) + tid];  // the last chunk's ragged end
    if (c == 0 && tid < head) acc += (double)v[tid];
    const double t = block_sum(acc, ws);
    if (tid == 0) partials[c] = t;
  }
  if (tid == 0) {
    __threadfence();
    last = (atomicAdd(&hdr->ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double acc = 0.0;
  for (int64_t c = tid; c < n_chunks; c += kThreads) acc += __ldcg(partials + c);
  const double t = block_sum(acc, ws);
  if (tid == 0) {
    hdr->sum = t;
    hdr->ticket = 0u;  // the next call finds a clean ticket even if its own clearing were dropped
  }
}

// Per-device set-up, done once: architecture check, SM count, and the DEVICE address of g_noalloc_scratch.
struct DeviceSlot {
  bool ready = false;
  int n_sms = 0;
  unsigned char *scratch = nullptr;
};
DeviceSlot g_slot[kMaxDevices];
std::mutex g_mu;  // guards g_slot and serialises the calls (they share the array)

struct RestoreDevice {
  int prev = -1;
  RestoreDevice() {
    if (cudaGetDevice(&prev) != cudaSuccess) {
      cudaGetLastError();
      prev = -1;
    }
  }
  ~RestoreDevice() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

#define NA_CU(call)                                                                                                  \
  do {                                                                                                               \
    cudaError_t e__ = (call);                                                                                        \
    if (e__ != cudaSuccess)                                                                                          \
      return smmc::set_error(SMMC_ERR_CUDA, "reduce_mean_device_noalloc: %s failed: %s (%s:%d)", #call,            \
                             cudaGetErrorString(e__), __FILE__, __LINE__);                                           \
  } while (0)

// `device` is current.  The symbol's address differs per device, hence one look-up per device.
smmc_status scratch_of(int device, DeviceSlot **out) {
  DeviceSlot &s = g_slot[device];
  if (!s.ready) {
    int major = 0, sms = 0;
    NA_CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10)
      return smmc::set_error(SMMC_ERR_NO_DEVICE, "reduce_mean_device_noalloc: device %d is not compute capability 10.x", device);
    NA_CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    void *p = nullptr;
    NA_CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255u))
      return smmc::set_error(SMMC_ERR_CUDA, "reduce_mean_device_noalloc: static scratch resolved to %p on device %d", p, device);
    s.scratch = static_cast<unsigned char *>(p);
    s.n_sms = sms;
    s.ready = true;
  }
  *out = &s;
  return SMMC_OK;
}

}  // namespace

extern "C" smmc_status smmc_reduce_mean_device_noalloc(const float *vec_dev, int64_t n, int32_t device, void *stream,
                                                       float *mean_out) {
  if (!vec_dev || !mean_out || n <= 0)
    return smmc::set_error(SMMC_ERR_INVALID_ARG, "reduce_mean_device_noalloc: need n > 0 and non-null pointers");
  if (reinterpret_cast<uintptr_t>(vec_dev) & 3u)
    return smmc::set_error(SMMC_ERR_INVALID_ARG, "reduce_mean_device_noalloc: pointer must be 4-byte aligned");
  if (n > SMMC_NOALLOC_MAX_N)
    return smmc::set_error(SMMC_ERR_INVALID_ARG,
                           "reduce_mean_device_noalloc: n = %lld exceeds the fixed scratch of this entry point (at most %lld "
                           "floats); use smmc_reduce_mean_device, which allocates, or reduce the vector in pieces",
                           (long long)n, (long long)SMMC_NOALLOC_MAX_N);
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
    cudaGetLastError();
    return smmc::set_error(SMMC_ERR_NO_DEVICE, "reduce_mean_device_noalloc: no CUDA device is visible; this library has no CPU fallback");
  }
  if (device < 0 || device >= n_dev || device >= kMaxDevices)
    return smmc::set_error(SMMC_ERR_INVALID_ARG, "reduce_mean_device_noalloc: device %d out of range [0,%d)", device,
                           n_dev < kMaxDevices ? n_dev : kMaxDevices);
  RestoreDevice restore;  // the caller's current device comes back on every return path
  std::lock_guard<std::mutex> lk(g_mu);
  NA_CU(cudaSetDevice(device));
  DeviceSlot *slot = nullptr;
  smmc_status rc = scratch_of(device, &slot);
  if (rc != SMMC_OK) return rc;

  int64_t head = (int64_t)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(vec_dev) & 15u)) & 15u) >> 2);
  if (head > n) head = n;
  int64_t n_chunks = (n - head + kChunkFloats - 1) / kChunkFloats;
  if (n_chunks < 1) n_chunks = 1;  // fewer than 4 floats, all in the head
  const int64_t max_grid = (int64_t)slot->n_sms * 4;
  const int grid = (int)(n_chunks < max_grid ? n_chunks : max_grid);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  NA_CU(cudaMemsetAsync(slot->scratch, 0, kHeaderBytes, st));
  reduce_sum_noalloc_kernel<<<grid, kThreads, 0, st>>>(vec_dev, n, n_chunks, slot->scratch);
  NA_CU(cudaGetLastError());
  double sum = 0.0;
  NA_CU(cudaMemcpyAsync(&sum, slot->scratch, sizeof(double), cudaMemcpyDeviceToHost, st));
  NA_CU(cudaStreamSynchronize(st));
  *mean_out = (float)(sum / (double)n);
  return SMMC_OK;
}
