// Gaussian inner-loop variants (sm_100a), development tool: the shipped Box-Muller pair (4 MUFU) against a pair whose
// sin/cos come from a shared-memory table of 16384 (cos, sin) cell centres + a first-order rotation (2 MUFU + LDS.64).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../stock_market_monte_carlo_b200/csrc -o gauss_variants gauss_variants.cu
#include <cmath>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#include "smmc_device.cuh"

using namespace smmc;

constexpr int kCells = 16384;

__device__ __forceinline__ void gauss_pair_tbl(uint32_t w, uint32_t one, const GaussConsts gc, uint32_t tbl_base, float kap,
                                               float kap0, float &g0, float &g1) {
  const float mr = __uint_as_float((w & 0xFFFFFu) | one);
  const float u = fmaf(mr, 8.0f, -7.99999952316284179688f);
  const uint32_t h = w * kAngleHash;
#ifdef CONFLICT_FREE  // emulates a bank-conflict-free gather (wrong values, same instruction mix): lane-private 8-byte column
  const uint32_t addr = ((h >> 15) & 0x1FF00u) + tbl_base;
#else
  const uint32_t addr = ((h >> 15) & 0x1FFF8u) + tbl_base;
#endif
  float c, s;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c), "=f"(s) : "r"(addr));
  const float beta = fmaf(__uint_as_float((h & 0x3FFFFu) | one), kap, kap0);
  const float t = lg2_approx(u) * gc.k;
  const float rs = sqrt_approx(t);
  const float cv = fmaf(-s, beta, c), sv = fmaf(c, beta, s);
  g0 = fmaf(rs, cv, gc.a);
  g1 = fmaf(rs, sv, gc.a);
}

template <int VAR, int ILP>
__global__ void __launch_bounds__(1024, 1) k(float *out, const float2 *tbl_g, int nq, PhiloxKeys keys, GaussConsts gc,
                                             uint32_t one, float kap, float kap0, long long *cyc) {
  extern __shared__ __align__(16) unsigned char smem[];
  float2 *tbl = reinterpret_cast<float2 *>(smem);
  if (VAR >= 1) {
    for (int i = threadIdx.x; i < kCells; i += blockDim.x) tbl[i] = tbl_g[i];
    __syncthreads();
  }
  #ifdef CONFLICT_FREE
  const uint32_t tb = smem_u32(tbl) + (threadIdx.x & 31) * 8u;
#else
  const uint32_t tb = smem_u32(tbl);
#endif
  const uint32_t path = blockIdx.x * 8192u + threadIdx.x;
  float acc = 0.f;
  const long long t0 = clock64();
  for (int s = 0; s < 8; s += ILP) {
    float tot[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) tot[i] = 1000.f;
#pragma unroll 1
    for (int q = 0; q < nq; q++) {
#pragma unroll
      for (int i = 0; i < ILP; i++) {
        uint32_t w[4];
        stream_words(path + (s + i) * 1024u, 0u, (uint32_t)q, kTagGaussian, keys, w);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          float g0, g1;
          if (j >= VAR) gauss_pair(w[j], one, gc, g0, g1);  // VAR of the 4 pairs of a call take the table sin/cos
          else gauss_pair_tbl(w[j], one, gc, tb, kap, kap0, g0, g1);
          tot[i] = __fmul_rn(__fmul_rn(tot[i], g0), g1);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < ILP; i++) acc += tot[i];
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if ((threadIdx.x & 31) == 0) {
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2] = t0;
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2 + 1] = t1;
  }
}

template <int VAR, int ILP>
void run(const char *name, int nsm, float *out, const float2 *tbl, long long *cyc_d) {
  PhiloxKeys keys;
  uint32_t k0 = 1234u, k1 = 0u;
  for (int r = 0; r < 10; r++) {
    keys.k0[r] = k0;
    keys.k1[r] = k1;
    k0 += kPhiloxW0;
    k1 += kPhiloxW1;
  }
  const float sig = 0.01f * 10.f / 12.f;
  GaussConsts gc = {1.005f, -2.0f * 0.69314718f * sig * sig};
  const float kap = (float)(2.0 * M_PI * 8388608.0 / 4294967296.0);
  const float kap0 = (float)(-(1.0 + 131072.0 / 8388608.0) * (2.0 * M_PI * 8388608.0 / 4294967296.0));
  const int nq = 45;  // 360 months
  auto kern = k<VAR, ILP>;
  const size_t smem = VAR >= 1 ? kCells * sizeof(float2) : 0;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  kern<<<nsm, 1024, smem>>>(out, tbl, nq, keys, gc, 0x3F800000u, kap, kap0, cyc_d);
  cudaEventRecord(e0);
  kern<<<nsm, 1024, smem>>>(out, tbl, nq, keys, gc, 0x3F800000u, kap, kap0, cyc_d);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  std::vector<long long> cyc(nsm * 64);
  cudaMemcpy(cyc.data(), cyc_d, cyc.size() * 8, cudaMemcpyDeviceToHost);
  long long lo = cyc[0], hi = cyc[1];
  for (int i = 0; i < 64; i += 2) {  // CTA 0
    lo = std::min(lo, cyc[i]);
    hi = std::max(hi, cyc[i + 1]);
  }
  // 32 warps x 8 paths/lane x 360 months on 4 SMSPs -> warp-months per SMSP
  const double wm = 32.0 * 8 * 360 / 4;
  std::vector<float> o(8);
  cudaMemcpy(o.data(), out, 32, cudaMemcpyDeviceToHost);
  printf("%-34s ILP %d  cycles/warp-month/SMSP %6.2f  kernel %.3f ms  (%s)  out0 %.3f\n", name, ILP, (double)(hi - lo) / wm, ms,
         cudaGetErrorString(e), o[0]);
}

int main() {
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  float *out;
  long long *cyc_d;
  float2 *tbl;
  cudaMalloc(&out, (size_t)nsm * 1024 * 4);
  cudaMalloc(&cyc_d, (size_t)nsm * 64 * 8);
  cudaMalloc(&tbl, kCells * sizeof(float2));
  std::vector<float2> h(kCells);
  for (int i = 0; i < kCells; i++) {
    const double th = 2.0 * M_PI * (i + 0.5) / kCells;
    h[i] = make_float2((float)cos(th), (float)sin(th));
  }
  cudaMemcpy(tbl, h.data(), kCells * sizeof(float2), cudaMemcpyHostToDevice);
  run<0, 8>("0/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  run<1, 8>("1/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  run<2, 8>("2/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  run<3, 8>("3/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  run<4, 8>("4/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  run<2, 4>("2/4 pairs table sin/cos", nsm, out, tbl, cyc_d);
  return 0;
}
