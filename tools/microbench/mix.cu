// Pipe-interaction microbenchmark for sm_100a (B200): how much ALU / FP32 / LSU work rides along for free next to
// the Philox multiplies (IMAD.WIDE on the fmaheavy pipe).  Development tool only (numbers go to profiles/).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mix mix.cu && ./mix
// One CTA of 1024 threads per SM (8 warps per SMSP), 4 independent chains per thread, 4 steps per loop trip.
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CH 4
#define ITERS 1024

// one "step": NR Philox rounds (2 WIDE + 2 LOP3 each) or their HI/LO split, + NL extra LOP3, NS extra SHF,
// NF extra FFMA, NA extra IADD3, ND LDS(+addr LEA) per chain
template <int NR, int SPLIT, int NL, int NS, int NF, int NA, int ND>
__device__ __forceinline__ void step(unsigned &x, unsigned &y, unsigned &z, float &f, unsigned &a, const unsigned c1,
                                     const unsigned c2, const float fa, const float fb, const unsigned *smu, unsigned lane) {
#pragma unroll
  for (int r = 0; r < NR; r++) {
    unsigned lo0, hi0, lo1, hi1;
    if (SPLIT) {
      asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(hi0) : "r"(x), "r"(c1));
      asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo0) : "r"(x), "r"(c1));
      asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(hi1) : "r"(y), "r"(c2));
      asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo1) : "r"(y), "r"(c2));
    } else {
      asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo0), "=r"(hi0) : "r"(x), "r"(c1));
      asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo1), "=r"(hi1) : "r"(y), "r"(c2));
    }
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(x) : "r"(hi1), "r"(lo0), "r"(c2));
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(y) : "r"(hi0), "r"(lo1), "r"(c1));
  }
#pragma unroll
  for (int i = 0; i < NL; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z) : "r"(c1), "r"(c2));
#pragma unroll
  for (int i = 0; i < NS; i++) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(z) : "r"(c1));
#pragma unroll
  for (int i = 0; i < NF; i++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
#pragma unroll
  for (int i = 0; i < NA; i++) asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(c1));
#pragma unroll
  for (int i = 0; i < ND; i++) {
    unsigned v = smu[(a & 1023u) * 32 + lane];   // LOP3 + LEA + LDS
    a += v;                                     // IADD
  }
}

template <int NR, int SPLIT, int NL, int NS, int NF, int NA, int ND>
__global__ void __launch_bounds__(1024, 1) k(unsigned *out, long long *cyc, unsigned seed, float fa, float fb) {
  extern __shared__ unsigned smu[];
  unsigned x[CH], y[CH], z[CH], a[CH];
  float f[CH];
  const unsigned lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32 * 1024; i += blockDim.x) smu[i] = (i * 2654435761u + seed) >> 9;
#pragma unroll
  for (int j = 0; j < CH; j++) {
    x[j] = (seed + 12345u) * (threadIdx.x + 1) + j * 7919u;
    y[j] = x[j] * 3u + 1u;
    z[j] = y[j] ^ 0x55555555u;
    a[j] = threadIdx.x * 31 + j;
    f[j] = 1.0f + 1e-3f * j + 1e-6f * threadIdx.x;
  }
  __syncthreads();
  long long t0 = clock64();
  const unsigned c1 = seed | 0xD2511F53u, c2 = seed | 0xCD9E8D57u;
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int rep = 0; rep < 4; rep++)
#pragma unroll
      for (int j = 0; j < CH; j++) step<NR, SPLIT, NL, NS, NF, NA, ND>(x[j], y[j], z[j], f[j], a[j], c1, c2, fa, fb, smu, lane);
  }
  long long t1 = clock64();
  unsigned acc = 0;
#pragma unroll
  for (int j = 0; j < CH; j++) acc ^= x[j] ^ y[j] ^ z[j] ^ a[j] ^ __float_as_uint(f[j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (lane == 0) {
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2] = t0;
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2 + 1] = t1;
  }
}

template <int NR, int SPLIT, int NL, int NS, int NF, int NA, int ND>
void run(const char *name, unsigned *out, long long *cyc_d, int nsm) {
  auto kern = k<NR, SPLIT, NL, NS, NF, NA, ND>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
  kern<<<nsm, 1024, 128 * 1024>>>(out, cyc_d, 1u, 1.0000001f, 1e-9f);
  kern<<<nsm, 1024, 128 * 1024>>>(out, cyc_d, 1u, 1.0000001f, 1e-9f);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(err)); return; }
  std::vector<long long> h(nsm * 64);
  cudaMemcpy(h.data(), cyc_d, h.size() * 8, cudaMemcpyDeviceToHost);
  std::vector<double> per_block;
  for (int b = 0; b < nsm; b++) {
    long long lo = h[b * 64], hi = h[b * 64 + 1];
    for (int w = 0; w < 32; w++) { lo = std::min(lo, h[(b * 32 + w) * 2]); hi = std::max(hi, h[(b * 32 + w) * 2 + 1]); }
    per_block.push_back((double)(hi - lo));
  }
  std::sort(per_block.begin(), per_block.end());
  const double cyc = per_block[per_block.size() / 2];
  const double steps = 8.0 * ITERS * 4 * CH;  // warp-steps per SMSP
  const int ninstr = NR * (SPLIT ? 6 : 4) + NL + NS + NF + NA + ND * 4;
  printf("%-64s instr/step %2d  cyc/step %7.3f  cyc/instr %5.3f\n", name, ninstr, cyc / steps, cyc / steps / ninstr);
}

int main() {
  int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  unsigned *out; long long *cyc;
  cudaMalloc(&out, nsm * 1024 * 4); cudaMalloc(&cyc, nsm * 64 * 8);
  //      NR SPLIT NL NS NF NA ND
  run<0, 0, 4, 0, 0, 0, 0>("4 LOP3", out, cyc, nsm);
  run<0, 0, 0, 4, 0, 0, 0>("4 SHF", out, cyc, nsm);
  run<0, 0, 0, 0, 4, 0, 0>("4 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 4, 0>("4 IADD", out, cyc, nsm);
  run<0, 0, 2, 0, 2, 0, 0>("2 LOP3 + 2 FFMA", out, cyc, nsm);
  run<0, 0, 2, 0, 0, 2, 0>("2 LOP3 + 2 IADD", out, cyc, nsm);
  run<0, 0, 2, 0, 4, 0, 0>("2 LOP3 + 4 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 0, 1>("1 LDS(+LOP3,LEA,IADD)", out, cyc, nsm);
  run<1, 0, 0, 0, 0, 0, 0>("Philox round (2 WIDE + 2 LOP3)", out, cyc, nsm);
  run<1, 1, 0, 0, 0, 0, 0>("Philox round split (2 HI + 2 LO + 2 LOP3)", out, cyc, nsm);
  run<1, 0, 1, 0, 0, 0, 0>("round + 1 LOP3", out, cyc, nsm);
  run<1, 0, 2, 0, 0, 0, 0>("round + 2 LOP3", out, cyc, nsm);
  run<1, 0, 4, 0, 0, 0, 0>("round + 4 LOP3", out, cyc, nsm);
  run<1, 0, 0, 0, 2, 0, 0>("round + 2 FFMA", out, cyc, nsm);
  run<1, 0, 0, 0, 4, 0, 0>("round + 4 FFMA", out, cyc, nsm);
  run<1, 0, 0, 0, 8, 0, 0>("round + 8 FFMA", out, cyc, nsm);
  run<1, 0, 0, 0, 0, 4, 0>("round + 4 IADD", out, cyc, nsm);
  run<1, 0, 2, 0, 2, 0, 0>("round + 2 LOP3 + 2 FFMA", out, cyc, nsm);
  run<1, 0, 0, 0, 0, 0, 1>("round + 1 LDS(+3)", out, cyc, nsm);
  run<1, 0, 1, 0, 2, 0, 1>("round + 1 LOP3 + 2 FFMA + 1 LDS(+3)   ~bootstrap mix", out, cyc, nsm);
  run<2, 0, 0, 0, 0, 0, 0>("2 rounds", out, cyc, nsm);
  return 0;
}
