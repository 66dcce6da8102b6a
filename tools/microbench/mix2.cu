// Second pipe-interaction microbenchmark (sm_100a): is IMAD.WIDE 2 or 4 issue cycles, what rides along with MUFU,
// are subnormal FFMAs full rate.  Development tool only.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mix2 mix2.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CH 4
#define ITERS 1024

// per step and chain: NW x (WIDE, result folded by IADD), NH x IMAD.HI, NI x IMAD.LO, NM x MUFU.LG2(+FADD to stay in range),
// NL x LOP3, NF x FFMA, ND x subnormal FFMA.RZ, NC x I2FP
template <int NW, int NH, int NI, int NM, int NL, int NF, int ND, int NC>
__device__ __forceinline__ void step(unsigned &x, unsigned &h, unsigned &z, float &f, float &g, float &d, const unsigned c1,
                                     const unsigned c2, const float fa, const float fb, const float dn) {
#pragma unroll
  for (int i = 0; i < NW; i++) {
    unsigned lo, hi;
    asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(c1));
    asm volatile("add.u32 %0, %1, %2;" : "=r"(x) : "r"(lo), "r"(hi));
  }
#pragma unroll
  for (int i = 0; i < NH; i++) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(h) : "r"(c2));
#pragma unroll
  for (int i = 0; i < NI; i++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(h) : "r"(c1), "r"(c2));
#pragma unroll
  for (int i = 0; i < NM; i++) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(g));
#pragma unroll
  for (int i = 0; i < NL; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z) : "r"(c1), "r"(c2));
#pragma unroll
  for (int i = 0; i < NF; i++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
#pragma unroll
  for (int i = 0; i < ND; i++) asm volatile("fma.rz.f32 %0, %1, %2, %0;" : "+f"(d) : "f"(fa), "f"(dn));
#pragma unroll
  for (int i = 0; i < NC; i++) { float t; asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(t) : "r"(z)); z = __float_as_uint(t); }
}

template <int NW, int NH, int NI, int NM, int NL, int NF, int ND, int NC>
__global__ void __launch_bounds__(1024, 1) k(unsigned *out, long long *cyc, unsigned seed, float fa, float fb, float dn) {
  unsigned x[CH], h[CH], z[CH];
  float f[CH], g[CH], d[CH];
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < CH; j++) {
    x[j] = (seed + 12345u) * (threadIdx.x + 1) + j * 7919u;
    h[j] = x[j] * 3u + 1u;
    z[j] = h[j] ^ 0x55555555u;
    f[j] = 1.0f + 1e-3f * j + 1e-6f * threadIdx.x;
    g[j] = 1.5f + j;
    d[j] = __uint_as_float(1000u + j);  // subnormal
  }
  long long t0 = clock64();
  const unsigned c1 = seed | 0xD2511F53u, c2 = seed | 0xCD9E8D57u;
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int rep = 0; rep < 4; rep++)
#pragma unroll
      for (int j = 0; j < CH; j++) {
        step<NW, NH, NI, NM, NL, NF, ND, NC>(x[j], h[j], z[j], f[j], g[j], d[j], c1, c2, fa, fb, dn);
        if (NM) g[j] = fabsf(g[j]) + 1.5f;  // keep the MUFU argument positive (1 FADD per step, counted)
      }
  }
  long long t1 = clock64();
  unsigned acc = 0;
#pragma unroll
  for (int j = 0; j < CH; j++) acc ^= x[j] ^ h[j] ^ z[j] ^ __float_as_uint(f[j]) ^ __float_as_uint(g[j]) ^ __float_as_uint(d[j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (lane == 0) {
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2] = t0;
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2 + 1] = t1;
  }
}

template <int NW, int NH, int NI, int NM, int NL, int NF, int ND, int NC>
void run(const char *name, unsigned *out, long long *cyc_d, int nsm) {
  auto kern = k<NW, NH, NI, NM, NL, NF, ND, NC>;
  kern<<<nsm, 1024>>>(out, cyc_d, 1u, 1.0000001f, 1e-9f, 1e-42f);
  kern<<<nsm, 1024>>>(out, cyc_d, 1u, 1.0000001f, 1e-9f, 1e-42f);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(err)); return; }
  std::vector<long long> hh(nsm * 64);
  cudaMemcpy(hh.data(), cyc_d, hh.size() * 8, cudaMemcpyDeviceToHost);
  std::vector<double> per_block;
  for (int b = 0; b < nsm; b++) {
    long long lo = hh[b * 64], hi = hh[b * 64 + 1];
    for (int w = 0; w < 32; w++) { lo = std::min(lo, hh[(b * 32 + w) * 2]); hi = std::max(hi, hh[(b * 32 + w) * 2 + 1]); }
    per_block.push_back((double)(hi - lo));
  }
  std::sort(per_block.begin(), per_block.end());
  const double cyc = per_block[per_block.size() / 2];
  const double steps = 8.0 * ITERS * 4 * CH;
  printf("%-56s cyc/step %7.3f\n", name, cyc / steps);
}

int main() {
  int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  unsigned *out; long long *cyc;
  cudaMalloc(&out, nsm * 1024 * 4); cudaMalloc(&cyc, nsm * 64 * 8);
  //  NW NH NI NM NL NF ND NC
  run<2, 0, 0, 0, 0, 0, 0, 0>("2 x (WIDE + IADD)", out, cyc, nsm);
  run<2, 0, 0, 0, 2, 0, 0, 0>("2 x (WIDE + IADD) + 2 LOP3", out, cyc, nsm);
  run<2, 0, 0, 0, 0, 4, 0, 0>("2 x (WIDE + IADD) + 4 FFMA", out, cyc, nsm);
  run<0, 2, 0, 0, 0, 0, 0, 0>("2 IMAD.HI", out, cyc, nsm);
  run<0, 2, 0, 0, 2, 0, 0, 0>("2 IMAD.HI + 2 LOP3", out, cyc, nsm);
  run<0, 2, 0, 0, 0, 4, 0, 0>("2 IMAD.HI + 4 FFMA", out, cyc, nsm);
  run<0, 0, 2, 0, 0, 0, 0, 0>("2 IMAD", out, cyc, nsm);
  run<0, 0, 2, 0, 2, 0, 0, 0>("2 IMAD + 2 LOP3", out, cyc, nsm);
  run<0, 0, 2, 0, 0, 2, 0, 0>("2 IMAD + 2 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 4, 0, 0>("4 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 0, 4, 0>("4 FFMA.RZ subnormal", out, cyc, nsm);
  run<0, 0, 0, 0, 1, 1, 0, 0>("1 LOP3 + 1 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 1, 2, 0, 0>("1 LOP3 + 2 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 1, 3, 0, 0>("1 LOP3 + 3 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 0, 0, 2>("2 I2FP", out, cyc, nsm);
  run<0, 0, 0, 0, 0, 2, 0, 2>("2 I2FP + 2 FFMA", out, cyc, nsm);
  run<0, 0, 0, 0, 2, 0, 0, 2>("2 I2FP + 2 LOP3", out, cyc, nsm);
  run<0, 0, 0, 1, 0, 0, 0, 0>("1 MUFU (+FADD)", out, cyc, nsm);
  run<0, 0, 0, 1, 2, 0, 0, 0>("1 MUFU (+FADD) + 2 LOP3", out, cyc, nsm);
  run<0, 0, 0, 1, 3, 0, 0, 0>("1 MUFU (+FADD) + 3 LOP3", out, cyc, nsm);
  run<0, 0, 0, 1, 4, 0, 0, 0>("1 MUFU (+FADD) + 4 LOP3", out, cyc, nsm);
  run<0, 0, 0, 1, 0, 4, 0, 0>("1 MUFU (+FADD) + 4 FFMA", out, cyc, nsm);
  run<0, 0, 0, 1, 0, 6, 0, 0>("1 MUFU (+FADD) + 6 FFMA", out, cyc, nsm);
  run<0, 0, 0, 1, 2, 3, 0, 0>("1 MUFU (+FADD) + 2 LOP3 + 3 FFMA", out, cyc, nsm);
  run<1, 0, 0, 1, 0, 0, 0, 0>("1 MUFU (+FADD) + 1 (WIDE + IADD)", out, cyc, nsm);
  run<1, 0, 0, 1, 1, 2, 0, 0>("1 MUFU (+FADD) + 1 (WIDE + IADD) + 1 LOP3 + 2 FFMA  ~gauss", out, cyc, nsm);
  return 0;
}
