// Pipe-throughput microbenchmarks for sm_100a (B200).  Development tool only:
// measures the reciprocal issue throughput (cycles per warp-instruction per SMSP)
// of the SASS instructions the Monte Carlo kernels are built from, so that the
// per-path-month recipes in DESIGN.md can be balanced across the fma / alu / xu /
// lsu pipes.  One CTA of 1024 threads per SM (8 warps per SMSP), ILP 8 per thread.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define ILP 8
#ifndef ITERS
#define ITERS 32768  // long enough (ms per kernel) for the SM clock to settle at its maximum
#endif

enum Op {
  FFMA, FMUL, FADD_RZ, LOP3, SHF, IADD, IMAD, IMAD_HI, IMAD_WIDE, PHILOX_ROUND, I2FP, F2I,
  MUFU_LG2, MUFU_EX2, MUFU_RSQ, MUFU_SQRT, MUFU_SIN, MUFU_COS, LDS_LINEAR, LDS_RANDOM, LDS_REPL,
  MIX_FFMA_LOP3, MIX_FFMA_IMAD, MIX_IMAD_LOP3, MIX_FFMA_IMAD_LOP3, MIX_4FFMA_MUFU, XORSHIFT, COMPAT_MONTH,
  BOXMULLER_PAIR, NOPS
};
struct Info { const char *name; int ops; };
static const Info info[NOPS] = {
  {"FFMA", 1}, {"FMUL", 1}, {"FADD.RZ", 1}, {"LOP3", 1}, {"SHF.R", 1}, {"IADD3", 1}, {"IMAD", 1}, {"IMAD.HI", 1},
  {"IMAD.WIDE (+1 LOP3)", 2}, {"Philox round (2 WIDE + 2 LOP3)", 4}, {"I2FP.F32.U32", 1}, {"F2I.TRUNC", 1},
  {"MUFU.LG2", 1}, {"MUFU.EX2", 1}, {"MUFU.RSQ", 1}, {"MUFU.SQRT", 1}, {"MUFU.SIN (+FMUL)", 2}, {"MUFU.COS (+FMUL)", 2},
  {"LDS linear (conflict-free)", 1}, {"LDS random word", 1}, {"LDS bank-private row (+IMAD)", 2},
  {"FFMA+LOP3", 2}, {"FFMA+IMAD", 2}, {"IMAD+LOP3", 2}, {"FFMA+IMAD+LOP3", 3}, {"4 FFMA + MUFU.LG2", 5},
  {"xorshift step (6)", 6}, {"compat month (12: xs6,I2FP,FMUL,F2I,IMAD,LDS,FFMA)", 12},
  {"Box-Muller pair (15)", 15},
};

template <int OP>
__device__ __forceinline__ void step(unsigned &x, unsigned &y, float &f, const unsigned c1, const unsigned c2,
                                     const float fa, const float fb, const unsigned *smu, const unsigned lane) {
  if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
  if (OP == FMUL) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f) : "f"(fa));
  if (OP == FADD_RZ) asm volatile("add.rz.f32 %0, %0, %1;" : "+f"(f) : "f"(fb));
  if (OP == LOP3) asm volatile("xor.b32 %0, %0, %1;" : "+r"(x) : "r"(c1));
  if (OP == SHF) asm volatile("shr.u32 %0, %0, 1;" : "+r"(x));
  if (OP == IADD) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(c1));
  if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(c1), "r"(c2));
  if (OP == IMAD_HI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x) : "r"(c1));
  if (OP == IMAD_WIDE) {
    unsigned lo, hi;
    asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(c1));
    asm volatile("xor.b32 %0, %1, %2;" : "=r"(x) : "r"(lo), "r"(hi));
  }
  if (OP == PHILOX_ROUND) {
    // state (x, y) + two more words folded in: hi0^k^c1 pattern with 3-input xor (LOP3)
    unsigned lo0, hi0, lo1, hi1;
    asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo0), "=r"(hi0) : "r"(x), "r"(c1));
    asm volatile("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo1), "=r"(hi1) : "r"(y), "r"(c2));
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(x) : "r"(hi1), "r"(lo0), "r"(c2));
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(y) : "r"(hi0), "r"(lo1), "r"(c1));
  }
  if (OP == I2FP) { asm volatile("cvt.rn.f32.u32 %0, %1;" : "=f"(f) : "r"(x)); x = __float_as_uint(f); }
  if (OP == F2I) { asm volatile("cvt.rzi.s32.f32 %0, %1;" : "=r"(x) : "f"(f)); f = __uint_as_float(x); }
  if (OP == MUFU_LG2) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == MUFU_EX2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == MUFU_RSQ) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == MUFU_SQRT) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == MUFU_SIN) asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == MUFU_COS) asm volatile("cos.approx.ftz.f32 %0, %0;" : "+f"(f));
  if (OP == LDS_LINEAR || OP == LDS_RANDOM) x = smu[x];
  if (OP == LDS_REPL) x = smu[x * 32 + lane];
  if (OP == MIX_FFMA_LOP3) {
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("xor.b32 %0, %0, %1;" : "+r"(x) : "r"(c1));
  }
  if (OP == MIX_FFMA_IMAD) {
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(c1), "r"(c2));
  }
  if (OP == MIX_IMAD_LOP3) {
    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(c1), "r"(c2));
    asm volatile("xor.b32 %0, %0, %1;" : "+r"(y) : "r"(c1));
  }
  if (OP == MIX_FFMA_IMAD_LOP3) {
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(c1), "r"(c2));
    asm volatile("xor.b32 %0, %0, %1;" : "+r"(y) : "r"(c1));
  }
  if (OP == MIX_4FFMA_MUFU) {
    float g = __uint_as_float(y);
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
    asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(g));
    y = __float_as_uint(g);
  }
  if (OP == XORSHIFT) { x ^= x << 11; x ^= x >> 7; x ^= x >> 12; }
  if (OP == COMPAT_MONTH) {
    x ^= x << 11; x ^= x >> 7; x ^= x >> 12;
    int idx = (int)((float)x * fa);          // fa = n * 2^-32
    f = fmaf(f, __uint_as_float(smu[idx * 32 + lane]), f);
  }
  if (OP == BOXMULLER_PAIR) {
    // x,y: two u32 -> two gaussian growth factors multiplied into f
    float u = fmaf((float)x, 2.3283064e-10f, 1.1641532e-10f);
    float v = fmaf((float)y, 1.4629181e-9f, 7.3145904e-10f);
    float t = fmaf(__log2f(u), fa, fb);
    float r; asm volatile("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t));
    float s = __sinf(v), c = __cosf(v);
    f *= fmaf(r, c, 1.005f);
    f *= fmaf(r, s, 1.005f);
    x += c1; y += c2;   // stand-in for fresh random words (2 IADD)
  }
}

template <int OP>
__global__ void __launch_bounds__(1024, 1) k(unsigned *out, long long *cyc, unsigned seed, float fa, float fb) {
  extern __shared__ unsigned smu[];
  unsigned x[ILP], y[ILP];
  float f[ILP];
  const unsigned lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32 * 1024; i += blockDim.x) {
    unsigned v = (i * 2654435761u + seed) >> 7;
    if (OP == LDS_LINEAR) v = ((i + 32) & 32767);
    else if (OP == LDS_RANDOM) v = v & 32767;
    else if (OP == LDS_REPL) v = v & 1023;
    else v = __float_as_uint(1e-3f * (float)(v & 255) - 0.1f);
    smu[i] = v;
  }
#pragma unroll
  for (int j = 0; j < ILP; j++) {
    x[j] = (seed + 12345u) * (threadIdx.x + 1) + j * 7919u;
    y[j] = x[j] * 3u + 1u;
    f[j] = 1.0f + 1e-3f * j + 1e-6f * threadIdx.x;
    if (OP == LDS_LINEAR || OP == LDS_RANDOM) x[j] = (threadIdx.x + j * 1024) & 32767;
    if (OP == LDS_REPL) x[j] = (threadIdx.x * 31 + j) & 1023;
    if (OP >= MUFU_LG2 && OP <= MUFU_COS) f[j] = 1.5f + j;
  }
  __syncthreads();
  long long t0 = clock64();
  const unsigned c1 = seed | 0xD2511F53u, c2 = seed | 0xCD9E8D57u;
  // unrolled 8x: with "#pragma unroll 1" the three loop instructions (IADD3, ISETP, BRA) per 8 timed ones showed up as
  // 1.42-1.45 "cycles" for the full-rate instructions (FFMA, FMUL, IADD3) in profiles/r01_pipe_microbench.txt
  // (only for those rows: ptxas folds eight consecutive XORs / shifts of one register, so the ALU-pipe rows keep the loop form,
  // where the 2-cycle instructions hide the loop overhead anyway)
  constexpr int kUnroll = (OP == FFMA || OP == FMUL || OP == FADD_RZ || OP == IADD) ? 8 : 1;
#pragma unroll kUnroll
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int j = 0; j < ILP; j++) step<OP>(x[j], y[j], f[j], c1, c2, fa, fb, smu, lane);
  }
  long long t1 = clock64();
  unsigned acc = 0;
#pragma unroll
  for (int j = 0; j < ILP; j++) acc ^= x[j] ^ y[j] ^ __float_as_uint(f[j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (lane == 0) {
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2] = t0;
    cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2 + 1] = t1;
  }
}

template <int OP>
void run(unsigned *out, long long *cyc_d, int nsm) {
  float fa = 1.0000001f, fb = 1e-9f;
  if (OP == COMPAT_MONTH) fa = 1023.0f * 2.3283064e-10f;  // idx <= 1023 even when float(x) rounds up to 2^32 (the reference's idx == n quirk)
  if (OP == BOXMULLER_PAIR) { fa = -9.6e-5f; fb = 0.0f; }
  cudaFuncSetAttribute(k<OP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<nsm, 1024, 128 * 1024>>>(out, cyc_d, 1u, fa, fb);   // warm-up
  cudaEventRecord(e0);
  k<OP><<<nsm, 1024, 128 * 1024>>>(out, cyc_d, 1u, fa, fb);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("%s: CUDA error %s\n", info[OP].name, cudaGetErrorString(err)); exit(1); }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<long long> h(nsm * 64);
  cudaMemcpy(h.data(), cyc_d, h.size() * 8, cudaMemcpyDeviceToHost);
  std::vector<double> per_block;
  for (int b = 0; b < nsm; b++) {
    long long lo = h[b * 64], hi = h[b * 64 + 1];
    for (int w = 0; w < 32; w++) { lo = std::min(lo, h[(b * 32 + w) * 2]); hi = std::max(hi, h[(b * 32 + w) * 2 + 1]); }
    per_block.push_back((double)(hi - lo));
  }
  std::sort(per_block.begin(), per_block.end());
  double cyc = per_block[per_block.size() / 2];
  double winstr_per_smsp = 8.0 * ITERS * ILP * info[OP].ops;
  double mhz = cyc / (ms * 1e3);
  printf("%-56s ops/iter %2d  cyc/warp-instr/SMSP %6.3f  cyc/iter/SMSP-warp %7.3f  (kernel %.3f ms, ~%.0f MHz)\n",
         info[OP].name, info[OP].ops, cyc / winstr_per_smsp, cyc / (8.0 * ITERS * ILP), ms, mhz);
}

template <int OP> struct Runner { static void go(unsigned *o, long long *c, int n) { run<OP>(o, c, n); Runner<OP + 1>::go(o, c, n); } };
template <> struct Runner<NOPS> { static void go(unsigned *, long long *, int) {} };

int main() {
  int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("device %s, %d SMs, cc %d.%d\n", p.name, nsm, p.major, p.minor);
  unsigned *out; long long *cyc;
  cudaMalloc(&out, nsm * 1024 * 4); cudaMalloc(&cyc, nsm * 64 * 8);
  Runner<0>::go(out, cyc, nsm);
  return 0;
}
