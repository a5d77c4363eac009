// pipe_rates.cu — issue-rate microbenchmark for the integer/compare instructions the matching
// epilogues lean on (B200): ops per clock per SM with 32 resident warps, 8 independent chains per
// thread.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(1024) body(uint32_t* out, int iters, uint32_t seed) {
  uint32_t a[8], b = seed ^ threadIdx.x, c = seed * 3 + blockIdx.x;
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = seed + k * 977 + threadIdx.x;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (OP == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(b), "r"(c));
      if (OP == 1) asm volatile("max.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      if (OP == 2) { a[k] = __vimax3_u32(a[k], b, c); asm volatile("" : "+r"(a[k])); }
      if (OP == 3) asm volatile("max.f32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      if (OP == 4) asm volatile("max.f32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
      if (OP == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
      if (OP == 6) asm volatile("popc.b32 %0, %0;" : "+r"(a[k]));
      if (OP == 7) a[k] = __shfl_xor_sync(0xffffffffu, a[k], 1);
      if (OP == 8) a[k] = __reduce_max_sync(0xffffffffu, a[k]) + k;
      if (OP == 9) asm volatile("{.reg .pred p; setp.ne.u32 p, %2, 0; selp.b32 %0, %0, %1, p;}" : "+r"(a[k]) : "r"(b), "r"(c & 1));
      if (OP == 10) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      if (OP == 11) { a[k] = __viaddmax_u32(a[k], b, c); asm volatile("" : "+r"(a[k])); }
      // source-operand count: two register sources vs three
      if (OP == 12) asm volatile("xor.b32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      if (OP == 13) asm volatile("mad.lo.u32 %0, %0, %1, 12345;" : "+r"(a[k]) : "r"(b));
      if (OP == 14) asm volatile("mad.lo.u32 %0, %0, 129, %1;" : "+r"(a[k]) : "r"(c));
      if (OP == 15) asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; selp.b32 %0, %0, %1, p;}" : "+r"(a[k]) : "r"(b));
      if (OP == 16) asm volatile("lop3.b32 %0, %0, %1, 0x0f0f0f0f, 0x96;" : "+r"(a[k]) : "r"(b));
      if (OP == 17) {   // 1:1 mix of IMAD (3 sources) and a 2-source max
        if (k & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
        else asm volatile("max.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      }
      if (OP == 18) {   // 1:1 mix of IMAD and VIMNMX3 (both 3 sources)
        if (k & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
        else { a[k] = __vimax3_u32(a[k], b, c); asm volatile("" : "+r"(a[k])); }
      }
      if (OP == 19) {   // 1:1 mix of IMAD with an immediate addend and a 2-source max
        if (k & 1) asm volatile("mad.lo.u32 %0, %0, %1, 12345;" : "+r"(a[k]) : "r"(b));
        else asm volatile("max.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
      }
      if (OP == 20) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
      if (OP == 21) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[k]) : "r"(b));   // 3 sources, two of them the same register
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s ^= a[k];
  if (s == 0x12345678u) out[0] = s;
}

template <int OP>
double run(const char* name, uint32_t* d, int sms, double ghz) {
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  body<OP><<<sms, 1024>>>(d, 64, 1u);
  cudaEventRecord(e0);
  body<OP><<<sms, 1024>>>(d, iters, 7u);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double ops = (double)sms * 1024 * iters * 8;
  const double per_clk_sm = ops / (ms * 1e-3) / (ghz * 1e9) / sms;
  printf("%-14s %8.3f ms  %7.1f thread-ops/clk/SM  (warp instr every %.2f clk per SMSP)\n", name, ms,
         per_clk_sm, 128.0 / per_clk_sm);
  return per_clk_sm;
}

int main() {
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const double ghz = khz * 1e-6;
  uint32_t* d; cudaMalloc(&d, 64);
  printf("SMs %d, clock %.3f GHz (nominal; rates assume it)\n", sms, ghz);
  run<0>("LOP3", d, sms, ghz);
  run<10>("IADD", d, sms, ghz);
  run<1>("VIMNMX.U32", d, sms, ghz);
  run<2>("VIMNMX3.U32", d, sms, ghz);
  run<11>("VIADDMNMX", d, sms, ghz);
  run<3>("FMNMX", d, sms, ghz);
  run<4>("FMNMX3", d, sms, ghz);
  run<9>("SEL", d, sms, ghz);
  run<5>("IMAD", d, sms, ghz);
  run<6>("POPC", d, sms, ghz);
  run<7>("SHFL.BFLY", d, sms, ghz);
  run<8>("CREDUX.MAX", d, sms, ghz);
  run<12>("XOR 2src", d, sms, ghz);
  run<16>("LOP3 2src+imm", d, sms, ghz);
  run<13>("IMAD r,r,imm", d, sms, ghz);
  run<14>("IMAD r,imm,r", d, sms, ghz);
  run<21>("IMAD r,r,same", d, sms, ghz);
  run<20>("FFMA 3src", d, sms, ghz);
  run<15>("SETP+SEL", d, sms, ghz);
  run<17>("IMAD3+MAX2 mix", d, sms, ghz);
  run<18>("IMAD3+MAX3 mix", d, sms, ghz);
  run<19>("IMADimm+MAX2", d, sms, ghz);
  return 0;
}
