// popc_peak.cu — measures the POPC issue-rate ceiling of the GPU (denominator of the K1 roofline;
// MEASURED_PEAKS.json has no integer entry).  Three bodies:
//   popc  : independent POPC chains only                       -> raw pipe rate
//   xpa   : XOR + POPC + ADD per word (the Hamming inner body) -> what a perfect K1 could reach
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o popc_peak popc_peak.cu
// Prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) body(unsigned* out, unsigned seed, int iters) {
  unsigned a[8], acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = seed * (threadIdx.x + 1 + k * 977u) + blockIdx.x;
    acc[k] = k;
  }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (MODE == 0) {
          acc[k] = __popc(acc[k] + a[k]);   // popc + add dependent chain, 8 chains
        } else {
          acc[k] += __popc(a[k] ^ (acc[(k + 1) & 7]));
        }
      }
    }
  }
  unsigned s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += acc[k];
  if (s == 0x12345678u) out[0] = s;
}

template <int MODE>
double run(int sms, int iters) {
  unsigned* out;
  cudaMalloc(&out, 4);
  dim3 grid(sms * 8), block(256);
  body<MODE><<<grid, block>>>(out, 3u, 10);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 1e30;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    body<MODE><<<grid, block>>>(out, 3u, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  double popc = (double)grid.x * block.x * (double)iters * 32.0;
  cudaFree(out);
  return popc / (best * 1e-3);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int iters = 20000;
  double r0 = run<0>(p.multiProcessorCount, iters);
  double r1 = run<1>(p.multiProcessorCount, iters);
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d, \"popc32_per_s_chain\": %.4e, "
         "\"popc32_per_s_xor_popc_add\": %.4e, \"popc_per_clk_per_sm_at_max_clock\": %.2f}\n",
         p.name, p.multiProcessorCount, clk, r0, r1,
         (r1 > r0 ? r1 : r0) / ((double)p.multiProcessorCount * clk * 1e3));
  return 0;
}
