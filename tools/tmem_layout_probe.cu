// tmem_layout_probe.cu — which TMEM (lane, column) lands in which (thread, register) for
// tcgen05.ld shapes 16x256b.x4 and 16x128b.x8: write lane*1000+col with 32x32b stores, read back.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tmem_layout_probe tools/tmem_layout_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(uint32_t* out256, uint32_t* out128) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&slot)), "r"(64u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  // each warp writes its own 32 lanes x 32 columns: value = tmem_lane * 1000 + column
  const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
  uint32_t v[32];
  for (int c = 0; c < 32; ++c) v[c] = (uint32_t)((warp * 32 + lane) * 1000 + c);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // 16x256b.x4: 16 lanes x 32 columns -> 16 registers; once at lane offset 0, once at +16
  for (int half = 0; half < 2; ++half) {
    uint32_t r[16];
    const uint32_t a = taddr + ((uint32_t)(half * 16) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(a) : "memory");
    for (int k = 0; k < 16; ++k) out256[((warp * 2 + half) * 32 + lane) * 16 + k] = r[k];
  }
  for (int half = 0; half < 2; ++half) {
    uint32_t r[16];
    const uint32_t a = taddr + ((uint32_t)(half * 16) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.16x128b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(a) : "memory");
    for (int k = 0; k < 16; ++k) out128[((warp * 2 + half) * 32 + lane) * 16 + k] = r[k];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

int main() {
  uint32_t *d256, *d128;
  const int n = 4 * 2 * 32 * 16;
  cudaMalloc(&d256, n * 4); cudaMalloc(&d128, n * 4);
  probe<<<1, 128>>>(d256, d128);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  static uint32_t h[n];
  for (int which = 0; which < 2; ++which) {
    cudaMemcpy(h, which ? d128 : d256, n * 4, cudaMemcpyDeviceToHost);
    printf("== %s, warp 1 (TMEM lanes 32..63): thread -> regs as lane:col\n", which ? "16x128b.x8" : "16x256b.x4");
    for (int half = 0; half < 2; ++half)
      for (int t = 0; t < 32; t += (t < 8 ? 1 : 8)) {
        printf("half %d thread %2d:", half, t);
        for (int k = 0; k < 16; ++k) { uint32_t v = h[((1 * 2 + half) * 32 + t) * 16 + k]; printf(" %u:%u", v / 1000, v % 1000); }
        printf("\n");
      }
  }
  return 0;
}
