// util.cu — error plumbing + version of libvo_b200 (no exceptions cross the C ABI).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace vo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA error: %s", what, cudaGetErrorString(e));
    return VO_ERR_CUDA;
  }
  return VO_OK;
}

}  // namespace vo

extern "C" int vo_version(void) { return 100; }
extern "C" const char* vo_last_error_string(void) { return vo::g_err; }

// ---------------------------------------------------------------- profiling hooks
#include <mutex>
#include <vector>

#include <nvtx3/nvToolsExt.h>
namespace vo {
static int g_prof_level = 0;   // 1: the matching / scoring / solve kernels (ids < 5); 2: every scope
static std::mutex g_prof_mu;
struct ProfRec { int id; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_prof;

// NVTX range around every scope (named like the profile ids of include/vo_b200.h): visible in
// Nsight Systems / ncu --nvtx; costs one null-pointer test when no tool is attached.
static const char* scope_name(int id) {
  static const char* names[] = {"vo:match",         "vo:ransac_score",  "vo:l2_gemm",      "vo:ransac_solve",
                                "vo:hamming_expand", "vo:recover_pose",  "vo:triangulate",  "vo:chain",
                                "vo:match_compact",  "vo:ransac_control", "vo:sift_pyramid", "vo:sift_extrema",
                                "vo:sift_describe"};
  return id >= 0 && id < (int)(sizeof(names) / sizeof(names[0])) ? names[id] : "vo:scope";
}

ProfScope::ProfScope(int kernel_id, cudaStream_t stream) : id(kernel_id), st(stream) {
  nvtxRangePushA(scope_name(kernel_id));
  if (g_prof_level == 0 || (g_prof_level == 1 && kernel_id >= 5)) return;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
}
ProfScope::~ProfScope() {
  nvtxRangePop();
  if (!e0) return;
  cudaEventRecord(e1, st);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof.push_back({id, e0, e1});
}

template <int MODE>
__global__ void __launch_bounds__(256) popc_body(unsigned* out, unsigned seed, int iters) {
  unsigned a[8], acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { a[k] = seed * (threadIdx.x + 1 + k * 977u) + blockIdx.x; acc[k] = k; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += __popc(a[k] ^ (acc[(k + 1) & 7]));
  }
  unsigned s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += acc[k];
  if (s == 0x12345678u) out[0] = s;
}
}  // namespace vo

extern "C" int vo_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(vo::g_prof_mu);
  for (auto& r : vo::g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  vo::g_prof.clear();
  vo::g_prof_level = on;
  return VO_OK;
}

extern "C" int vo_profile_read(int kernel_id, double* total_ms, int* launches) {
  std::lock_guard<std::mutex> lk(vo::g_prof_mu);
  double tot = 0.0;
  int n = 0;
  for (auto& r : vo::g_prof) {
    if (r.id != kernel_id) continue;
    if (cudaEventSynchronize(r.e1) != cudaSuccess) return vo::check_launch("vo_profile_read");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    tot += ms;
    ++n;
  }
  if (total_ms) *total_ms = tot;
  if (launches) *launches = n;
  return VO_OK;
}

// XOR + POPC + ADD per 32-bit word, 8 independent chains per thread: the Hamming inner body with
// nothing else around it.  Best of 5 launches; result in popc32 per second (host pointer).
extern "C" int vo_measure_popc_peak(double* popc32_per_s, void* stream) {
  VO_REQUIRE(popc32_per_s, VO_ERR_INVALID_ARG, "vo_measure_popc_peak: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  unsigned* out = nullptr;
  if (cudaMalloc(&out, 4) != cudaSuccess) return vo::check_launch("vo_measure_popc_peak");
  const int iters = 20000;
  dim3 grid(sms * 8), block(256);
  vo::popc_body<1><<<grid, block, 0, st>>>(out, 3u, 10);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 1e30;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0, st);
    vo::popc_body<1><<<grid, block, 0, st>>>(out, 3u, iters);
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *popc32_per_s = (double)grid.x * block.x * (double)iters * 32.0 / (best * 1e-3);
  return vo::check_launch("vo_measure_popc_peak");
}

extern "C" int vo_host_alloc(void** ptr, size_t bytes, int write_combined) {
  VO_REQUIRE(ptr && bytes > 0, VO_ERR_INVALID_ARG, "vo_host_alloc: null pointer or zero size");
  *ptr = nullptr;
  unsigned flags = cudaHostAllocPortable | (write_combined ? cudaHostAllocWriteCombined : 0u);
  if (cudaHostAlloc(ptr, bytes, flags) != cudaSuccess) return vo::check_launch("vo_host_alloc");
  return VO_OK;
}

extern "C" int vo_host_free(void* ptr) {
  if (ptr && cudaFreeHost(ptr) != cudaSuccess) return vo::check_launch("vo_host_free");
  return VO_OK;
}
