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
