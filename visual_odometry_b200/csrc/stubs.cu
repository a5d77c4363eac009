// stubs.cu — entry points declared in vo_b200.h whose kernels are not written yet.
// Each returns VO_ERR_INVALID_ARG with a clear message (never a silent CPU path).
#include "common.cuh"
#define NOT_YET(name) vo::set_error(name ": not implemented yet"); return VO_ERR_INVALID_ARG;
extern "C" size_t vo_pnp_workspace_bytes(int, int) { return 0; }
extern "C" int vo_pnp_ransac(const double*, const double*, int, const double*, int, double, double, double*, double*, int32_t*, int32_t*, int32_t*, void*, size_t, void*) { NOT_YET("vo_pnp_ransac") }
