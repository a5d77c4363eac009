// mapops.cu — N2: reprojection errors of all map observations in one launch.
//
// Replaces the per-observation Python loops of the reference's Map
//   calculate_reprojection_error                                  (src/map.py:72-97)
//   remove_observations_with_reprojection_errors_above_threshold  (src/map.py:46-70)
//   calculate_reprojection_error_for_point                        (src/map.py:32-44)
// called several times per frame (src/visual_slam.py:181,266; src/map.py:239-241):
//   p_cam = pose(camera) @ [X,1];  t = K @ p_cam[:3];  t /= t[2];
//   sqerr = |dx*dx| + |dy*dy|  with (dx,dy) = t[:2] - image_coordinates.
// One thread per observation; the structure-of-arrays map lives in HBM, cameras and K are tiny and
// stay in L1/L2.  HBM-bound: 24 B of indices/coordinates + a 24 B point gather in, 9 B out.
// The total is a fixed-order tree sum (per block, then across blocks in block order), so it is
// reproducible run to run; it differs from Python's sequential sum only by fp64 re-association.
#include "common.cuh"

namespace vo {

constexpr int MAPOPS_THREADS = 256;

__global__ void __launch_bounds__(MAPOPS_THREADS)
reproj_kernel(const double* __restrict__ points, const double* __restrict__ poses,
              const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
              const double* __restrict__ obs_xy, int n_obs, int n_points, int n_cams,
              const double* __restrict__ K, double threshold, double* __restrict__ sqerr,
              uint8_t* __restrict__ keep, double* __restrict__ block_sums,
              int32_t* __restrict__ block_keep) {
  __shared__ double red[MAPOPS_THREADS / 32];
  __shared__ int redk[MAPOPS_THREADS / 32];
  const int i = blockIdx.x * MAPOPS_THREADS + threadIdx.x;
  double e = 0.0;
  int kept = 0;
  if (i < n_obs) {
    const int c = obs_cam[i], p = obs_pt[i];
    double err = __longlong_as_double(0x7ff8000000000000ll);   // NaN: bad index
    if (c >= 0 && c < n_cams && p >= 0 && p < n_points) {
      const double* T = poses + (size_t)c * 16;
      const double X = __ldg(points + 3 * (size_t)p), Y = __ldg(points + 3 * (size_t)p + 1),
                   Z = __ldg(points + 3 * (size_t)p + 2);
      // row-times-column in index order, no FMA contraction (numpy's small matmul order)
      double pc[3];
#pragma unroll
      for (int r = 0; r < 3; ++r)
        pc[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(__ldg(T + 4 * r), X),
                                              __dmul_rn(__ldg(T + 4 * r + 1), Y)),
                                    __dmul_rn(__ldg(T + 4 * r + 2), Z)),
                          __ldg(T + 4 * r + 3));
      double t[3];
#pragma unroll
      for (int r = 0; r < 3; ++r)
        t[r] = __dadd_rn(__dadd_rn(__dmul_rn(__ldg(K + 3 * r), pc[0]), __dmul_rn(__ldg(K + 3 * r + 1), pc[1])),
                         __dmul_rn(__ldg(K + 3 * r + 2), pc[2]));
      const double2 xy = reinterpret_cast<const double2*>(obs_xy)[i];
      const double dx = __dsub_rn(__ddiv_rn(t[0], t[2]), xy.x);
      const double dy = __dsub_rn(__ddiv_rn(t[1], t[2]), xy.y);
      err = __dadd_rn(fabs(__dmul_rn(dx, dx)), fabs(__dmul_rn(dy, dy)));
    }
    sqerr[i] = err;
    kept = err < threshold;        // NaN compares false, like the Python `if`
    if (keep) keep[i] = (uint8_t)kept;
    e = err;
  }
  // fixed-order block sum: warp shuffle tree, then warps in order
  for (int o = 16; o > 0; o >>= 1) {
    e += __shfl_down_sync(0xffffffffu, e, o);
    kept += __shfl_down_sync(0xffffffffu, kept, o);
  }
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = e; redk[threadIdx.x >> 5] = kept; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    int k = 0;
    for (int w = 0; w < MAPOPS_THREADS / 32; ++w) { s += red[w]; k += redk[w]; }
    block_sums[blockIdx.x] = s;
    block_keep[blockIdx.x] = k;
  }
}

__global__ void reproj_total_kernel(const double* __restrict__ block_sums,
                                    const int32_t* __restrict__ block_keep, int n_blocks,
                                    double* __restrict__ total, int32_t* __restrict__ n_keep) {
  // one warp, strided partial sums in fixed order, then a shuffle tree
  const int lane = threadIdx.x;
  double s = 0.0;
  int k = 0;
  for (int b = lane; b < n_blocks; b += 32) { s += block_sums[b]; k += block_keep[b]; }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, o);
    k += __shfl_down_sync(0xffffffffu, k, o);
  }
  if (lane == 0) { *total = s; if (n_keep) *n_keep = k; }
}

}  // namespace vo

using namespace vo;

extern "C" size_t vo_reprojection_workspace_bytes(int n_obs) {
  const size_t nb = (size_t)((n_obs > 0 ? n_obs : 0) + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  return nb * (sizeof(double) + sizeof(int32_t)) + 256;
}

extern "C" int vo_reprojection_errors(const double* points, int n_points, const double* poses,
                                      int n_cams, const int32_t* obs_cam, const int32_t* obs_pt,
                                      const double* obs_xy, int n_obs, const double* K,
                                      double threshold, double* sqerr, uint8_t* keep,
                                      double* total, int32_t* n_keep, void* ws, size_t ws_bytes,
                                      void* stream) {
  VO_REQUIRE(n_obs >= 0 && n_points >= 0 && n_cams >= 0, VO_ERR_INVALID_ARG,
             "vo_reprojection_errors: negative size");
  VO_REQUIRE(total && K, VO_ERR_INVALID_ARG, "vo_reprojection_errors: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_obs == 0) {
    cudaMemsetAsync(total, 0, sizeof(double), st);
    if (n_keep) cudaMemsetAsync(n_keep, 0, sizeof(int32_t), st);
    return check_launch("vo_reprojection_errors");
  }
  VO_REQUIRE(points && poses && obs_cam && obs_pt && obs_xy && sqerr && ws, VO_ERR_INVALID_ARG,
             "vo_reprojection_errors: null pointer");
  VO_REQUIRE(ws_bytes >= vo_reprojection_workspace_bytes(n_obs), VO_ERR_WORKSPACE,
             "vo_reprojection_errors: workspace too small");
  const int nb = (n_obs + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  double* block_sums = (double*)ws;
  int32_t* block_keep = (int32_t*)(block_sums + nb);
  reproj_kernel<<<nb, MAPOPS_THREADS, 0, st>>>(points, poses, obs_cam, obs_pt, obs_xy, n_obs,
                                              n_points, n_cams, K, threshold, sqerr, keep,
                                              block_sums, block_keep);
  reproj_total_kernel<<<1, 32, 0, st>>>(block_sums, block_keep, nb, total, n_keep);
  return check_launch("vo_reprojection_errors");
}
