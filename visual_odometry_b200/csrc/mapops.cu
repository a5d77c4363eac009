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
constexpr int REPROJ_SMEM_CAMS = 128;

__global__ void __launch_bounds__(MAPOPS_THREADS, 8)
reproj_kernel(const double* __restrict__ points, const double* __restrict__ poses,
              const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
              const double* __restrict__ obs_xy, int n_obs, int n_points, int n_cams,
              const double* __restrict__ K, double threshold, double* __restrict__ sqerr,
              uint8_t* __restrict__ keep, double* __restrict__ block_sums,
              int32_t* __restrict__ block_keep) {
  __shared__ double red[MAPOPS_THREADS / 32];
  __shared__ int redk[MAPOPS_THREADS / 32];
  // camera poses (rows 0..2) and K staged in shared memory, pose stride 13 doubles so lanes that
  // look at different cameras hit different banks (global loads of 18 scattered 128-byte pose
  // rows per instruction made the kernel LSU-bound: 1.0 of 6.5 TB/s)
  __shared__ double sT[REPROJ_SMEM_CAMS * 13];
  __shared__ double sK[9];
  const bool staged = n_cams <= REPROJ_SMEM_CAMS;
  if (staged)
    for (int e = threadIdx.x; e < n_cams * 12; e += MAPOPS_THREADS)
      sT[(e / 12) * 13 + e % 12] = poses[(size_t)(e / 12) * 16 + e % 12];
  if (threadIdx.x < 9) sK[threadIdx.x] = K[threadIdx.x];
  __syncthreads();
  // persistent blocks: tile = 256 consecutive observations; the per-tile partial sums (and so the
  // summation order of the total) do not depend on the grid size
  const int n_tiles = (n_obs + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
  const int i = tile * MAPOPS_THREADS + threadIdx.x;
  double e = 0.0;
  int kept = 0;
  if (i < n_obs) {
    const int c = obs_cam[i], p = obs_pt[i];
    double err = __longlong_as_double(0x7ff8000000000000ll);   // NaN: bad index
    if (c >= 0 && c < n_cams && p >= 0 && p < n_points) {
      double T[12];
      if (staged) {
#pragma unroll
        for (int k = 0; k < 12; ++k) T[k] = sT[c * 13 + k];
      } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) T[k] = __ldg(poses + (size_t)c * 16 + k);
      }
      const double X = __ldg(points + 3 * (size_t)p), Y = __ldg(points + 3 * (size_t)p + 1),
                   Z = __ldg(points + 3 * (size_t)p + 2);
      // row-times-column in index order, no FMA contraction (numpy's small matmul order)
      double pc[3];
#pragma unroll
      for (int r = 0; r < 3; ++r)
        pc[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[4 * r], X), __dmul_rn(T[4 * r + 1], Y)),
                                    __dmul_rn(T[4 * r + 2], Z)),
                          T[4 * r + 3]);
      double t[3];
#pragma unroll
      for (int r = 0; r < 3; ++r)
        t[r] = __dadd_rn(__dadd_rn(__dmul_rn(sK[3 * r], pc[0]), __dmul_rn(sK[3 * r + 1], pc[1])),
                         __dmul_rn(sK[3 * r + 2], pc[2]));
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
    block_sums[tile] = s;
    block_keep[tile] = k;
  }
  __syncthreads();
  }
}

__global__ void __launch_bounds__(1024)
reproj_total_kernel(const double* __restrict__ block_sums, const int32_t* __restrict__ block_keep,
                    int n_blocks, double* __restrict__ total, int32_t* __restrict__ n_keep) {
  // one block: strided partial sums in a fixed order, shuffle tree, warps in order
  __shared__ double red[32];
  __shared__ int redk[32];
  const int tid = threadIdx.x;
  double s = 0.0;
  int k = 0;
  for (int b = tid; b < n_blocks; b += 1024) { s += block_sums[b]; k += block_keep[b]; }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, o);
    k += __shfl_down_sync(0xffffffffu, k, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = s; redk[tid >> 5] = k; }
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    int c = 0;
    for (int w = 0; w < 32; ++w) { t += red[w]; c += redk[w]; }
    *total = t;
    if (n_keep) *n_keep = c;
  }
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
  const int grid = nb < 148 * 8 ? nb : 148 * 8;       // 8 resident blocks per SM
  reproj_kernel<<<grid, MAPOPS_THREADS, 0, st>>>(points, poses, obs_cam, obs_pt, obs_xy, n_obs,
                                              n_points, n_cams, K, threshold, sqerr, keep,
                                              block_sums, block_keep);
  reproj_total_kernel<<<1, 1024, 0, st>>>(block_sums, block_keep, nb, total, n_keep);
  return check_launch("vo_reprojection_errors");
}

// ---------------------------------------------------------------------------------------------
// Map growth and pruning on the device-resident structure-of-arrays map.
//
// vo_map_extend = the loop of VisualSlam.add_information_to_map (src/visual_slam.py:172-179) with
// add_point_observation_to_map / add_new_observation_of_existing_point / add_new_match_to_map
// (:100-150), and, with max_norm = +inf on an empty map, the two loops of initialize_map
// (:49-52, :69-83).  For every 3-D match, in match order:
//     if |point| > 50: skip
//     root = track_feature_back_in_time(featureid2)
//     root in mappointdict ? append Observation(point, camera2, keypoint2)
//                          : append TrackedPoint(point, featureid1), Observation(.., camera1,
//                            keypoint1), Observation(.., camera2, keypoint2)
// mappointdict is built BEFORE the loop (:154-156) and not touched inside it, so the lookups of
// one call never see the points the same call appends; the new points are published to
// root_point[] after the loop (what the rebuilds at :193-196 / :238-240 do next).
// One block: an ordered (stable) compaction with two running offsets, n <= a few thousand.
namespace vo {

constexpr int EXT_THREADS = 1024;

__device__ __forceinline__ int2 block_excl_scan2(int a, int b, int2* warp_tot, int2& total) {
  // exclusive prefix sums of (a, b) over the block, in thread order
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int sa = a, sb = b;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int ua = __shfl_up_sync(0xffffffffu, sa, o), ub = __shfl_up_sync(0xffffffffu, sb, o);
    if (lane >= o) { sa += ua; sb += ub; }
  }
  if (lane == 31) warp_tot[wid] = make_int2(sa, sb);
  __syncthreads();
  int oa = 0, ob = 0, ta = 0, tb = 0;
  for (int w = 0; w < EXT_THREADS / 32; ++w) {
    const int2 v = warp_tot[w];
    if (w < wid) { oa += v.x; ob += v.y; }
    ta += v.x; tb += v.y;
  }
  __syncthreads();
  total = make_int2(ta, tb);
  return make_int2(oa + sa - a, ob + sb - b);
}

__global__ void __launch_bounds__(EXT_THREADS)
map_extend_kernel(const int32_t* __restrict__ q, const int32_t* __restrict__ t,
                  const uint8_t* __restrict__ mask, const int32_t* __restrict__ n_match, int f1,
                  int f2, int n_cap, const int32_t* __restrict__ root, int32_t* root_point,
                  const float* __restrict__ kp, const double* __restrict__ X, int x_stride,
                  int cam1, int cam2, double max_norm, double* __restrict__ points,
                  int32_t* point_fid, int32_t* n_points, int cap_points,
                  double* __restrict__ obs_xy, int32_t* __restrict__ obs_cam,
                  int32_t* __restrict__ obs_pt, int32_t* n_obs, int cap_obs, int publish,
                  int32_t* __restrict__ status) {
  __shared__ int2 warp_tot[EXT_THREADS / 32];
  const int tid = threadIdx.x;
  const int n = *n_match;
  const int p0 = *n_points, o0 = *n_obs;
  int rank_base = 0, p_base = 0, o_base = 0;
  bool overflow = false;
  for (int i0 = 0; i0 < n; i0 += EXT_THREADS) {
    const int i = i0 + tid;
    const bool inl = i < n && (!mask || mask[i]);
    int2 tot;
    const int rank = rank_base + block_excl_scan2(inl, 0, warp_tot, tot).x;
    rank_base += tot.x;
    bool live = false, existing = false;
    int prow = -1;
    double P[3] = {0.0, 0.0, 0.0};
    if (inl) {
      const double* x = X + (size_t)rank * x_stride;
      P[0] = x[0]; P[1] = x[1]; P[2] = x[2];
      // np.linalg.norm: sqrt of the sum of squares; `norm > 50` is false for NaN
      const double nrm = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(P[0], P[0]), __dmul_rn(P[1], P[1])),
                                        __dmul_rn(P[2], P[2])));
      live = !(nrm > max_norm);
      if (live) {
        prow = root_point[root[f2 * n_cap + t[i]]];
        existing = prow >= 0;
      }
    }
    const int no = live ? (existing ? 1 : 2) : 0, np = (live && !existing) ? 1 : 0;
    const int2 off = block_excl_scan2(no, np, warp_tot, tot);
    const int oi = o0 + o_base + off.x, pi = p0 + p_base + off.y;
    o_base += tot.x;
    p_base += tot.y;
    if (live) {
      const float2 k2 = reinterpret_cast<const float2*>(kp)[(size_t)f2 * n_cap + t[i]];
      if (existing) {
        if (oi < cap_obs) {
          obs_pt[oi] = prow; obs_cam[oi] = cam2;
          obs_xy[2 * oi] = (double)k2.x; obs_xy[2 * oi + 1] = (double)k2.y;
        } else overflow = true;
      } else {
        if (pi < cap_points && oi + 1 < cap_obs) {
          const float2 k1 = reinterpret_cast<const float2*>(kp)[(size_t)f1 * n_cap + q[i]];
          points[3 * (size_t)pi] = P[0]; points[3 * (size_t)pi + 1] = P[1];
          points[3 * (size_t)pi + 2] = P[2];
          point_fid[pi] = f1 * n_cap + q[i];
          obs_pt[oi] = pi; obs_cam[oi] = cam1;
          obs_xy[2 * oi] = (double)k1.x; obs_xy[2 * oi + 1] = (double)k1.y;
          obs_pt[oi + 1] = pi; obs_cam[oi + 1] = cam2;
          obs_xy[2 * oi + 2] = (double)k2.x; obs_xy[2 * oi + 3] = (double)k2.y;
        } else overflow = true;
      }
    }
  }
  const int any_over = __syncthreads_or(overflow);
  const int p1 = min(p0 + p_base, cap_points), o1 = min(o0 + o_base, cap_obs);
  if (publish && !any_over)
    for (int r = p0 + tid; r < p1; r += EXT_THREADS) root_point[point_fid[r]] = r;
  if (tid == 0) {
    if (any_over) {            // nothing is committed when the arrays are too small
      *status = 1;
    } else {
      *n_points = p1;
      *n_obs = o1;
    }
  }
}

// remove_observations_with_reprojection_errors_above_threshold (src/map.py:46-70) on the flat
// arrays: stable compaction of the observation rows by keep flags (the flags come from
// vo_reprojection_errors).  count -> scan -> scatter, output in a second set of arrays.
__global__ void __launch_bounds__(MAPOPS_THREADS)
prune_count_kernel(const uint8_t* __restrict__ keep, int n, int32_t* __restrict__ block_cnt) {
  const int i = blockIdx.x * MAPOPS_THREADS + threadIdx.x;
  const int c = __syncthreads_count(i < n && keep[i]);
  if (threadIdx.x == 0) block_cnt[blockIdx.x] = c;
}

__global__ void __launch_bounds__(1024)
prune_scan_kernel(int32_t* __restrict__ block_cnt, int nb, int32_t* __restrict__ n_out) {
  // in-place exclusive scan of the block counts, one block, tiles of 1024 with a running base
  __shared__ int wsum[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int b0 = 0; b0 < nb; b0 += 1024) {
    const int b = b0 + tid;
    const int v = b < nb ? block_cnt[b] : 0;
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += u;
    }
    if (lane == 31) wsum[wid] = s;
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < wid; ++w) off += wsum[w];
    if (b < nb) block_cnt[b] = off + s - v;
    __syncthreads();
    if (tid == 1023) base_s = off + s;
    __syncthreads();
  }
  if (tid == 0) *n_out = base_s;
}

__global__ void __launch_bounds__(MAPOPS_THREADS)
prune_scatter_kernel(const uint8_t* __restrict__ keep, int n, const int32_t* __restrict__ block_off,
                     const int32_t* __restrict__ cam_in, const int32_t* __restrict__ pt_in,
                     const double* __restrict__ xy_in, int32_t* __restrict__ cam_out,
                     int32_t* __restrict__ pt_out, double* __restrict__ xy_out) {
  __shared__ int wcnt[MAPOPS_THREADS / 32];
  const int i = blockIdx.x * MAPOPS_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const bool k = i < n && keep[i];
  const unsigned bal = __ballot_sync(0xffffffffu, k);
  if (lane == 0) wcnt[wid] = __popc(bal);
  __syncthreads();
  if (!k) return;
  int off = block_off[blockIdx.x] + __popc(bal & ((1u << lane) - 1u));
  for (int w = 0; w < wid; ++w) off += wcnt[w];
  cam_out[off] = cam_in[i];
  pt_out[off] = pt_in[i];
  reinterpret_cast<double2*>(xy_out)[off] = reinterpret_cast<const double2*>(xy_in)[i];
}

}  // namespace vo

extern "C" int vo_map_extend(const int32_t* q, const int32_t* t, const uint8_t* mask,
                             const int32_t* n_match, int frame1, int frame2, int n_cap,
                             const int32_t* root, int32_t* root_point, const float* kp,
                             const double* X, int x_stride, int cam1_row, int cam2_row,
                             double max_norm, double* points, int32_t* point_fid,
                             int32_t* n_points, int cap_points, double* obs_xy, int32_t* obs_cam,
                             int32_t* obs_pt, int32_t* n_obs, int cap_obs, int publish,
                             int32_t* status, void* stream) {
  VO_REQUIRE(q && t && n_match && root && root_point && kp && X && points && point_fid &&
                 n_points && obs_xy && obs_cam && obs_pt && n_obs && status,
             VO_ERR_INVALID_ARG, "vo_map_extend: null pointer");
  VO_REQUIRE(n_cap > 0 && frame1 >= 0 && frame2 >= 0 && x_stride >= 3 && cap_points >= 0 &&
                 cap_obs >= 0 && cam1_row >= 0 && cam2_row >= 0,
             VO_ERR_INVALID_ARG, "vo_map_extend: bad sizes");
  map_extend_kernel<<<1, EXT_THREADS, 0, (cudaStream_t)stream>>>(
      q, t, mask, n_match, frame1, frame2, n_cap, root, root_point, kp, X, x_stride, cam1_row,
      cam2_row, max_norm, points, point_fid, n_points, cap_points, obs_xy, obs_cam, obs_pt, n_obs,
      cap_obs, publish, status);
  return check_launch("vo_map_extend");
}

extern "C" size_t vo_map_prune_workspace_bytes(int n_obs) {
  const size_t nb = (size_t)((n_obs > 0 ? n_obs : 0) + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  return nb * sizeof(int32_t) + 256;
}

extern "C" int vo_map_prune_observations(const uint8_t* keep, int n_obs, const int32_t* obs_cam_in,
                                         const int32_t* obs_pt_in, const double* obs_xy_in,
                                         int32_t* obs_cam_out, int32_t* obs_pt_out,
                                         double* obs_xy_out, int32_t* n_out, void* ws,
                                         size_t ws_bytes, void* stream) {
  VO_REQUIRE(n_obs >= 0 && n_out, VO_ERR_INVALID_ARG, "vo_map_prune_observations: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_obs == 0) {
    cudaMemsetAsync(n_out, 0, sizeof(int32_t), st);
    return check_launch("vo_map_prune_observations");
  }
  VO_REQUIRE(keep && obs_cam_in && obs_pt_in && obs_xy_in && obs_cam_out && obs_pt_out &&
                 obs_xy_out && ws,
             VO_ERR_INVALID_ARG, "vo_map_prune_observations: null pointer");
  VO_REQUIRE(obs_cam_in != obs_cam_out && obs_pt_in != obs_pt_out && obs_xy_in != obs_xy_out,
             VO_ERR_INVALID_ARG, "vo_map_prune_observations: output must not alias the input");
  VO_REQUIRE(ws_bytes >= vo_map_prune_workspace_bytes(n_obs), VO_ERR_WORKSPACE,
             "vo_map_prune_observations: workspace too small");
  const int nb = (n_obs + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  int32_t* cnt = (int32_t*)ws;
  prune_count_kernel<<<nb, MAPOPS_THREADS, 0, st>>>(keep, n_obs, cnt);
  prune_scan_kernel<<<1, 1024, 0, st>>>(cnt, nb, n_out);
  prune_scatter_kernel<<<nb, MAPOPS_THREADS, 0, st>>>(keep, n_obs, cnt, obs_cam_in, obs_pt_in,
                                                     obs_xy_in, obs_cam_out, obs_pt_out, obs_xy_out);
  return check_launch("vo_map_prune_observations");
}

// ---------------------------------------------------------------------------------------------
// vo_map_remove_camera = Map.remove_camera_from_map (reference src/map.py:188-232), what
// limit_number_of_camera_in_map(18) (:301-318, called at src/visual_slam.py:307) does to the
// oldest camera: drop the camera and its observations, drop every point that is left with exactly
// ONE observation (points left with none are kept: the reference only looks at points that still
// occur in an observation), drop the observations of the dropped points; everything else keeps
// its order.  Rows are renumbered (cameras above the removed one move down by one, points are
// compacted), and root_point[] (mappointdict) is rebuilt for the new point rows.
namespace vo {

__global__ void rc_count_kernel(const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
                                int n_obs, int cam, int32_t* cnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_obs && obs_cam[i] != cam) atomicAdd(&cnt[obs_pt[i]], 1);
}

__global__ void rc_flags_kernel(const int32_t* __restrict__ cnt, int n_points,
                                const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
                                int n_obs, int cam, uint8_t* __restrict__ keep_pt,
                                uint8_t* __restrict__ keep_obs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_points) keep_pt[i] = cnt[i] != 1;
  if (i < n_obs) keep_obs[i] = obs_cam[i] != cam && cnt[obs_pt[i]] != 1;
}

__global__ void __launch_bounds__(MAPOPS_THREADS)
rc_scatter_points_kernel(const uint8_t* __restrict__ keep, int n, const int32_t* __restrict__ block_off,
                         const double* __restrict__ pts_in, const int32_t* __restrict__ fid_in,
                         double* __restrict__ pts_out, int32_t* __restrict__ fid_out,
                         int32_t* __restrict__ new_row, int32_t* __restrict__ root_point) {
  __shared__ int wcnt[MAPOPS_THREADS / 32];
  const int i = blockIdx.x * MAPOPS_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const bool k = i < n && keep[i];
  const unsigned bal = __ballot_sync(0xffffffffu, k);
  if (lane == 0) wcnt[wid] = __popc(bal);
  __syncthreads();
  if (i >= n) return;
  root_point[fid_in[i]] = -1;          // mappointdict is rebuilt from the surviving points below
  if (!k) { new_row[i] = -1; return; }
  int off = block_off[blockIdx.x] + __popc(bal & ((1u << lane) - 1u));
  for (int w = 0; w < wid; ++w) off += wcnt[w];
  new_row[i] = off;
  pts_out[3 * (size_t)off] = pts_in[3 * (size_t)i];
  pts_out[3 * (size_t)off + 1] = pts_in[3 * (size_t)i + 1];
  pts_out[3 * (size_t)off + 2] = pts_in[3 * (size_t)i + 2];
  fid_out[off] = fid_in[i];
}

__global__ void rc_publish_kernel(const int32_t* __restrict__ fid, const int32_t* __restrict__ n_ptr,
                                  int32_t* root_point) {
  // mappointdict[point.feature_id] = point in list order: the LAST point with an id wins
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < *n_ptr) atomicMax(&root_point[fid[i]], i);
}

__global__ void __launch_bounds__(MAPOPS_THREADS)
rc_scatter_obs_kernel(const uint8_t* __restrict__ keep, int n, const int32_t* __restrict__ block_off,
                      int cam, const int32_t* __restrict__ new_row, const int32_t* __restrict__ cam_in,
                      const int32_t* __restrict__ pt_in, const double* __restrict__ xy_in,
                      int32_t* __restrict__ cam_out, int32_t* __restrict__ pt_out,
                      double* __restrict__ xy_out) {
  __shared__ int wcnt[MAPOPS_THREADS / 32];
  const int i = blockIdx.x * MAPOPS_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const bool k = i < n && keep[i];
  const unsigned bal = __ballot_sync(0xffffffffu, k);
  if (lane == 0) wcnt[wid] = __popc(bal);
  __syncthreads();
  if (!k) return;
  int off = block_off[blockIdx.x] + __popc(bal & ((1u << lane) - 1u));
  for (int w = 0; w < wid; ++w) off += wcnt[w];
  const int c = cam_in[i];
  cam_out[off] = c - (c > cam);
  pt_out[off] = new_row[pt_in[i]];
  reinterpret_cast<double2*>(xy_out)[off] = reinterpret_cast<const double2*>(xy_in)[i];
}

__global__ void __launch_bounds__(512)
rc_shift_cameras_kernel(double* poses, uint8_t* fixed, int n_cams, int cam) {
  // rows above `cam` move down by one; one block, values pass through registers
  const int t = threadIdx.x;
  const int n = (n_cams - 1 - cam) * 16;
  double v = 0.0;
  uint8_t f = 0;
  if (t < n) v = poses[(cam + 1) * 16 + t];
  if (t < n_cams - 1 - cam) f = fixed[cam + 1 + t];
  __syncthreads();
  if (t < n) poses[cam * 16 + t] = v;
  if (t < n_cams - 1 - cam) fixed[cam + t] = f;
}

}  // namespace vo

extern "C" size_t vo_map_remove_camera_workspace_bytes(int n_points, int n_obs) {
  const size_t P = n_points > 0 ? n_points : 0, M = n_obs > 0 ? n_obs : 0;
  const size_t nbp = (P + MAPOPS_THREADS - 1) / MAPOPS_THREADS, nbo = (M + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  auto up = [](size_t n) { return (n + 255) / 256 * 256; };
  return up(P * 4) + up(P * 4) + up(P) + up(M) + up(nbp * 4 + 4) + up(nbo * 4 + 4) + 512;
}

extern "C" int vo_map_remove_camera(int cam_row, double* poses, uint8_t* fixed, int n_cams,
                                    const double* points_in, const int32_t* fid_in, int n_points,
                                    double* points_out, int32_t* fid_out, int32_t* n_points_out,
                                    const int32_t* obs_cam_in, const int32_t* obs_pt_in,
                                    const double* obs_xy_in, int n_obs, int32_t* obs_cam_out,
                                    int32_t* obs_pt_out, double* obs_xy_out, int32_t* n_obs_out,
                                    int32_t* root_point, void* ws, size_t ws_bytes, void* stream) {
  VO_REQUIRE(n_cams > 0 && n_cams <= 32 && cam_row >= 0 && cam_row < n_cams && n_points >= 0 && n_obs >= 0,
             VO_ERR_INVALID_ARG, "vo_map_remove_camera: bad sizes");
  VO_REQUIRE(poses && fixed && n_points_out && n_obs_out && root_point && ws, VO_ERR_INVALID_ARG,
             "vo_map_remove_camera: null pointer");
  VO_REQUIRE(n_points == 0 || (points_in && fid_in && points_out && fid_out && points_in != points_out),
             VO_ERR_INVALID_ARG, "vo_map_remove_camera: point arrays");
  VO_REQUIRE(n_obs == 0 || (obs_cam_in && obs_pt_in && obs_xy_in && obs_cam_out && obs_pt_out &&
                            obs_xy_out && obs_cam_in != obs_cam_out),
             VO_ERR_INVALID_ARG, "vo_map_remove_camera: observation arrays");
  VO_REQUIRE(ws_bytes >= vo_map_remove_camera_workspace_bytes(n_points, n_obs), VO_ERR_WORKSPACE,
             "vo_map_remove_camera: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int P = n_points, M = n_obs;
  const int nbp = (P + MAPOPS_THREADS - 1) / MAPOPS_THREADS, nbo = (M + MAPOPS_THREADS - 1) / MAPOPS_THREADS;
  char* p = (char*)(((uintptr_t)ws + 255) / 256 * 256);
  auto take = [&](size_t n) { char* r = p; p += (n + 255) / 256 * 256; return (void*)r; };
  int32_t* cnt = (int32_t*)take((size_t)P * 4);
  int32_t* new_row = (int32_t*)take((size_t)P * 4);
  uint8_t* keep_pt = (uint8_t*)take(P);
  uint8_t* keep_obs = (uint8_t*)take(M);
  int32_t* off_pt = (int32_t*)take((size_t)nbp * 4 + 4);
  int32_t* off_obs = (int32_t*)take((size_t)nbo * 4 + 4);
  if (P > 0) cudaMemsetAsync(cnt, 0, (size_t)P * 4, st);
  if (M > 0) rc_count_kernel<<<nbo, MAPOPS_THREADS, 0, st>>>(obs_cam_in, obs_pt_in, M, cam_row, cnt);
  const int nmax = P > M ? P : M;
  if (nmax > 0)
    rc_flags_kernel<<<(nmax + 255) / 256, 256, 0, st>>>(cnt, P, obs_cam_in, obs_pt_in, M, cam_row, keep_pt,
                                                       keep_obs);
  if (P > 0) {
    prune_count_kernel<<<nbp, MAPOPS_THREADS, 0, st>>>(keep_pt, P, off_pt);
    prune_scan_kernel<<<1, 1024, 0, st>>>(off_pt, nbp, n_points_out);
    rc_scatter_points_kernel<<<nbp, MAPOPS_THREADS, 0, st>>>(keep_pt, P, off_pt, points_in, fid_in, points_out,
                                                            fid_out, new_row, root_point);
    // every surviving point publishes its new row (the new count stays on the device)
    rc_publish_kernel<<<nbp, MAPOPS_THREADS, 0, st>>>(fid_out, n_points_out, root_point);
  } else {
    cudaMemsetAsync(n_points_out, 0, 4, st);
  }
  if (M > 0) {
    prune_count_kernel<<<nbo, MAPOPS_THREADS, 0, st>>>(keep_obs, M, off_obs);
    prune_scan_kernel<<<1, 1024, 0, st>>>(off_obs, nbo, n_obs_out);
    rc_scatter_obs_kernel<<<nbo, MAPOPS_THREADS, 0, st>>>(keep_obs, M, off_obs, cam_row, new_row, obs_cam_in,
                                                         obs_pt_in, obs_xy_in, obs_cam_out, obs_pt_out,
                                                         obs_xy_out);
  } else {
    cudaMemsetAsync(n_obs_out, 0, 4, st);
  }
  rc_shift_cameras_kernel<<<1, 512, 0, st>>>(poses, fixed, n_cams, cam_row);
  return check_launch("vo_map_remove_camera");
}
