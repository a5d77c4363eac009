// tracks.cu — N1: feature-track bookkeeping and map association on the device.
//
// Replaces the Python dict chains of the reference's VisualSlam:
//   feature_mapper[featureid2] = featureid1               (src/visual_slam.py:46-50, 183-188)
//   track_feature_back_in_time: follow feature_mapper      (src/visual_slam.py:94-99, 204-207)
//   mappointdict[point.feature_id] = point                 (src/visual_slam.py:154-156, 193-196)
//   MatchWithMap assembly for solvePnPRansac               (src/visual_slam.py:202-229)
// A feature id (frame_id, idx) (src/frame_generator.py:34-36) is the flat index
// frame * n_cap + idx.  Links always point one frame back and the older feature's chain is
// complete when the link is made, so the root is carried forward at insertion time
// (root[b] = root[a]) and "track back in time" is one load instead of a chain walk.
// The gather kernel emits, in match order (stable compaction), the image / map coordinates of
// every current match whose track root owns a map point: PnP's inputs never leave the GPU.
#include "common.cuh"

namespace vo {

__global__ void tracks_reset_kernel(int32_t* __restrict__ parent, int32_t* __restrict__ root,
                                    int32_t* __restrict__ root_point, long n) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    parent[i] = -1;
    root[i] = (int32_t)i;
    root_point[i] = -1;
  }
}

// The reference inserts the links sequentially (feature_mapper[featureid2] = featureid1 in match
// order), so when several matches share a train feature (knn + ratio mode; never with the
// cross-check) the LAST one wins.  Two passes keep that deterministic: pass 0 elects the highest
// match index per train feature (atomicMax of a tagged index: the tag bit puts it above -1 and
// above any feature id), pass 1 lets the elected match write parent and root.
constexpr int32_t TRACK_TAG = 0x40000000;
__global__ void tracks_update_kernel(const int32_t* __restrict__ q, const int32_t* __restrict__ t,
                                     const uint8_t* __restrict__ mask,
                                     const int32_t* __restrict__ n_match, int f1, int f2, int n_cap,
                                     int pass, int32_t* __restrict__ parent, int32_t* __restrict__ root) {
  const int n = *n_match;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    const int a = f1 * n_cap + q[i], b = f2 * n_cap + t[i];
    if (pass == 0) {
      atomicMax(&parent[b], TRACK_TAG | i);
    } else if (parent[b] == (TRACK_TAG | i)) {
      parent[b] = a;
      root[b] = root[a];
    }
  }
}

__global__ void tracks_set_points_kernel(const int32_t* __restrict__ feature_ids,
                                         const int32_t* __restrict__ point_ids, int n,
                                         int32_t* __restrict__ root_point) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) root_point[feature_ids[i]] = point_ids ? point_ids[i] : i;
}

constexpr int TM_THREADS = 1024;
__global__ void __launch_bounds__(TM_THREADS)
tracks_match_map_kernel(const int32_t* __restrict__ q, const int32_t* __restrict__ t,
                        const uint8_t* __restrict__ mask, const int32_t* __restrict__ n_match,
                        int f1, int f2, int n_cap, const int32_t* __restrict__ root,
                        const int32_t* __restrict__ root_point, const float* __restrict__ kp,
                        const double* __restrict__ map_points, double* __restrict__ img_xy,
                        double* __restrict__ map_xyz, int32_t* __restrict__ feat2,
                        int32_t* __restrict__ point_row, int32_t* __restrict__ n_out) {
  __shared__ int warp_tot[TM_THREADS / 32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int n = *n_match;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += TM_THREADS) {
    const int i = i0 + tid;
    int b = -1, prow = -1;
    if (i < n && (!mask || mask[i])) {
      b = f2 * n_cap + t[i];
      prow = root_point[root[b]];       // track back in time, then mappointdict lookup
    }
    const bool keep = prow >= 0;
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_tot[wid] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < wid; ++w) off += warp_tot[w];
    off += __popc(bal & ((1u << lane) - 1u));
    if (keep) {
      const float2 p2 = reinterpret_cast<const float2*>(kp)[(size_t)f2 * n_cap + t[i]];
      img_xy[2 * off] = (double)p2.x;
      img_xy[2 * off + 1] = (double)p2.y;
      map_xyz[3 * off] = map_points[3 * (size_t)prow];
      map_xyz[3 * off + 1] = map_points[3 * (size_t)prow + 1];
      map_xyz[3 * off + 2] = map_points[3 * (size_t)prow + 2];
      if (feat2) feat2[off] = b;
      if (point_row) point_row[off] = prow;
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < TM_THREADS / 32; ++w) tot += warp_tot[w];
      base_s += tot;
    }
    __syncthreads();
  }
  if (tid == 0) *n_out = base_s;
}

}  // namespace vo

using namespace vo;

extern "C" int vo_tracks_reset(int32_t* parent, int32_t* root, int32_t* root_point,
                               long n_features, void* stream) {
  VO_REQUIRE(parent && root && root_point && n_features >= 0 && n_features < (1L << 30),
             VO_ERR_INVALID_ARG, "vo_tracks_reset: bad arguments");
  if (n_features == 0) return VO_OK;
  const int blocks = (int)((n_features + 255) / 256 < 1184 ? (n_features + 255) / 256 : 1184);
  tracks_reset_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(parent, root, root_point, n_features);
  return check_launch("vo_tracks_reset");
}

extern "C" int vo_tracks_update(const int32_t* q, const int32_t* t, const uint8_t* mask,
                                const int32_t* n_match, int frame1, int frame2, int n_cap,
                                int32_t* parent, int32_t* root, void* stream) {
  VO_REQUIRE(q && t && n_match && parent && root && n_cap > 0 && frame1 >= 0 && frame2 >= 0,
             VO_ERR_INVALID_ARG, "vo_tracks_update: bad arguments");
  const int blocks = (n_cap + 255) / 256;
  for (int pass = 0; pass < 2; ++pass)
    tracks_update_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(q, t, mask, n_match, frame1, frame2,
                                                                 n_cap, pass, parent, root);
  return check_launch("vo_tracks_update");
}

extern "C" int vo_tracks_set_points(const int32_t* feature_ids, const int32_t* point_rows, int n,
                                    int32_t* root_point, void* stream) {
  VO_REQUIRE(root_point && n >= 0, VO_ERR_INVALID_ARG, "vo_tracks_set_points: bad arguments");
  if (n == 0) return VO_OK;
  VO_REQUIRE(feature_ids, VO_ERR_INVALID_ARG, "vo_tracks_set_points: null feature ids");
  tracks_set_points_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(feature_ids, point_rows,
                                                                            n, root_point);
  return check_launch("vo_tracks_set_points");
}

extern "C" int vo_tracks_match_map(const int32_t* q, const int32_t* t, const uint8_t* mask,
                                   const int32_t* n_match, int frame1, int frame2, int n_cap,
                                   const int32_t* root, const int32_t* root_point, const float* kp,
                                   const double* map_points, double* img_xy, double* map_xyz,
                                   int32_t* feat2, int32_t* point_row, int32_t* n_out,
                                   void* stream) {
  VO_REQUIRE(q && t && n_match && root && root_point && kp && map_points && img_xy && map_xyz &&
                 n_out && n_cap > 0,
             VO_ERR_INVALID_ARG, "vo_tracks_match_map: bad arguments");
  tracks_match_map_kernel<<<1, TM_THREADS, 0, (cudaStream_t)stream>>>(
      q, t, mask, n_match, frame1, frame2, n_cap, root, root_point, kp, map_points, img_xy, map_xyz,
      feat2, point_row, n_out);
  return check_launch("vo_tracks_match_map");
}
