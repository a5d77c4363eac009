// pose.cu — K4 DLT triangulation, K5 recoverPose cheirality, trajectory chaining.
//
// Replaces cv2.recoverPose(E, p1, p2, K) (reference src/image_pair.py:304-308) and
// cv2.triangulatePoints(P, P0, p1.T, p2.T) + dehomogenisation (src/image_pair.py:332-339).
// Both are batched per-point 4x4 null-vector solves (one-sided Jacobi in registers);
// recoverPose = 3x3 SVD of E -> four [R|t] candidates -> DLT of every inlier under each
// candidate -> OpenCV's three depth tests -> first maximum of the four counts.
#include "linalg.cuh"

namespace vo {

constexpr int POSE_THREADS = 256;   // two blocks per SM: one block's serial head (3x3 SVD) overlaps another's points (512: 0.168, 256: 0.160, 128: 0.162 ms)

__global__ void __launch_bounds__(POSE_THREADS)
recover_pose_kernel(const double* __restrict__ E, const double* __restrict__ pts,
                    const int32_t* __restrict__ m, const uint8_t* __restrict__ mask, int n_cap,
                    const double* __restrict__ K, double dist, double* __restrict__ R_out,
                    double* __restrict__ t_out, uint8_t* __restrict__ pose_mask,
                    int32_t* __restrict__ n_good, uint8_t* __restrict__ flags) {
  __shared__ double sP[4][12];  // the four candidate [R|t]
  __shared__ double sR[2][9], st[3];
  __shared__ int cnt[4];
  __shared__ int win;
  const int p = blockIdx.x;
  const int tid = threadIdx.x;
  const int mp = m[p];
  if (tid == 0) {
    decompose_essential(E + (size_t)p * 9, sR[0], sR[1], st);
    for (int c = 0; c < 4; ++c) {
      const double* R = sR[c & 1];
      const double sg = (c < 2) ? 1.0 : -1.0;
      for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) sP[c][4 * i + j] = R[3 * i + j];
        sP[c][4 * i + 3] = sg * st[i];
      }
    }
  }
  if (tid < 4) cnt[tid] = 0;
  __syncthreads();
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  uint8_t* fl = flags + (size_t)p * n_cap;
  int lp = 0, lm = 0;   // this thread's rotation index is fixed: c = tid & 1
  // work item = (point, rotation c in {0,1}): candidates c and c+2 differ only in the sign of t,
  // and [R|-t] has the DLT matrix of [R|t] with its fourth column negated, so its null vector is
  // (Q0,Q1,Q2,-Q3): ONE 4x4 SVD serves both.  2 consecutive threads share a point.
  const int n_items = 2 * mp;
  for (int it0 = 0; it0 < n_items; it0 += POSE_THREADS) {
    const int it = it0 + tid;
    const int i = it >> 1, c = it & 1;
    bool good_p = false, good_m = false;
    if (it < n_items && (!mask || mask[(size_t)p * n_cap + i])) {
      const double4 v = reinterpret_cast<const double4*>(pts)[(size_t)p * n_cap + i];
      const double x1 = (v.x - cx) / fx, y1 = (v.y - cy) / fy;
      const double x2 = (v.z - cx) / fx, y2 = (v.w - cy) / fy;
      double Q[4];
      dlt_point(P0, sP[c], x1, y1, x2, y2, Q);
      const double* P = sP[c];
      {  // candidate c: [R | +t]
        good_p = (Q[2] * Q[3]) > 0;
        const double qx = Q[0] / Q[3], qy = Q[1] / Q[3], qz = Q[2] / Q[3];
        good_p = good_p && (qz < dist);
        // Q = P1 * Q with Q[3] = w/w = 1
        const double z2 = ((P[8] * qx + P[9] * qy) + P[10] * qz) + P[11] * (Q[3] / Q[3]);
        good_p = good_p && (z2 > 0) && (z2 < dist);
      }
      {  // candidate c+2: [R | -t], null vector (Q0,Q1,Q2,-Q3)
        const double w = -Q[3];
        good_m = (Q[2] * w) > 0;
        const double qx = Q[0] / w, qy = Q[1] / w, qz = Q[2] / w;
        good_m = good_m && (qz < dist);
        const double z2 = ((P[8] * qx + P[9] * qy) + P[10] * qz) + (-P[11]) * (w / w);
        good_m = good_m && (z2 > 0) && (z2 < dist);
      }
    }
    // gather the 4 candidate bits of this point into its even lane and store them
    const unsigned lane = tid & 31;
    const unsigned bal_p = __ballot_sync(0xffffffffu, good_p);
    const unsigned bal_m = __ballot_sync(0xffffffffu, good_m);
    if (c == 0 && it < n_items)
      fl[i] = (uint8_t)(((bal_p >> lane) & 3u) | (((bal_m >> lane) & 3u) << 2));
    lp += good_p;
    lm += good_m;
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    int v = ((tid & 1) == (c & 1)) ? (c < 2 ? lp : lm) : 0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0 && v) atomicAdd(&cnt[c], v);
  }
  __syncthreads();
  if (tid == 0) {
    const int g1 = cnt[0], g2 = cnt[1], g3 = cnt[2], g4 = cnt[3];
    int k;  // OpenCV's if-chain: first candidate that is >= all others
    if (g1 >= g2 && g1 >= g3 && g1 >= g4) k = 0;
    else if (g2 >= g1 && g2 >= g3 && g2 >= g4) k = 1;
    else if (g3 >= g1 && g3 >= g2 && g3 >= g4) k = 2;
    else k = 3;
    win = k;
    n_good[p] = cnt[k];
    for (int e = 0; e < 9; ++e) R_out[(size_t)p * 9 + e] = sR[k & 1][e];
    for (int e = 0; e < 3; ++e) t_out[(size_t)p * 3 + e] = (k < 2 ? 1.0 : -1.0) * st[e];
  }
  __syncthreads();
  if (pose_mask) {
    const int k = win;
    for (int i = tid; i < n_cap; i += POSE_THREADS) {
      uint8_t v = 0;
      if (i < mp && (!mask || mask[(size_t)p * n_cap + i])) v = ((fl[i] >> k) & 1) ? 255 : 0;
      pose_mask[(size_t)p * n_cap + i] = v;
    }
  }
}

__global__ void triangulate_kernel(const double* __restrict__ P1, const double* __restrict__ P2,
                                   const double* __restrict__ x1, const double* __restrict__ x2,
                                   int n, double* __restrict__ X, int dehomog) {
  __shared__ double sP[24];
  if (threadIdx.x < 24) sP[threadIdx.x] = threadIdx.x < 12 ? P1[threadIdx.x] : P2[threadIdx.x - 12];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double2 a = reinterpret_cast<const double2*>(x1)[i];
  const double2 b = reinterpret_cast<const double2*>(x2)[i];
  double Q[4];
  dlt_point(sP, sP + 12, a.x, a.y, b.x, b.y, Q);
  if (dehomog) {
    const double w = Q[3];
    Q[0] /= w; Q[1] /= w; Q[2] /= w; Q[3] /= w;
  }
  double2* o = reinterpret_cast<double2*>(X) + 2 * (size_t)i;
  o[0] = make_double2(Q[0], Q[1]);
  o[1] = make_double2(Q[2], Q[3]);
}

// reference flavour (image_pair.py:316-339): P_first = K [R^T | -R^T t] with frame-1 points,
// P_second = K [I | 0] with frame-2 points; inliers compacted in match order; X /= w.
constexpr int TRI_THREADS = 256;
__global__ void __launch_bounds__(TRI_THREADS)
triangulate_pairs_kernel(const double* __restrict__ R, const double* __restrict__ t,
                         const double* __restrict__ pts, const int32_t* __restrict__ m,
                         const uint8_t* __restrict__ mask, int n_cap, const double* __restrict__ K,
                         double* __restrict__ X, int32_t* __restrict__ n_x) {
  __shared__ double sP[24];
  __shared__ int warp_tot[TRI_THREADS / 32];
  __shared__ int base_s;
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid < 12) {
    // estimated_transform = [R^T | -R^T t]; row i, col j
    const int i = tid / 4, j = tid % 4;
    const double* Rp = R + (size_t)p * 9;
    const double* tp = t + (size_t)p * 3;
    double T[3][4];
    for (int a = 0; a < 3; ++a) {
      for (int b = 0; b < 3; ++b) T[a][b] = Rp[3 * b + a];
      T[a][3] = -((Rp[a] * tp[0] + Rp[3 + a] * tp[1]) + Rp[6 + a] * tp[2]);
    }
    sP[tid] = (K[3 * i] * T[0][j] + K[3 * i + 1] * T[1][j]) + K[3 * i + 2] * T[2][j];
    sP[12 + tid] = (j < 3) ? K[3 * i + j] : 0.0;
  }
  if (tid == 0) base_s = 0;
  __syncthreads();
  const int mp = m[p];
  for (int i0 = 0; i0 < mp; i0 += TRI_THREADS) {
    const int i = i0 + tid;
    const bool keep = i < mp && (!mask || mask[(size_t)p * n_cap + i]);
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_tot[wid] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < wid; ++w) off += warp_tot[w];
    off += __popc(bal & ((1u << lane) - 1u));
    if (keep) {
      const double4 v = reinterpret_cast<const double4*>(pts)[(size_t)p * n_cap + i];
      double Q[4];
      dlt_point(sP, sP + 12, v.x, v.y, v.z, v.w, Q);
      const double w = Q[3];
      double2* o = reinterpret_cast<double2*>(X) + 2 * ((size_t)p * n_cap + off);
      o[0] = make_double2(Q[0] / w, Q[1] / w);
      o[1] = make_double2(Q[2] / w, Q[3] / w);
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < TRI_THREADS / 32; ++w) tot += warp_tot[w];
      base_s += tot;
    }
    __syncthreads();
  }
  if (tid == 0) n_x[p] = base_s;
}

// P_0 given (or I); P_k = rel_k * P_{k-1}, rel_k = [R_{k-1} t_{k-1}; 0 1].
// SE(3) composition is associative, so the chain is a prefix product: every thread multiplies a
// contiguous run of CHAIN_RUN relative poses, the run totals are scanned across the block
// (Hillis-Steele in shared memory), and each thread replays its run from its carried-in prefix.
// The multiplication order inside every product is the sequential one (rel_k on the left of the
// accumulated pose), so results differ from a serial loop only by fp64 re-association (~1e-16).
constexpr int CHAIN_THREADS = 128;
struct Rt { double m[12]; };   // [R|t] row-major 3x4

// out = a * b  (a applied after b): R = Ra Rb, t = Ra tb + ta
__device__ __forceinline__ Rt rt_mul(const Rt& a, const Rt& b) {
  Rt o;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      o.m[4 * i + j] = a.m[4 * i] * b.m[j] + a.m[4 * i + 1] * b.m[4 + j] + a.m[4 * i + 2] * b.m[8 + j] +
                       (j == 3 ? a.m[4 * i + 3] : 0.0);
  return o;
}
__device__ __forceinline__ Rt rt_load(const double* __restrict__ R, const double* __restrict__ t, int k) {
  Rt o;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o.m[4 * i + j] = R[(size_t)k * 9 + 3 * i + j];
    o.m[4 * i + 3] = t[(size_t)k * 3 + i];
  }
  return o;
}

// packed rows [R row-major (9) | t (3)]: the all-gather payload of the multi-GPU path
__device__ __forceinline__ Rt rt_load_packed(const double* __restrict__ rt, int k) {
  Rt o;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o.m[4 * i + j] = rt[(size_t)k * 12 + 3 * i + j];
    o.m[4 * i + 3] = rt[(size_t)k * 12 + 9 + i];
  }
  return o;
}

__global__ void pack_poses_kernel(const double* __restrict__ R, const double* __restrict__ t, int n,
                                  double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * 12) return;
  const int k = i / 12, e = i % 12;
  out[i] = e < 9 ? R[(size_t)k * 9 + e] : t[(size_t)k * 3 + (e - 9)];
}

template <bool PACKED>
__global__ void __launch_bounds__(CHAIN_THREADS)
chain_kernel(const double* __restrict__ R, const double* __restrict__ t, int n,
             const double* __restrict__ P0, double* __restrict__ poses) {
  __shared__ Rt sc[2][CHAIN_THREADS];
  const int tid = threadIdx.x;
  const int run = (n + CHAIN_THREADS - 1) / CHAIN_THREADS;
  const int k0 = tid * run, k1 = min(n, k0 + run);
  Rt ident;
#pragma unroll
  for (int e = 0; e < 12; ++e) ident.m[e] = (e % 5 == 0) ? 1.0 : 0.0;
  // 1) product of this thread's run: rel_{k1-1} ... rel_{k0}
  Rt acc = ident;
  for (int k = k0; k < k1; ++k) acc = rt_mul(PACKED ? rt_load_packed(R, k) : rt_load(R, t, k), acc);
  sc[0][tid] = acc;
  __syncthreads();
  // 2) inclusive scan of run totals (later runs multiply on the left)
  int cur = 0;
  for (int off = 1; off < CHAIN_THREADS; off <<= 1) {
    Rt v = sc[cur][tid];
    if (tid >= off) v = rt_mul(v, sc[cur][tid - off]);
    sc[cur ^ 1][tid] = v;
    cur ^= 1;
    __syncthreads();
  }
  // 3) carried-in prefix = P0 composed with everything before this run
  Rt pre = ident;
  if (P0) {
#pragma unroll
    for (int e = 0; e < 12; ++e) pre.m[e] = P0[e];
  }
  if (tid > 0) pre = rt_mul(sc[cur][tid - 1], pre);
  if (tid == 0) {
#pragma unroll
    for (int e = 0; e < 12; ++e) poses[e] = pre.m[e];
    poses[12] = poses[13] = poses[14] = 0.0; poses[15] = 1.0;
  }
  for (int k = k0; k < k1; ++k) {
    pre = rt_mul(PACKED ? rt_load_packed(R, k) : rt_load(R, t, k), pre);
    double* o = poses + (size_t)(k + 1) * 16;
#pragma unroll
    for (int e = 0; e < 12; ++e) o[e] = pre.m[e];
    o[12] = o[13] = o[14] = 0.0; o[15] = 1.0;
  }
}

}  // namespace vo

using namespace vo;

extern "C" size_t vo_pose_workspace_bytes(int n_pairs, int n_cap) {
  return (size_t)(n_pairs > 0 ? n_pairs : 0) * (n_cap > 0 ? n_cap : 0) + 256;
}

extern "C" int vo_recover_pose_batch(const double* E, const double* pts, const int32_t* m,
                                     const uint8_t* mask, int n_cap, int n_pairs, const double* K,
                                     double dist_thresh, double* R, double* t, uint8_t* pose_mask,
                                     int32_t* n_good, void* ws, size_t ws_bytes, void* stream) {
  VO_REQUIRE(E && pts && m && K && R && t && n_good && ws, VO_ERR_INVALID_ARG,
             "vo_recover_pose_batch: null pointer");
  VO_REQUIRE(n_cap > 0 && n_pairs >= 0, VO_ERR_INVALID_ARG, "vo_recover_pose_batch: bad sizes");
  VO_REQUIRE(ws_bytes >= vo_pose_workspace_bytes(n_pairs, n_cap), VO_ERR_WORKSPACE,
             "vo_recover_pose_batch: workspace too small");
  if (n_pairs == 0) return VO_OK;
  ProfScope prof(5, (cudaStream_t)stream);
  recover_pose_kernel<<<n_pairs, POSE_THREADS, 0, (cudaStream_t)stream>>>(
      E, pts, m, mask, n_cap, K, dist_thresh, R, t, pose_mask, n_good, (uint8_t*)ws);
  return check_launch("vo_recover_pose_batch");
}

extern "C" int vo_triangulate(const double* P1, const double* P2, const double* x1,
                              const double* x2, int n, double* X, int dehomog, void* stream) {
  VO_REQUIRE(P1 && P2 && x1 && x2 && X, VO_ERR_INVALID_ARG, "vo_triangulate: null pointer");
  if (n <= 0) return VO_OK;
  triangulate_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P1, P2, x1, x2, n, X,
                                                                       dehomog);
  return check_launch("vo_triangulate");
}

extern "C" int vo_triangulate_pairs(const double* R, const double* t, const double* pts,
                                    const int32_t* m, const uint8_t* mask, int n_cap, int n_pairs,
                                    const double* K, double* X, int32_t* n_x, void* stream) {
  VO_REQUIRE(R && t && pts && m && K && X && n_x, VO_ERR_INVALID_ARG,
             "vo_triangulate_pairs: null pointer");
  if (n_pairs <= 0) return VO_OK;
  ProfScope prof(6, (cudaStream_t)stream);
  triangulate_pairs_kernel<<<n_pairs, TRI_THREADS, 0, (cudaStream_t)stream>>>(R, t, pts, m, mask,
                                                                             n_cap, K, X, n_x);
  return check_launch("vo_triangulate_pairs");
}

extern "C" int vo_chain_poses(const double* R, const double* t, int n_pairs, const double* P0,
                              double* poses, void* stream) {
  VO_REQUIRE(R && t && poses && n_pairs >= 0, VO_ERR_INVALID_ARG, "vo_chain_poses: bad arguments");
  ProfScope prof(7, (cudaStream_t)stream);
  chain_kernel<false><<<1, CHAIN_THREADS, 0, (cudaStream_t)stream>>>(R, t, n_pairs, P0, poses);
  return check_launch("vo_chain_poses");
}

extern "C" int vo_pack_poses(const double* R, const double* t, int n_pairs, double* rt, void* stream) {
  VO_REQUIRE(R && t && rt && n_pairs >= 0, VO_ERR_INVALID_ARG, "vo_pack_poses: bad arguments");
  if (n_pairs == 0) return VO_OK;
  pack_poses_kernel<<<(n_pairs * 12 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(R, t, n_pairs, rt);
  return check_launch("vo_pack_poses");
}

extern "C" int vo_chain_poses_packed(const double* rt, int n_pairs, const double* P0, double* poses,
                                     void* stream) {
  VO_REQUIRE(rt && poses && n_pairs >= 0, VO_ERR_INVALID_ARG, "vo_chain_poses_packed: bad arguments");
  ProfScope prof(7, (cudaStream_t)stream);
  chain_kernel<true><<<1, CHAIN_THREADS, 0, (cudaStream_t)stream>>>(rt, nullptr, n_pairs, P0, poses);
  return check_launch("vo_chain_poses_packed");
}
