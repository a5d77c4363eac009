// ba.cu — N3: local bundle adjustment on the flat device map.
//
// Replaces Map.optimize_map (reference src/map.py:104-186): g2o SparseOptimizer with
// BlockSolverSE3 + LinearSolverEigenSE3 under OptimizationAlgorithmLevenberg, VertexSE3Expmap per
// camera (fixed flag), marginalised VertexPointXYZ per point, EdgeProjectXYZ2UV + Huber per
// observation, optimize(40).  g2o is a third-party dependency absent from the reference tree; the
// algorithm is restated from its published sources (see oracle/ba_restate.py for the list):
//   error      e = uv - (f * (R X + t).xy / z + c)              (one focal length, cam_map)
//   robust     Huber on chi2 = |e|^2: rho0, weight rho1 (first-order only, as g2o)
//   system     Hpp (6x6 per free camera), Hll (3x3 per point), Hpl (6x3 per observation), b
//   LM trial   (H + lambda I) dx = b by the Schur complement on the points, dense Cholesky of the
//              reduced camera system (<= 24 cameras), back-substitution for the points
//   update     T <- exp(dx_pose) * T  (omega first), X <- X + dx_point
//   control    g2o's lambda policy: init tau*max diag, rho = dChi / (dx.(lambda dx + b) + 1e-3),
//              good step: lambda *= max(1/3, min(2/3, 1-(2rho-1)^3)); bad: lambda *= ni, ni *= 2,
//              <= 10 trials per iteration
// Every sum runs in a fixed order (lists sorted by observation index, tree reductions, no
// floating-point atomics), so results are reproducible run to run.  The accept / reject decision
// is taken on the host from one 32-byte read-back per trial: this entry point synchronises its
// stream (documented in include/vo_b200.h), unlike the rest of the library.
#include <float.h>
#include <math.h>
#include <string.h>

#include "common.cuh"

namespace vo {

constexpr int BA_THREADS = 256;
constexpr int BA_MAX_CAMS = 24;
constexpr int BA_LM_MAX_TRIALS = 10;
constexpr int BA_PT_LANES = 4;            // lanes that share one point's observation list

constexpr int BA_TARGET_BLOCKS = 296;     // 2 x 148 SMs: how far the per-camera loops are split
constexpr int BA_MAX_SPLIT = 64;
constexpr int BA_SPART_ROWS = BA_TARGET_BLOCKS + BA_MAX_CAMS * (BA_MAX_CAMS + 1) / 2 + 8;
constexpr int BA_CPART_ROWS = BA_TARGET_BLOCKS + BA_MAX_CAMS + 8;

struct BAWs {
  int32_t *pt_start, *pt_cursor, *pt_obs, *cam_start, *cam_obs, *cam_count, *bad, *flist;
  double *OA, *OB, *OW;                   // per-observation records [M,10] [M,12] [M,18]
  double *Hll, *bl, *Hpp, *bp;            // [P,6] [P,3] [C,21] [C,6]
  double *Spart, *Cpart, *xp;             // [pairs*NS,42] [nf*NSc,27] [6C]
  double *trial_points, *trial_poses;     // [P,3] [C,16]
  double *chi_part, *scale_part, *stats;  // block partials; stats[0..7]
  unsigned long long* maxdiag;
  unsigned int* ticket;
  size_t bytes;
};

static BAWs ba_carve(void* base, int C, int P, int M) {
  BAWs w;
  char* p = (char*)base;
  auto take = [&](size_t n) { char* r = p; p += (n + 255) / 256 * 256; return (void*)r; };
  const int nbo = (M + BA_THREADS - 1) / BA_THREADS, nbp = (P + BA_THREADS - 1) / BA_THREADS;
  w.pt_start = (int32_t*)take(sizeof(int32_t) * (P + 1));
  w.pt_cursor = (int32_t*)take(sizeof(int32_t) * (P + 1));
  w.pt_obs = (int32_t*)take(sizeof(int32_t) * (size_t)M);
  w.cam_start = (int32_t*)take(sizeof(int32_t) * (C + 1));
  w.cam_obs = (int32_t*)take(sizeof(int32_t) * (size_t)M);
  w.cam_count = (int32_t*)take(sizeof(int32_t) * (C + 1));
  w.bad = (int32_t*)take(sizeof(int32_t));
  w.flist = (int32_t*)take(sizeof(int32_t) * (C + 1));
  w.OA = (double*)take(sizeof(double) * 10 * (size_t)M);
  w.OB = (double*)take(sizeof(double) * 12 * (size_t)M);
  w.OW = (double*)take(sizeof(double) * 18 * (size_t)M);
  w.Hll = (double*)take(sizeof(double) * 6 * (size_t)P);
  w.bl = (double*)take(sizeof(double) * 3 * (size_t)P);
  w.Hpp = (double*)take(sizeof(double) * 21 * C);
  w.bp = (double*)take(sizeof(double) * 6 * C);
  w.Spart = (double*)take(sizeof(double) * 42 * BA_SPART_ROWS);
  w.Cpart = (double*)take(sizeof(double) * 27 * BA_CPART_ROWS);
  w.xp = (double*)take(sizeof(double) * 6 * C);
  w.trial_points = (double*)take(sizeof(double) * 3 * (size_t)P);
  w.trial_poses = (double*)take(sizeof(double) * 16 * C);
  w.chi_part = (double*)take(sizeof(double) * (nbo + 1));
  w.scale_part = (double*)take(sizeof(double) * ((size_t)nbp * BA_PT_LANES + 2));
  w.stats = (double*)take(sizeof(double) * 8);
  w.maxdiag = (unsigned long long*)take(sizeof(unsigned long long));
  w.ticket = (unsigned int*)take(sizeof(unsigned int));
  w.bytes = (size_t)(p - (char*)base);
  return w;
}

// ------------------------------------------------------------------ adjacency lists (once)
__global__ void ba_count_kernel(const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
                                int M, int C, int P, int32_t* pt_count, int32_t* cam_count,
                                int32_t* bad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  const int c = obs_cam[i], p = obs_pt[i];
  if (c < 0 || c >= C || p < 0 || p >= P) { *bad = 1; return; }
  atomicAdd(&pt_count[p], 1);
  atomicAdd(&cam_count[c], 1);
}

__global__ void __launch_bounds__(1024)
ba_scan_kernel(int32_t* __restrict__ pt_start /* in: counts, out: exclusive scan [P+1] */, int P,
               int32_t* __restrict__ pt_cursor, const int32_t* __restrict__ cam_count,
               int32_t* __restrict__ cam_start, int C, const uint8_t* __restrict__ fixed,
               int32_t* __restrict__ flist /* [0] = count, then the camera rows */) {
  __shared__ int wsum[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {
    base_s = 0;
    int s = 0;
    int nf = 0;
    for (int c = 0; c < C; ++c) {
      cam_start[c] = s;
      s += cam_count[c];
      if (!fixed[c] && cam_count[c] > 0) flist[1 + nf++] = c;
    }
    cam_start[C] = s;
    flist[0] = nf;
  }
  __syncthreads();
  for (int b0 = 0; b0 <= P; b0 += 1024) {
    const int b = b0 + tid;
    const int v = b < P ? pt_start[b] : 0;
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
    if (b <= P) { pt_start[b] = off + s - v; pt_cursor[b] = off + s - v; }
    __syncthreads();
    if (tid == 1023) base_s = off + s;
    __syncthreads();
  }
}

__global__ void ba_fill_pts_kernel(const int32_t* __restrict__ obs_pt, int M, int32_t* pt_cursor,
                                   int32_t* __restrict__ pt_obs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < M) pt_obs[atomicAdd(&pt_cursor[obs_pt[i]], 1)] = i;
}

__global__ void ba_sort_pts_kernel(const int32_t* __restrict__ pt_start, int P, int32_t* pt_obs) {
  // insertion sort of each (short) list by observation index: a fixed summation order
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int s = pt_start[p], e = pt_start[p + 1];
  for (int a = s + 1; a < e; ++a) {
    const int v = pt_obs[a];
    int b = a - 1;
    while (b >= s && pt_obs[b] > v) { pt_obs[b + 1] = pt_obs[b]; --b; }
    pt_obs[b + 1] = v;
  }
}

__global__ void __launch_bounds__(1024)
ba_fill_cams_kernel(const int32_t* __restrict__ obs_cam, int M, const int32_t* __restrict__ cam_start,
                    int32_t* __restrict__ cam_obs) {
  // block = camera: stable compaction of the observation indices of that camera
  __shared__ int wcnt[32];
  __shared__ int base_s;
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) base_s = cam_start[c];
  __syncthreads();
  for (int i0 = 0; i0 < M; i0 += 1024) {
    const int i = i0 + tid;
    const bool k = i < M && obs_cam[i] == c;
    const unsigned bal = __ballot_sync(0xffffffffu, k);
    if (lane == 0) wcnt[wid] = __popc(bal);
    __syncthreads();
    int off = base_s + __popc(bal & ((1u << lane) - 1u));
    for (int w = 0; w < wid; ++w) off += wcnt[w];
    if (k) cam_obs[off] = i;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += wcnt[w];
      base_s += t;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ per-observation arithmetic
struct Proj { double x, y, z, ex, ey; };

__device__ __forceinline__ Proj ba_project(const double* __restrict__ T, const double* __restrict__ X,
                                           double u, double v, double f, double cx, double cy) {
  Proj r;
  r.x = T[0] * X[0] + T[1] * X[1] + T[2] * X[2] + T[3];
  r.y = T[4] * X[0] + T[5] * X[1] + T[6] * X[2] + T[7];
  r.z = T[8] * X[0] + T[9] * X[1] + T[10] * X[2] + T[11];
  r.ex = u - (r.x / r.z * f + cx);
  r.ey = v - (r.y / r.z * f + cy);
  return r;
}

__device__ __forceinline__ void ba_huber(double e2, double delta, double& rho0, double& w) {
  if (delta <= 0.0 || e2 <= delta * delta) { rho0 = e2; w = 1.0; return; }
  const double s = sqrt(e2);
  rho0 = 2.0 * s * delta - delta * delta;
  w = delta / s;
}

__device__ __forceinline__ double block_sum_ordered(double v, double* red) {
  // shuffle tree inside each warp, warps added in order by thread 0 (valid in thread 0)
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
  __syncthreads();
  return s;
}

template <int N>
__device__ __forceinline__ void block_sum_ordered_n(double (&acc)[N], double* red /* [8][N] smem */) {
  // N sums at once: shuffle tree per warp, one barrier, then threads 0..N-1 add the warps in
  // order; the totals come back in acc[] of thread k for value k (k < N)
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double v = acc[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[(threadIdx.x >> 5) * N + k] = v;
  }
  __syncthreads();
}

template <int N2>
__device__ __forceinline__ void load_rec(const double* __restrict__ src, double* dst) {
  // 16-byte vector loads of one observation record (records are 16 B aligned)
  const double2* s2 = reinterpret_cast<const double2*>(src);
#pragma unroll
  for (int k = 0; k < N2; ++k) { const double2 v = s2[k]; dst[2 * k] = v.x; dst[2 * k + 1] = v.y; }
}

__global__ void __launch_bounds__(BA_THREADS)
ba_chi_kernel(const double* __restrict__ poses, const double* __restrict__ points,
              const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
              const double* __restrict__ obs_xy, int M, double f, double cx, double cy,
              double delta, double* chi_part, const double* scale_part, int nbp, double* stats,
              unsigned int* ticket) {
  // robust chi2 of a state; the LAST block to finish adds the block partials (and the partials of
  // g2o's computeScale from ba_back_kernel) in a fixed order: stats[0] = chi2, stats[1] = scale
  __shared__ double red[BA_THREADS / 32];
  __shared__ int last_s;
  const int i = blockIdx.x * BA_THREADS + threadIdx.x;
  double r0 = 0.0;
  if (i < M) {
    const Proj p = ba_project(poses + 16 * obs_cam[i], points + 3 * (size_t)obs_pt[i],
                              obs_xy[2 * (size_t)i], obs_xy[2 * (size_t)i + 1], f, cx, cy);
    double w;
    ba_huber(p.ex * p.ex + p.ey * p.ey, delta, r0, w);
  }
  const double s = block_sum_ordered(r0, red);
  if (threadIdx.x == 0) {
    chi_part[blockIdx.x] = s;
    __threadfence();
    last_s = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last_s || threadIdx.x >= 32) return;
  __threadfence();
  const int lane = threadIdx.x;
  double c = 0.0, sc = 0.0;
  for (int b = lane; b < (int)gridDim.x; b += 32) c += __ldcg(chi_part + b);
  for (int b = lane; b < nbp; b += 32) sc += __ldcg(scale_part + b);
  for (int o = 16; o > 0; o >>= 1) {
    c += __shfl_down_sync(0xffffffffu, c, o);
    sc += __shfl_down_sync(0xffffffffu, sc, o);
  }
  if (lane == 0) { stats[0] = c; stats[1] = sc + stats[3]; *ticket = 0u; }
}

// per-observation records (gathered through the point / camera lists, so array-of-records):
//   OA[i][10] = d e / d point (2x3), Huber weight, e_x, e_y, pad      (80 B)
//   OB[i][12] = d e / d pose (2x6)                                    (96 B)
//   OW[i][18] = w * B^T A (6x3), zero for a fixed camera              (144 B)
__global__ void __launch_bounds__(BA_THREADS)
ba_linearize_kernel(const double* __restrict__ poses, const uint8_t* __restrict__ fixed,
                    const double* __restrict__ points, const int32_t* __restrict__ obs_cam,
                    const int32_t* __restrict__ obs_pt, const double* __restrict__ obs_xy, int M,
                    double f, double cx, double cy, double delta, double* __restrict__ OA,
                    double* __restrict__ OB, double* __restrict__ OW) {
  const int i = blockIdx.x * BA_THREADS + threadIdx.x;
  if (i >= M) return;
  const int c = obs_cam[i];
  const double* T = poses + 16 * c;
  const Proj p = ba_project(T, points + 3 * (size_t)obs_pt[i], obs_xy[2 * (size_t)i],
                            obs_xy[2 * (size_t)i + 1], f, cx, cy);
  double rho0, w;
  ba_huber(p.ex * p.ex + p.ey * p.ey, delta, rho0, w);
  const double x = p.x, y = p.y, z = p.z, z2 = z * z, iz = -1.0 / z;
  const double t02 = -x / z * f, t12 = -y / z * f;
  double A[2][3], B[2][6];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    A[0][k] = iz * (f * T[k] + t02 * T[8 + k]);
    A[1][k] = iz * (f * T[4 + k] + t12 * T[8 + k]);
  }
  B[0][0] = x * y / z2 * f;        B[0][1] = -(1.0 + x * x / z2) * f; B[0][2] = y / z * f;
  B[0][3] = -1.0 / z * f;          B[0][4] = 0.0;                     B[0][5] = x / z2 * f;
  B[1][0] = (1.0 + y * y / z2) * f; B[1][1] = -x * y / z2 * f;        B[1][2] = -x / z * f;
  B[1][3] = 0.0;                   B[1][4] = -1.0 / z * f;            B[1][5] = y / z2 * f;
  double2* oa = reinterpret_cast<double2*>(OA + 10 * (size_t)i);
  oa[0] = make_double2(A[0][0], A[0][1]); oa[1] = make_double2(A[0][2], A[1][0]);
  oa[2] = make_double2(A[1][1], A[1][2]); oa[3] = make_double2(w, p.ex);
  oa[4] = make_double2(p.ey, 0.0);
  double2* ob = reinterpret_cast<double2*>(OB + 12 * (size_t)i);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    ob[k] = make_double2(B[0][2 * k], B[0][2 * k + 1]);
    ob[3 + k] = make_double2(B[1][2 * k], B[1][2 * k + 1]);
  }
  const bool free_cam = !fixed[c];
  double wv[18];
#pragma unroll
  for (int r = 0; r < 6; ++r)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      wv[r * 3 + k] = free_cam ? w * (B[0][r] * A[0][k] + B[1][r] * A[1][k]) : 0.0;
  double2* ow = reinterpret_cast<double2*>(OW + 18 * (size_t)i);
#pragma unroll
  for (int k = 0; k < 9; ++k) ow[k] = make_double2(wv[2 * k], wv[2 * k + 1]);
}

// points are handled by groups of 4 lanes: the lanes stride over the point's observation list,
// partial sums are combined by two xor-shuffles ((l0+l1)+(l2+l3)): a fixed order
__device__ __forceinline__ double group4_sum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}

__global__ void __launch_bounds__(BA_THREADS)
ba_point_blocks_kernel(const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs, int P,
                       const double* __restrict__ OA, double* __restrict__ Hll,
                       double* __restrict__ bl) {
  const int t = blockIdx.x * BA_THREADS + threadIdx.x, p = t / BA_PT_LANES, sub = t % BA_PT_LANES;
  double h[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
  if (p < P) {
    for (int a = pt_start[p] + sub; a < pt_start[p + 1]; a += BA_PT_LANES) {
      double r[10];
      load_rec<5>(OA + 10 * (size_t)pt_obs[a], r);
      const double w = r[6], ex = r[7], ey = r[8];
      h[0] += w * (r[0] * r[0] + r[3] * r[3]);
      h[1] += w * (r[0] * r[1] + r[3] * r[4]);
      h[2] += w * (r[0] * r[2] + r[3] * r[5]);
      h[3] += w * (r[1] * r[1] + r[4] * r[4]);
      h[4] += w * (r[1] * r[2] + r[4] * r[5]);
      h[5] += w * (r[2] * r[2] + r[5] * r[5]);
#pragma unroll
      for (int k = 0; k < 3; ++k) b[k] -= w * (r[k] * ex + r[3 + k] * ey);
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) h[k] = group4_sum(h[k]);
#pragma unroll
  for (int k = 0; k < 3; ++k) b[k] = group4_sum(b[k]);
  if (p < P && sub == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) Hll[6 * (size_t)p + k] = h[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) bl[3 * (size_t)p + k] = b[k];
  }
}

__global__ void __launch_bounds__(BA_THREADS)
ba_camera_blocks_kernel(const int32_t* __restrict__ cam_start, const int32_t* __restrict__ cam_obs,
                        const int32_t* __restrict__ flist, int NS, const double* __restrict__ OA,
                        const double* __restrict__ OB, double* __restrict__ Cpart) {
  // block (f, s): chunk s of the observation list of the f-th free camera; 21 + 6 ordered sums
  __shared__ double red[(BA_THREADS / 32) * 27];
  const int c = flist[1 + blockIdx.x], sp = blockIdx.y;
  const int beg = cam_start[c], end = cam_start[c + 1], chunk = (end - beg + NS - 1) / NS;
  const int lo = beg + sp * chunk, hi = min(lo + chunk, end);
  double acc[27];
#pragma unroll
  for (int k = 0; k < 27; ++k) acc[k] = 0.0;
  for (int a = lo + threadIdx.x; a < hi; a += BA_THREADS) {
    const int i = cam_obs[a];
    double B[12], r[4];
    load_rec<6>(OB + 12 * (size_t)i, B);
    load_rec<2>(OA + 10 * (size_t)i + 6, r);
    const double w = r[0], ex = r[1], ey = r[2];
    int q = 0;
#pragma unroll
    for (int rr = 0; rr < 6; ++rr)
#pragma unroll
      for (int s = rr; s < 6; ++s) acc[q++] += w * (B[rr] * B[s] + B[6 + rr] * B[6 + s]);
#pragma unroll
    for (int rr = 0; rr < 6; ++rr) acc[21 + rr] -= w * (B[rr] * ex + B[6 + rr] * ey);
  }
  block_sum_ordered_n<27>(acc, red);
  if (threadIdx.x < 27) {
    double s = 0.0;
    for (int w = 0; w < BA_THREADS / 32; ++w) s += red[w * 27 + threadIdx.x];
    Cpart[((size_t)blockIdx.x * NS + sp) * 27 + threadIdx.x] = s;
  }
}

__global__ void ba_camera_fold_kernel(const int32_t* __restrict__ flist, int NS,
                                      const double* __restrict__ Cpart, double* __restrict__ Hpp,
                                      double* __restrict__ bp) {
  const int c = flist[1 + blockIdx.x], k = threadIdx.x;
  if (k >= 27) return;
  double s = 0.0;
  for (int sp = 0; sp < NS; ++sp) s += Cpart[((size_t)blockIdx.x * NS + sp) * 27 + k];
  if (k < 21) Hpp[21 * c + k] = s; else bp[6 * c + (k - 21)] = s;
}

__global__ void ba_maxdiag_kernel(const double* __restrict__ Hll, const int32_t* __restrict__ pt_start,
                                  int P, const double* __restrict__ Hpp,
                                  const int32_t* __restrict__ cam_start,
                                  const uint8_t* __restrict__ fixed, int C,
                                  unsigned long long* maxdiag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double m = 0.0;
  if (i < P) {
    if (pt_start[i + 1] > pt_start[i])
      m = fmax(fabs(Hll[6 * (size_t)i]), fmax(fabs(Hll[6 * (size_t)i + 3]), fabs(Hll[6 * (size_t)i + 5])));
  } else if (i < P + C) {
    const int c = i - P;
    if (!fixed[c] && cam_start[c + 1] > cam_start[c]) {
      const int d[6] = {0, 6, 11, 15, 18, 20};
      for (int k = 0; k < 6; ++k) m = fmax(m, fabs(Hpp[21 * c + d[k]]));
    }
  } else {
    return;
  }
  if (m > 0.0) atomicMax(maxdiag, (unsigned long long)__double_as_longlong(m));  // order-free
}

// ------------------------------------------------------------------ one LM trial
__device__ __forceinline__ void sym3_inverse(const double* h, double lam, double* inv) {
  // (H + lam I)^-1 of a symmetric 3x3 in packed form [xx xy xz yy yz zz], by cofactors
  const double a = h[0] + lam, b = h[1], c = h[2], d = h[3] + lam, e = h[4], g = h[5] + lam;
  const double c00 = d * g - e * e, c01 = c * e - b * g, c02 = b * e - c * d;
  const double det = a * c00 + b * c01 + c * c02, id = 1.0 / det;
  inv[0] = c00 * id; inv[1] = c01 * id; inv[2] = c02 * id;
  inv[3] = (a * g - c * c) * id; inv[4] = (b * c - a * e) * id; inv[5] = (a * d - b * b) * id;
}

__global__ void __launch_bounds__(BA_THREADS)
ba_schur_kernel(const int32_t* __restrict__ cam_start, const int32_t* __restrict__ cam_obs,
                const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs,
                const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
                const int32_t* __restrict__ flist, int NS, const double* __restrict__ OW,
                const double* __restrict__ Hll, const double* __restrict__ bl, double lam,
                double* __restrict__ Spart) {
  // block (fa, fb, s), fa <= fb over the FREE cameras: chunk s of camera a's observation list.
  //   partial of  sum_{i in obs(a)} Y_i * sum_{j in obs(point_i), cam_j == b} W_j^T   (36 values)
  //   and, on the diagonal, of  sum_i Y_i bl[point_i]  (6 values),  Y_i = W_i (Hll + lam I)^-1
  __shared__ double red[(BA_THREADS / 32) * 42];
  const int fa = blockIdx.x, fb = blockIdx.y, sp = blockIdx.z, nf = gridDim.x;
  if (fa > fb) return;
  const int a = flist[1 + fa], b = flist[1 + fb];
  const int beg = cam_start[a], end = cam_start[a + 1], chunk = (end - beg + NS - 1) / NS;
  const int lo = beg + sp * chunk, hi = min(lo + chunk, end);
  double acc[42];
#pragma unroll
  for (int k = 0; k < 42; ++k) acc[k] = 0.0;
  for (int idx = lo + threadIdx.x; idx < hi; idx += BA_THREADS) {
    const int i = cam_obs[idx], p = obs_pt[i];
    double v[6], wi[18], y[18];
    sym3_inverse(Hll + 6 * (size_t)p, lam, v);
    load_rec<9>(OW + 18 * (size_t)i, wi);
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double w0 = wi[r * 3], w1 = wi[r * 3 + 1], w2 = wi[r * 3 + 2];
      y[r * 3] = w0 * v[0] + w1 * v[1] + w2 * v[2];
      y[r * 3 + 1] = w0 * v[1] + w1 * v[3] + w2 * v[4];
      y[r * 3 + 2] = w0 * v[2] + w1 * v[4] + w2 * v[5];
    }
    if (a == b) {
      const double b0 = bl[3 * (size_t)p], b1 = bl[3 * (size_t)p + 1], b2 = bl[3 * (size_t)p + 2];
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[36 + r] += y[r * 3] * b0 + y[r * 3 + 1] * b1 + y[r * 3 + 2] * b2;
    }
    for (int l = pt_start[p]; l < pt_start[p + 1]; ++l) {
      const int j = pt_obs[l];
      if (obs_cam[j] != b) continue;
      double wj[18];
      load_rec<9>(OW + 18 * (size_t)j, wj);
#pragma unroll
      for (int s = 0; s < 6; ++s)
#pragma unroll
        for (int r = 0; r < 6; ++r)
          acc[r * 6 + s] += y[r * 3] * wj[s * 3] + y[r * 3 + 1] * wj[s * 3 + 1] + y[r * 3 + 2] * wj[s * 3 + 2];
    }
  }
  const size_t pair = (size_t)fa * nf - (size_t)fa * (fa - 1) / 2 + (fb - fa);
  block_sum_ordered_n<42>(acc, red);
  if (threadIdx.x < (a == b ? 42 : 36)) {
    double s = 0.0;
    for (int w = 0; w < BA_THREADS / 32; ++w) s += red[w * 42 + threadIdx.x];
    Spart[(pair * NS + sp) * 42 + threadIdx.x] = s;
  }
}

__device__ void ba_se3_exp_apply(const double* dx, const double* T, double* out) {
  // out = exp(dx) * T,  dx = (omega, upsilon); rows 0..2 of the 4x4, row 3 = (0,0,0,1)
  const double wx = dx[0], wy = dx[1], wz = dx[2];
  const double theta = sqrt(wx * wx + wy * wy + wz * wz);
  const double Om[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double Om2[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      Om2[3 * r + c] = Om[3 * r] * Om[c] + Om[3 * r + 1] * Om[3 + c] + Om[3 * r + 2] * Om[6 + c];
  double ka, kb, va, vb;
  if (theta < 0.00001) {
    ka = 1.0; kb = 0.5; va = 0.5; vb = 1.0 / 6.0;
  } else {
    const double s = sin(theta), c = cos(theta), t2 = theta * theta;
    ka = s / theta; kb = (1.0 - c) / t2; va = (1.0 - c) / t2; vb = (theta - s) / (t2 * theta);
  }
  double R[9], V[9];
  for (int k = 0; k < 9; ++k) {
    const double I = (k % 4 == 0) ? 1.0 : 0.0;
    R[k] = I + ka * Om[k] + kb * Om2[k];
    V[k] = I + va * Om[k] + vb * Om2[k];
  }
  double dt[3];
  for (int r = 0; r < 3; ++r) dt[r] = V[3 * r] * dx[3] + V[3 * r + 1] * dx[4] + V[3 * r + 2] * dx[5];
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c)
      out[4 * r + c] = R[3 * r] * T[c] + R[3 * r + 1] * T[4 + c] + R[3 * r + 2] * T[8 + c];
    out[4 * r + 3] = R[3 * r] * T[3] + R[3 * r + 1] * T[7] + R[3 * r + 2] * T[11] + dt[r];
  }
  out[12] = 0.0; out[13] = 0.0; out[14] = 0.0; out[15] = 1.0;
}

__global__ void __launch_bounds__(BA_THREADS)
ba_solve_kernel(const double* __restrict__ Spart, int NS, const int32_t* __restrict__ flist, int C,
                const double* __restrict__ Hpp, const double* __restrict__ bp, double lam,
                const double* __restrict__ poses, double* __restrict__ trial_poses,
                double* __restrict__ xp, double* __restrict__ stats) {
  // reduced camera system over the nf free cameras (n = 6 nf <= 144) in shared memory:
  // S = Hpp + lam I - sum of the Schur partials (folded in split order), dense Cholesky (two
  // barriers per column), triangular solves by warp 0, pose update exp(dx) * T, camera part of
  // g2o's computeScale()
  extern __shared__ double sm[];
  const int nf = flist[0], n = 6 * nf, ld = n + 1;
  double* A = sm;                        // [n][ld]
  double* x = sm + (size_t)n * ld;       // [n]
  double* diag = x + n;                  // [n]
  const int tid = threadIdx.x;
  for (int e = tid; e < n * n; e += BA_THREADS) {
    const int r = e / n, c = e % n;
    int fa = r / 6, fb = c / 6, rr = r % 6, cc = c % 6;
    if (fa > fb) { int t = fa; fa = fb; fb = t; t = rr; rr = cc; cc = t; }
    const size_t pair = (size_t)fa * nf - (size_t)fa * (fa - 1) / 2 + (fb - fa);
    double sub = 0.0;
    for (int sp = 0; sp < NS; ++sp) sub += Spart[(pair * NS + sp) * 42 + rr * 6 + cc];
    double h = 0.0;
    if (fa == fb) {
      const int lo = rr < cc ? rr : cc, hi = rr < cc ? cc : rr;
      h = Hpp[21 * flist[1 + fa] + (lo * 6 - lo * (lo - 1) / 2 + (hi - lo))] + (rr == cc ? lam : 0.0);
    }
    A[r * ld + c] = h - sub;
  }
  for (int r = tid; r < n; r += BA_THREADS) {
    const int fa = r / 6;
    const size_t pair = (size_t)fa * nf - (size_t)fa * (fa - 1) / 2;
    double sub = 0.0;
    for (int sp = 0; sp < NS; ++sp) sub += Spart[(pair * NS + sp) * 42 + 36 + r % 6];
    x[r] = bp[6 * flist[1 + fa] + r % 6] - sub;
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;
  int ok = 1;
  for (int k = 0; k < n; ++k) {
    const double d = A[k * ld + k];          // final since the barrier that ended step k-1
    if (!(d > 0.0)) { ok = 0; break; }       // same value in every thread
    const double sd = sqrt(d);
    if (tid == 0) diag[k] = sd;
    for (int i = k + 1 + tid; i < n; i += BA_THREADS) A[i * ld + k] /= sd;
    __syncthreads();
    for (int i = k + 1 + ty; i < n; i += 16)
      for (int j = k + 1 + tx; j <= i; j += 16) A[i * ld + j] -= A[i * ld + k] * A[j * ld + k];
    __syncthreads();
  }
  if (tid < 32) {
    if (ok) {
      for (int k = 0; k < n; ++k) {           // L y = b
        if (tid == 0) x[k] /= diag[k];
        __syncwarp();
        const double yk = x[k];
        for (int i = k + 1 + tid; i < n; i += 32) x[i] -= A[i * ld + k] * yk;
        __syncwarp();
      }
      for (int k = n - 1; k >= 0; --k) {      // L^T x = y
        if (tid == 0) x[k] /= diag[k];
        __syncwarp();
        const double xk = x[k];
        for (int i = tid; i < k; i += 32) x[i] -= A[k * ld + i] * xk;
        __syncwarp();
      }
    } else {
      for (int i = tid; i < n; i += 32) x[i] = 0.0;
    }
  }
  __syncthreads();
  for (int r = tid; r < 6 * C; r += BA_THREADS) xp[r] = 0.0;
  for (int r = tid; r < 16 * C; r += BA_THREADS) trial_poses[r] = poses[r];
  __syncthreads();
  for (int r = tid; r < n; r += BA_THREADS) xp[6 * flist[1 + r / 6] + r % 6] = x[r];
  if (tid < nf && ok) {
    const int c = flist[1 + tid];
    ba_se3_exp_apply(x + 6 * tid, poses + 16 * c, trial_poses + 16 * c);
  }
  if (tid == 0) {
    double sc = 0.0;
    for (int f = 0; f < nf; ++f)
      for (int r = 0; r < 6; ++r)
        sc += x[6 * f + r] * (lam * x[6 * f + r] + bp[6 * flist[1 + f] + r]);
    stats[3] = sc;
    stats[2] = (double)ok;
  }
}

__global__ void __launch_bounds__(BA_THREADS)
ba_back_kernel(const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs,
               const int32_t* __restrict__ obs_cam, const uint8_t* __restrict__ fixed, int P,
               const double* __restrict__ OW, const double* __restrict__ Hll,
               const double* __restrict__ bl, double lam, const double* __restrict__ xp,
               const double* __restrict__ points, double* __restrict__ trial_points,
               double* __restrict__ scale_part) {
  // 4 lanes per point: dx_l = (Hll + lam I)^-1 (bl - sum_j W_j^T dx_cam(j)); X <- X + dx_l;
  // point part of g2o's computeScale()
  __shared__ double red[BA_THREADS / 32];
  const int t = blockIdx.x * BA_THREADS + threadIdx.x, p = t / BA_PT_LANES, sub = t % BA_PT_LANES;
  double cp[3] = {0.0, 0.0, 0.0};
  bool active = false;
  if (p < P) {
    active = pt_start[p + 1] > pt_start[p];
    for (int l = pt_start[p] + sub; l < pt_start[p + 1]; l += BA_PT_LANES) {
      const int j = pt_obs[l], c = obs_cam[j];
      if (fixed[c]) continue;              // W_j = 0 and dx = 0 there
      double wj[18];
      load_rec<9>(OW + 18 * (size_t)j, wj);
      const double* d = xp + 6 * c;
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const double dr = d[r];
        cp[0] -= wj[r * 3] * dr;
        cp[1] -= wj[r * 3 + 1] * dr;
        cp[2] -= wj[r * 3 + 2] * dr;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) cp[k] = group4_sum(cp[k]);
  double sc = 0.0;
  if (p < P && sub == 0) {
    double X[3] = {points[3 * (size_t)p], points[3 * (size_t)p + 1], points[3 * (size_t)p + 2]};
    if (active) {
      const double b0 = bl[3 * (size_t)p], b1 = bl[3 * (size_t)p + 1], b2 = bl[3 * (size_t)p + 2];
      cp[0] += b0; cp[1] += b1; cp[2] += b2;
      double v[6];
      sym3_inverse(Hll + 6 * (size_t)p, lam, v);
      const double d0 = v[0] * cp[0] + v[1] * cp[1] + v[2] * cp[2];
      const double d1 = v[1] * cp[0] + v[3] * cp[1] + v[4] * cp[2];
      const double d2 = v[2] * cp[0] + v[4] * cp[1] + v[5] * cp[2];
      X[0] += d0; X[1] += d1; X[2] += d2;
      sc = d0 * (lam * d0 + b0) + d1 * (lam * d1 + b1) + d2 * (lam * d2 + b2);
    }
    trial_points[3 * (size_t)p] = X[0];
    trial_points[3 * (size_t)p + 1] = X[1];
    trial_points[3 * (size_t)p + 2] = X[2];
  }
  const double s = block_sum_ordered(sc, red);
  if (threadIdx.x == 0) scale_part[blockIdx.x] = s;
}

}  // namespace vo

using namespace vo;

extern "C" size_t vo_bundle_adjust_workspace_bytes(int n_cams, int n_points, int n_obs) {
  if (n_cams < 0 || n_points < 0 || n_obs < 0) return 0;
  return ba_carve(nullptr, n_cams, n_points, n_obs).bytes + 256;
}

extern "C" int vo_bundle_adjust(double* poses, const uint8_t* fixed, int n_cams, double* points,
                                int n_points, const int32_t* obs_cam, const int32_t* obs_pt,
                                const double* obs_xy, int n_obs, double focal, double cx, double cy,
                                double huber_delta, int iterations, double* log_chi2,
                                double* log_lambda, int32_t* log_trials, int* n_iterations,
                                double* chi2_initial, void* ws, size_t ws_bytes, void* stream) {
  VO_REQUIRE(n_cams >= 0 && n_points >= 0 && n_obs >= 0 && iterations >= 0, VO_ERR_INVALID_ARG,
             "vo_bundle_adjust: negative size");
  VO_REQUIRE(n_cams <= BA_MAX_CAMS, VO_ERR_INVALID_ARG,
             "vo_bundle_adjust: more than %d cameras (the reference keeps <= 18)", BA_MAX_CAMS);
  if (n_iterations) *n_iterations = 0;
  if (chi2_initial) *chi2_initial = 0.0;
  if (n_obs == 0 || n_cams == 0 || n_points == 0 || iterations == 0) return VO_OK;
  VO_REQUIRE(poses && fixed && points && obs_cam && obs_pt && obs_xy && ws, VO_ERR_INVALID_ARG,
             "vo_bundle_adjust: null pointer");
  VO_REQUIRE(ws_bytes >= vo_bundle_adjust_workspace_bytes(n_cams, n_points, n_obs), VO_ERR_WORKSPACE,
             "vo_bundle_adjust: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int C = n_cams, P = n_points, M = n_obs;
  char* base = (char*)(((uintptr_t)ws + 255) / 256 * 256);
  BAWs w = ba_carve(base, C, P, M);
  const int nbo = (M + BA_THREADS - 1) / BA_THREADS, nbp = (P + BA_THREADS - 1) / BA_THREADS;
  const int nbp4 = (int)(((size_t)P * BA_PT_LANES + BA_THREADS - 1) / BA_THREADS);
  static bool attr_done[64] = {false};     // cudaFuncSetAttribute is per device
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  cur_dev &= 63;
  if (!attr_done[cur_dev]) {
    cudaFuncSetAttribute(ba_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)(sizeof(double) * ((size_t)6 * BA_MAX_CAMS * (6 * BA_MAX_CAMS + 1) + 12 * BA_MAX_CAMS)));
    attr_done[cur_dev] = true;
  }

  // adjacency lists
  cudaMemsetAsync(w.pt_start, 0, sizeof(int32_t) * (P + 1), st);
  cudaMemsetAsync(w.cam_count, 0, sizeof(int32_t) * (C + 1), st);
  cudaMemsetAsync(w.bad, 0, sizeof(int32_t), st);
  cudaMemsetAsync(w.maxdiag, 0, sizeof(unsigned long long), st);
  cudaMemsetAsync(w.stats, 0, sizeof(double) * 8, st);
  cudaMemsetAsync(w.ticket, 0, sizeof(unsigned int), st);
  ba_count_kernel<<<nbo, BA_THREADS, 0, st>>>(obs_cam, obs_pt, M, C, P, w.pt_start, w.cam_count, w.bad);
  int32_t h_bad = 0, h_count[BA_MAX_CAMS];
  uint8_t h_fixed[BA_MAX_CAMS];
  cudaMemcpyAsync(&h_bad, w.bad, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  cudaMemcpyAsync(h_count, w.cam_count, sizeof(int32_t) * C, cudaMemcpyDeviceToHost, st);
  cudaMemcpyAsync(h_fixed, fixed, C, cudaMemcpyDeviceToHost, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(count)");
  VO_REQUIRE(!h_bad, VO_ERR_INVALID_ARG, "vo_bundle_adjust: observation index out of range");
  // free, observed cameras: the unknowns of the reduced system; how far their lists are split
  int nf = 0;
  for (int c = 0; c < C; ++c) nf += (!h_fixed[c] && h_count[c] > 0);
  const int npairs = nf * (nf + 1) / 2;
  auto split_for = [](int units) {
    int ns = units > 0 ? (BA_TARGET_BLOCKS + units - 1) / units : 1;
    return ns < 1 ? 1 : (ns > BA_MAX_SPLIT ? BA_MAX_SPLIT : ns);
  };
  const int NS = split_for(npairs), NSc = split_for(nf);
  const size_t solve_smem = sizeof(double) * ((size_t)6 * nf * (6 * nf + 1) + 12 * nf + 2);
  ba_scan_kernel<<<1, 1024, 0, st>>>(w.pt_start, P, w.pt_cursor, w.cam_count, w.cam_start, C, fixed,
                                     w.flist);
  ba_fill_pts_kernel<<<nbo, BA_THREADS, 0, st>>>(obs_pt, M, w.pt_cursor, w.pt_obs);
  ba_sort_pts_kernel<<<nbp, BA_THREADS, 0, st>>>(w.pt_start, P, w.pt_obs);
  ba_fill_cams_kernel<<<C, 1024, 0, st>>>(obs_cam, M, w.cam_start, w.cam_obs);

  double* cur_pts = points;  double* tri_pts = w.trial_points;
  double* cur_pos = poses;   double* tri_pos = w.trial_poses;
  double h_stats[4];
  auto chi_of = [&](const double* ps, const double* pt, int n_scale_parts) {
    ba_chi_kernel<<<nbo, BA_THREADS, 0, st>>>(ps, pt, obs_cam, obs_pt, obs_xy, M, focal, cx, cy,
                                              huber_delta, w.chi_part, w.scale_part, n_scale_parts,
                                              w.stats, w.ticket);
  };
  chi_of(cur_pos, cur_pts, 0);
  cudaMemcpyAsync(h_stats, w.stats, sizeof(double) * 4, cudaMemcpyDeviceToHost, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(setup)");
  double current = h_stats[0];
  if (chi2_initial) *chi2_initial = current;

  double lambda = 0.0, ni = 2.0;
  int it = 0;
  for (; it < iterations; ++it) {
    ba_linearize_kernel<<<nbo, BA_THREADS, 0, st>>>(cur_pos, fixed, cur_pts, obs_cam, obs_pt, obs_xy,
                                                    M, focal, cx, cy, huber_delta, w.OA, w.OB, w.OW);
    ba_point_blocks_kernel<<<nbp4, BA_THREADS, 0, st>>>(w.pt_start, w.pt_obs, P, w.OA, w.Hll, w.bl);
    if (nf > 0) {
      ba_camera_blocks_kernel<<<dim3(nf, NSc), BA_THREADS, 0, st>>>(w.cam_start, w.cam_obs, w.flist, NSc,
                                                                    w.OA, w.OB, w.Cpart);
      ba_camera_fold_kernel<<<nf, 32, 0, st>>>(w.flist, NSc, w.Cpart, w.Hpp, w.bp);
    }
    if (it == 0) {
      ba_maxdiag_kernel<<<(P + C + 255) / 256, 256, 0, st>>>(w.Hll, w.pt_start, P, w.Hpp, w.cam_start,
                                                            fixed, C, w.maxdiag);
      unsigned long long bits = 0;
      cudaMemcpyAsync(&bits, w.maxdiag, sizeof(bits), cudaMemcpyDeviceToHost, st);
      if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(lambda)");
      double md;
      memcpy(&md, &bits, sizeof(md));
      lambda = 1e-5 * md;
      ni = 2.0;
    }
    double rho = 0.0;
    int qmax = 0;
    do {
      if (nf > 0)
        ba_schur_kernel<<<dim3(nf, nf, NS), BA_THREADS, 0, st>>>(
            w.cam_start, w.cam_obs, w.pt_start, w.pt_obs, obs_cam, obs_pt, w.flist, NS, w.OW, w.Hll,
            w.bl, lambda, w.Spart);
      ba_solve_kernel<<<1, BA_THREADS, solve_smem, st>>>(w.Spart, NS, w.flist, C, w.Hpp, w.bp, lambda,
                                                        cur_pos, tri_pos, w.xp, w.stats);
      ba_back_kernel<<<nbp4, BA_THREADS, 0, st>>>(w.pt_start, w.pt_obs, obs_cam, fixed, P, w.OW, w.Hll,
                                                 w.bl, lambda, w.xp, cur_pts, tri_pts, w.scale_part);
      chi_of(tri_pos, tri_pts, nbp4);
      cudaMemcpyAsync(h_stats, w.stats, sizeof(double) * 4, cudaMemcpyDeviceToHost, st);
      if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(trial)");
      const bool ok = h_stats[2] != 0.0;
      const double temp = ok ? h_stats[0] : DBL_MAX;
      const double scale = (ok ? h_stats[1] : 0.0) + 1e-3;
      rho = (current - temp) / scale;
      if (rho > 0.0 && isfinite(temp)) {
        double alpha = 1.0 - pow(2.0 * rho - 1.0, 3.0);
        alpha = alpha < 2.0 / 3.0 ? alpha : 2.0 / 3.0;
        lambda *= alpha > 1.0 / 3.0 ? alpha : 1.0 / 3.0;
        ni = 2.0;
        current = temp;
        double* t;
        t = cur_pts; cur_pts = tri_pts; tri_pts = t;
        t = cur_pos; cur_pos = tri_pos; tri_pos = t;
      } else {
        lambda *= ni;
        ni *= 2.0;
        if (!isfinite(lambda)) break;
      }
      ++qmax;
    } while (rho < 0.0 && qmax < BA_LM_MAX_TRIALS);
    if (log_chi2) log_chi2[it] = current;
    if (log_lambda) log_lambda[it] = lambda;
    if (log_trials) log_trials[it] = qmax;
    if (qmax == BA_LM_MAX_TRIALS || rho == 0.0 || !isfinite(lambda)) { ++it; break; }
  }
  if (n_iterations) *n_iterations = it;
  if (cur_pts != points) cudaMemcpyAsync(points, cur_pts, sizeof(double) * 3 * (size_t)P, cudaMemcpyDeviceToDevice, st);
  if (cur_pos != poses) cudaMemcpyAsync(poses, cur_pos, sizeof(double) * 16 * C, cudaMemcpyDeviceToDevice, st);
  return check_launch("vo_bundle_adjust");
}
