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

struct BAWs {
  int32_t *pt_start, *pt_cursor, *pt_obs, *cam_start, *cam_obs, *cam_count, *bad;
  double *JA, *JB, *we, *W, *Y;           // per observation, component-major [k * M + i]
  double *Hll, *bl, *Hpp, *bp;            // [P,6] [P,3] [C,21] [C,6]
  double *S, *bs, *xp;                    // [C,C,36] [6C] [6C]
  double *trial_points, *trial_poses;     // [P,3] [C,16]
  double *chi_part, *scale_part, *stats;  // block partials; stats[0..7]
  unsigned long long* maxdiag;
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
  w.JA = (double*)take(sizeof(double) * 6 * (size_t)M);
  w.JB = (double*)take(sizeof(double) * 12 * (size_t)M);
  w.we = (double*)take(sizeof(double) * 3 * (size_t)M);
  w.W = (double*)take(sizeof(double) * 18 * (size_t)M);
  w.Y = (double*)take(sizeof(double) * 18 * (size_t)M);
  w.Hll = (double*)take(sizeof(double) * 6 * (size_t)P);
  w.bl = (double*)take(sizeof(double) * 3 * (size_t)P);
  w.Hpp = (double*)take(sizeof(double) * 21 * C);
  w.bp = (double*)take(sizeof(double) * 6 * C);
  w.S = (double*)take(sizeof(double) * 36 * C * C);
  w.bs = (double*)take(sizeof(double) * 6 * C);
  w.xp = (double*)take(sizeof(double) * 6 * C);
  w.trial_points = (double*)take(sizeof(double) * 3 * (size_t)P);
  w.trial_poses = (double*)take(sizeof(double) * 16 * C);
  w.chi_part = (double*)take(sizeof(double) * (nbo + 1));
  w.scale_part = (double*)take(sizeof(double) * (nbp + 1));
  w.stats = (double*)take(sizeof(double) * 8);
  w.maxdiag = (unsigned long long*)take(sizeof(unsigned long long));
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
               int32_t* __restrict__ cam_start, int C) {
  __shared__ int wsum[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {
    base_s = 0;
    int s = 0;
    for (int c = 0; c < C; ++c) { cam_start[c] = s; s += cam_count[c]; }
    cam_start[C] = s;
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

__global__ void __launch_bounds__(BA_THREADS)
ba_chi_kernel(const double* __restrict__ poses, const double* __restrict__ points,
              const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
              const double* __restrict__ obs_xy, int M, double f, double cx, double cy,
              double delta, double* __restrict__ chi_part) {
  __shared__ double red[BA_THREADS / 32];
  const int i = blockIdx.x * BA_THREADS + threadIdx.x;
  double r0 = 0.0;
  if (i < M) {
    const Proj p = ba_project(poses + 16 * obs_cam[i], points + 3 * (size_t)obs_pt[i],
                              obs_xy[2 * (size_t)i], obs_xy[2 * (size_t)i + 1], f, cx, cy);
    double w;
    ba_huber(p.ex * p.ex + p.ey * p.ey, delta, r0, w);
  }
  const double s = block_sum_ordered(r0, red);
  if (threadIdx.x == 0) chi_part[blockIdx.x] = s;
}

__global__ void __launch_bounds__(BA_THREADS)
ba_linearize_kernel(const double* __restrict__ poses, const uint8_t* __restrict__ fixed,
                    const double* __restrict__ points, const int32_t* __restrict__ obs_cam,
                    const int32_t* __restrict__ obs_pt, const double* __restrict__ obs_xy, int M,
                    double f, double cx, double cy, double delta, double* __restrict__ JA,
                    double* __restrict__ JB, double* __restrict__ we, double* __restrict__ W) {
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
#pragma unroll
  for (int k = 0; k < 3; ++k) { JA[(size_t)k * M + i] = A[0][k]; JA[(size_t)(3 + k) * M + i] = A[1][k]; }
#pragma unroll
  for (int k = 0; k < 6; ++k) { JB[(size_t)k * M + i] = B[0][k]; JB[(size_t)(6 + k) * M + i] = B[1][k]; }
  we[i] = w; we[(size_t)M + i] = p.ex; we[2 * (size_t)M + i] = p.ey;
  const bool free_cam = !fixed[c];
#pragma unroll
  for (int r = 0; r < 6; ++r)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      W[(size_t)(r * 3 + k) * M + i] = free_cam ? w * (B[0][r] * A[0][k] + B[1][r] * A[1][k]) : 0.0;
}

__global__ void __launch_bounds__(BA_THREADS)
ba_point_blocks_kernel(const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs, int P,
                       int M, const double* __restrict__ JA, const double* __restrict__ we,
                       double* __restrict__ Hll, double* __restrict__ bl) {
  const int p = blockIdx.x * BA_THREADS + threadIdx.x;
  if (p >= P) return;
  double h[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
  for (int a = pt_start[p]; a < pt_start[p + 1]; ++a) {
    const int i = pt_obs[a];
    double A0[3], A1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { A0[k] = JA[(size_t)k * M + i]; A1[k] = JA[(size_t)(3 + k) * M + i]; }
    const double w = we[i], ex = we[(size_t)M + i], ey = we[2 * (size_t)M + i];
    h[0] += w * (A0[0] * A0[0] + A1[0] * A1[0]);
    h[1] += w * (A0[0] * A0[1] + A1[0] * A1[1]);
    h[2] += w * (A0[0] * A0[2] + A1[0] * A1[2]);
    h[3] += w * (A0[1] * A0[1] + A1[1] * A1[1]);
    h[4] += w * (A0[1] * A0[2] + A1[1] * A1[2]);
    h[5] += w * (A0[2] * A0[2] + A1[2] * A1[2]);
#pragma unroll
    for (int k = 0; k < 3; ++k) b[k] -= w * (A0[k] * ex + A1[k] * ey);
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) Hll[6 * (size_t)p + k] = h[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) bl[3 * (size_t)p + k] = b[k];
}

__global__ void __launch_bounds__(BA_THREADS)
ba_camera_blocks_kernel(const int32_t* __restrict__ cam_start, const int32_t* __restrict__ cam_obs,
                        const uint8_t* __restrict__ fixed, int M, const double* __restrict__ JB,
                        const double* __restrict__ we, double* __restrict__ Hpp,
                        double* __restrict__ bp) {
  // block = camera; 21 + 6 sums, each thread strides over the camera's list, ordered block sums
  __shared__ double red[BA_THREADS / 32];
  const int c = blockIdx.x;
  double acc[27];
#pragma unroll
  for (int k = 0; k < 27; ++k) acc[k] = 0.0;
  if (!fixed[c]) {
    for (int a = cam_start[c] + threadIdx.x; a < cam_start[c + 1]; a += BA_THREADS) {
      const int i = cam_obs[a];
      double B0[6], B1[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) { B0[k] = JB[(size_t)k * M + i]; B1[k] = JB[(size_t)(6 + k) * M + i]; }
      const double w = we[i], ex = we[(size_t)M + i], ey = we[2 * (size_t)M + i];
      int q = 0;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int s = r; s < 6; ++s) acc[q++] += w * (B0[r] * B0[s] + B1[r] * B1[s]);
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[21 + r] -= w * (B0[r] * ex + B1[r] * ey);
    }
  }
#pragma unroll
  for (int k = 0; k < 27; ++k) {
    const double s = block_sum_ordered(acc[k], red);
    if (threadIdx.x == 0) {
      if (k < 21) Hpp[21 * c + k] = s; else bp[6 * c + (k - 21)] = s;
    }
  }
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
ba_y_kernel(const int32_t* __restrict__ obs_pt, int M, const double* __restrict__ Hll, double lam,
            const double* __restrict__ W, double* __restrict__ Y) {
  const int i = blockIdx.x * BA_THREADS + threadIdx.x;
  if (i >= M) return;
  double v[6];
  sym3_inverse(Hll + 6 * (size_t)obs_pt[i], lam, v);
#pragma unroll
  for (int r = 0; r < 6; ++r) {
    const double w0 = W[(size_t)(r * 3) * M + i], w1 = W[(size_t)(r * 3 + 1) * M + i],
                 w2 = W[(size_t)(r * 3 + 2) * M + i];
    Y[(size_t)(r * 3) * M + i] = w0 * v[0] + w1 * v[1] + w2 * v[2];
    Y[(size_t)(r * 3 + 1) * M + i] = w0 * v[1] + w1 * v[3] + w2 * v[4];
    Y[(size_t)(r * 3 + 2) * M + i] = w0 * v[2] + w1 * v[4] + w2 * v[5];
  }
}

__global__ void __launch_bounds__(BA_THREADS)
ba_schur_kernel(const int32_t* __restrict__ cam_start, const int32_t* __restrict__ cam_obs,
                const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs,
                const int32_t* __restrict__ obs_cam, const int32_t* __restrict__ obs_pt,
                const uint8_t* __restrict__ fixed, int C, int M, const double* __restrict__ W,
                const double* __restrict__ Y, const double* __restrict__ bl,
                const double* __restrict__ Hpp, const double* __restrict__ bp, double lam,
                double* __restrict__ S, double* __restrict__ bs) {
  // block (a, b), a <= b: S_ab = [a==b](Hpp_a + lam I) - sum_{i in obs(a)} Y_i * sum_{j in
  // obs(point_i), cam_j == b} W_j^T;  bs_a = bp_a - sum_i Y_i bl[point_i]
  __shared__ double red[BA_THREADS / 32];
  const int a = blockIdx.x, b = blockIdx.y;
  if (a > b || fixed[a] || fixed[b]) return;
  if (cam_start[a + 1] == cam_start[a] || cam_start[b + 1] == cam_start[b]) return;
  double acc[42];
#pragma unroll
  for (int k = 0; k < 42; ++k) acc[k] = 0.0;
  for (int idx = cam_start[a] + threadIdx.x; idx < cam_start[a + 1]; idx += BA_THREADS) {
    const int i = cam_obs[idx], p = obs_pt[i];
    double y[18];
#pragma unroll
    for (int k = 0; k < 18; ++k) y[k] = Y[(size_t)k * M + i];
    for (int l = pt_start[p]; l < pt_start[p + 1]; ++l) {
      const int j = pt_obs[l];
      if (obs_cam[j] != b) continue;
#pragma unroll
      for (int s = 0; s < 6; ++s) {
        const double w0 = W[(size_t)(s * 3) * M + j], w1 = W[(size_t)(s * 3 + 1) * M + j],
                     w2 = W[(size_t)(s * 3 + 2) * M + j];
#pragma unroll
        for (int r = 0; r < 6; ++r) acc[r * 6 + s] += y[r * 3] * w0 + y[r * 3 + 1] * w1 + y[r * 3 + 2] * w2;
      }
    }
    if (a == b) {
      const double b0 = bl[3 * (size_t)p], b1 = bl[3 * (size_t)p + 1], b2 = bl[3 * (size_t)p + 2];
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[36 + r] += y[r * 3] * b0 + y[r * 3 + 1] * b1 + y[r * 3 + 2] * b2;
    }
  }
#pragma unroll
  for (int k = 0; k < 42; ++k) {
    if (k >= 36 && a != b) break;
    const double s = block_sum_ordered(acc[k], red);
    if (threadIdx.x == 0) {
      if (k < 36) {
        double h = 0.0;
        if (a == b) {
          const int r = k / 6, c = k % 6, lo = r < c ? r : c, hi = r < c ? c : r;
          h = Hpp[21 * a + (lo * 6 - lo * (lo - 1) / 2 + (hi - lo))] + (r == c ? lam : 0.0);
        }
        S[((size_t)a * C + b) * 36 + k] = h - s;
      } else {
        bs[6 * a + (k - 36)] = bp[6 * a + (k - 36)] - s;
      }
    }
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
ba_solve_kernel(const double* __restrict__ S, const double* __restrict__ bs,
                const int32_t* __restrict__ cam_start, const uint8_t* __restrict__ fixed, int C,
                const double* __restrict__ bp, double lam, const double* __restrict__ poses,
                double* __restrict__ trial_poses, double* __restrict__ xp,
                double* __restrict__ stats) {
  // dense Cholesky of the reduced camera system in shared memory (n = 6C <= 144), triangular
  // solves by warp 0, pose update exp(dx) * T, camera part of g2o's computeScale()
  extern __shared__ double sm[];
  const int n = 6 * C, ld = n + 1;
  double* A = sm;               // [n][ld]
  double* x = sm + (size_t)n * ld;   // [n]
  __shared__ int ok_s;
  const int tid = threadIdx.x;
  if (tid == 0) ok_s = 1;
  for (int e = tid; e < n * n; e += BA_THREADS) {
    const int r = e / n, c = e % n, a = r / 6, b = c / 6;
    const bool act_a = !fixed[a] && cam_start[a + 1] > cam_start[a];
    const bool act_b = !fixed[b] && cam_start[b + 1] > cam_start[b];
    double v;
    if (!act_a || !act_b) v = (r == c) ? 1.0 : 0.0;
    else if (a <= b) v = S[((size_t)a * C + b) * 36 + (r % 6) * 6 + (c % 6)];
    else v = S[((size_t)b * C + a) * 36 + (c % 6) * 6 + (r % 6)];
    A[r * ld + c] = v;
  }
  for (int r = tid; r < n; r += BA_THREADS) {
    const int a = r / 6;
    x[r] = (!fixed[a] && cam_start[a + 1] > cam_start[a]) ? bs[r] : 0.0;
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;
  for (int k = 0; k < n; ++k) {
    if (tid == 0) {
      const double d = A[k * ld + k];
      if (!(d > 0.0)) ok_s = 0;
      A[k * ld + k] = sqrt(d);
    }
    __syncthreads();
    if (!ok_s) break;
    const double dk = A[k * ld + k];
    __syncthreads();
    for (int i = k + 1 + tid; i < n; i += BA_THREADS) A[i * ld + k] /= dk;
    __syncthreads();
    for (int i = k + 1 + ty; i < n; i += 16)
      for (int j = k + 1 + tx; j <= i; j += 16) A[i * ld + j] -= A[i * ld + k] * A[j * ld + k];
    __syncthreads();
  }
  const int ok = ok_s;
  if (tid < 32) {
    if (ok) {
      for (int k = 0; k < n; ++k) {           // L y = b
        if (tid == 0) x[k] /= A[k * ld + k];
        __syncwarp();
        const double yk = x[k];
        for (int i = k + 1 + tid; i < n; i += 32) x[i] -= A[i * ld + k] * yk;
        __syncwarp();
      }
      for (int k = n - 1; k >= 0; --k) {      // L^T x = y
        if (tid == 0) x[k] /= A[k * ld + k];
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
  for (int r = tid; r < n; r += BA_THREADS) xp[r] = x[r];
  if (tid < C) {
    const int c = tid;
    const bool act = !fixed[c] && cam_start[c + 1] > cam_start[c];
    if (act && ok) {
      ba_se3_exp_apply(x + 6 * c, poses + 16 * c, trial_poses + 16 * c);
    } else {
      for (int k = 0; k < 16; ++k) trial_poses[16 * c + k] = poses[16 * c + k];
    }
  }
  if (tid == 0) {
    double sc = 0.0;
    for (int c = 0; c < C; ++c) {
      if (fixed[c] || cam_start[c + 1] == cam_start[c]) continue;
      for (int r = 0; r < 6; ++r) sc += x[6 * c + r] * (lam * x[6 * c + r] + bp[6 * c + r]);
    }
    stats[3] = sc;
    stats[2] = (double)ok;
  }
}

__global__ void __launch_bounds__(BA_THREADS)
ba_back_kernel(const int32_t* __restrict__ pt_start, const int32_t* __restrict__ pt_obs,
               const int32_t* __restrict__ obs_cam, int P, int M, const double* __restrict__ W,
               const double* __restrict__ Hll, const double* __restrict__ bl, double lam,
               const double* __restrict__ xp, const double* __restrict__ points,
               double* __restrict__ trial_points, double* __restrict__ scale_part) {
  __shared__ double red[BA_THREADS / 32];
  const int p = blockIdx.x * BA_THREADS + threadIdx.x;
  double sc = 0.0;
  if (p < P) {
    double X[3] = {points[3 * (size_t)p], points[3 * (size_t)p + 1], points[3 * (size_t)p + 2]};
    if (pt_start[p + 1] > pt_start[p]) {
      double cp[3] = {bl[3 * (size_t)p], bl[3 * (size_t)p + 1], bl[3 * (size_t)p + 2]};
      const double b0 = cp[0], b1 = cp[1], b2 = cp[2];
      for (int l = pt_start[p]; l < pt_start[p + 1]; ++l) {
        const int j = pt_obs[l];
        const double* d = xp + 6 * obs_cam[j];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          const double dr = d[r];
          cp[0] -= W[(size_t)(r * 3) * M + j] * dr;
          cp[1] -= W[(size_t)(r * 3 + 1) * M + j] * dr;
          cp[2] -= W[(size_t)(r * 3 + 2) * M + j] * dr;
        }
      }
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

__global__ void ba_final_kernel(const double* __restrict__ chi_part, int nbo,
                                const double* __restrict__ scale_part, int nbp,
                                double* __restrict__ stats) {
  // one warp: strided partial sums in fixed order, shuffle tree
  const int lane = threadIdx.x;
  double c = 0.0, s = 0.0;
  for (int b = lane; b < nbo; b += 32) c += chi_part[b];
  for (int b = lane; b < nbp; b += 32) s += scale_part[b];
  for (int o = 16; o > 0; o >>= 1) {
    c += __shfl_down_sync(0xffffffffu, c, o);
    s += __shfl_down_sync(0xffffffffu, s, o);
  }
  if (lane == 0) { stats[0] = c; stats[1] = s + stats[3]; }
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
  const size_t solve_smem = sizeof(double) * ((size_t)6 * C * (6 * C + 1) + 6 * C);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(ba_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)(sizeof(double) * ((size_t)6 * BA_MAX_CAMS * (6 * BA_MAX_CAMS + 1) + 6 * BA_MAX_CAMS)));
    attr_done = true;
  }

  // adjacency lists
  cudaMemsetAsync(w.pt_start, 0, sizeof(int32_t) * (P + 1), st);
  cudaMemsetAsync(w.cam_count, 0, sizeof(int32_t) * (C + 1), st);
  cudaMemsetAsync(w.bad, 0, sizeof(int32_t), st);
  cudaMemsetAsync(w.maxdiag, 0, sizeof(unsigned long long), st);
  cudaMemsetAsync(w.stats, 0, sizeof(double) * 8, st);
  ba_count_kernel<<<nbo, BA_THREADS, 0, st>>>(obs_cam, obs_pt, M, C, P, w.pt_start, w.cam_count, w.bad);
  int32_t h_bad = 0;
  cudaMemcpyAsync(&h_bad, w.bad, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(count)");
  VO_REQUIRE(!h_bad, VO_ERR_INVALID_ARG, "vo_bundle_adjust: observation index out of range");
  ba_scan_kernel<<<1, 1024, 0, st>>>(w.pt_start, P, w.pt_cursor, w.cam_count, w.cam_start, C);
  ba_fill_pts_kernel<<<nbo, BA_THREADS, 0, st>>>(obs_pt, M, w.pt_cursor, w.pt_obs);
  ba_sort_pts_kernel<<<nbp, BA_THREADS, 0, st>>>(w.pt_start, P, w.pt_obs);
  ba_fill_cams_kernel<<<C, 1024, 0, st>>>(obs_cam, M, w.cam_start, w.cam_obs);

  double* cur_pts = points;  double* tri_pts = w.trial_points;
  double* cur_pos = poses;   double* tri_pos = w.trial_poses;
  double h_stats[4];
  auto chi_of = [&](const double* ps, const double* pt) {
    ba_chi_kernel<<<nbo, BA_THREADS, 0, st>>>(ps, pt, obs_cam, obs_pt, obs_xy, M, focal, cx, cy,
                                              huber_delta, w.chi_part);
  };
  chi_of(cur_pos, cur_pts);
  ba_final_kernel<<<1, 32, 0, st>>>(w.chi_part, nbo, w.scale_part, 0, w.stats);
  cudaMemcpyAsync(h_stats, w.stats, sizeof(double) * 4, cudaMemcpyDeviceToHost, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) return check_launch("vo_bundle_adjust(setup)");
  double current = h_stats[0];
  if (chi2_initial) *chi2_initial = current;

  double lambda = 0.0, ni = 2.0;
  int it = 0;
  for (; it < iterations; ++it) {
    ba_linearize_kernel<<<nbo, BA_THREADS, 0, st>>>(cur_pos, fixed, cur_pts, obs_cam, obs_pt, obs_xy,
                                                    M, focal, cx, cy, huber_delta, w.JA, w.JB, w.we, w.W);
    ba_point_blocks_kernel<<<nbp, BA_THREADS, 0, st>>>(w.pt_start, w.pt_obs, P, M, w.JA, w.we, w.Hll, w.bl);
    ba_camera_blocks_kernel<<<C, BA_THREADS, 0, st>>>(w.cam_start, w.cam_obs, fixed, M, w.JB, w.we,
                                                      w.Hpp, w.bp);
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
      ba_y_kernel<<<nbo, BA_THREADS, 0, st>>>(obs_pt, M, w.Hll, lambda, w.W, w.Y);
      ba_schur_kernel<<<dim3(C, C), BA_THREADS, 0, st>>>(w.cam_start, w.cam_obs, w.pt_start, w.pt_obs,
                                                        obs_cam, obs_pt, fixed, C, M, w.W, w.Y, w.bl,
                                                        w.Hpp, w.bp, lambda, w.S, w.bs);
      ba_solve_kernel<<<1, BA_THREADS, solve_smem, st>>>(w.S, w.bs, w.cam_start, fixed, C, w.bp, lambda,
                                                        cur_pos, tri_pos, w.xp, w.stats);
      ba_back_kernel<<<nbp, BA_THREADS, 0, st>>>(w.pt_start, w.pt_obs, obs_cam, P, M, w.W, w.Hll, w.bl,
                                                lambda, w.xp, cur_pts, tri_pts, w.scale_part);
      chi_of(tri_pos, tri_pts);
      ba_final_kernel<<<1, 32, 0, st>>>(w.chi_part, nbo, w.scale_part, nbp, w.stats);
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
