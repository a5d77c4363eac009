// pnp.cu — K7: PnP-RANSAC for map tracking.
//
// Replaces cv2.solvePnPRansac(X, x, K, zeros(4)) with default arguments (reference
// src/visual_slam.py:231-235): points converted to float32, RANSAC over 5-point samples replayed
// from cv::RNG(2^64-1), EPnP minimal solver (OpenCV epnp.cpp), error = float32 squared
// reprojection distance vs 8 px, 100 iterations with the adaptive bound, strict-greater best
// update; then the SOLVEPNP_ITERATIVE refit (Levenberg-Marquardt on the reprojection error from
// the RANSAC model) on the inliers.  Returned inliers = those of the RANSAC model.
//
// GPU shape: all hypotheses of a problem are solved in parallel (one thread per EPnP solve —
// 12x12 Jacobi eigen-decomposition in local memory), scored in parallel (one warp per
// hypothesis, coalesced point reads), then OpenCV's sequential rule is applied as an ordered
// scan; the refit accumulates the 6x6 normal equations with a block reduction per iteration.
#include <float.h>

#include "linalg.cuh"

namespace vo {

constexpr int PNP_MAX_ITERS = 1024;

struct PnpWs {
  double* Xf;         // [n_total][3] float32-rounded object points
  double* xf;         // [n_total][2] float32-rounded image points
  double* models;     // [F][iters][12]  R (9) + t (3)
  int32_t* samples;   // [F][iters][5]
  int32_t* counts;    // [F][iters]
  int32_t* valid;     // [F][iters]
  int32_t* best;      // [F][4] best index, niters, max_good, n_inl
  uint8_t* mask;      // [n_total]
  int iters;
  size_t bytes;
};

static PnpWs pnp_carve(void* base, long n_total, int n_frames, int iters) {
  PnpWs w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += (bytes + 255) & ~(size_t)255;
    return p;
  };
  const size_t fi = (size_t)n_frames * iters;
  w.Xf = (double*)take((size_t)n_total * 3 * sizeof(double));
  w.xf = (double*)take((size_t)n_total * 2 * sizeof(double));
  w.models = (double*)take(fi * 12 * sizeof(double));
  w.samples = (int32_t*)take(fi * 5 * sizeof(int32_t));
  w.counts = (int32_t*)take(fi * sizeof(int32_t));
  w.valid = (int32_t*)take(fi * sizeof(int32_t));
  w.best = (int32_t*)take((size_t)n_frames * 4 * sizeof(int32_t));
  w.mask = (uint8_t*)take((size_t)n_total);
  w.iters = iters;
  w.bytes = off;
  return w;
}

// frame f of a batch: observations [base, base + n) of the concatenated arrays.  off == nullptr:
// a single problem of n_single observations.
struct PnpFrame { int base, n; };
__device__ __forceinline__ PnpFrame pnp_frame(const int32_t* __restrict__ off, int n_single, int f) {
  PnpFrame fr;
  fr.base = off ? off[f] : 0;
  fr.n = off ? off[f + 1] - off[f] : n_single;
  return fr;
}
// solvepnp.cpp: npoints == model_points -> one plain solvePnP call, no RANSAC: 4 points P3P,
// 5 points EPnP; fewer than 4: error
__device__ __forceinline__ bool pnp_is_ransac(int n) { return n >= 6; }

// ------------------------------------------------------------------ small dense helpers (runtime sizes)
// cyclic Jacobi eigen-decomposition of a symmetric n x n matrix (row-major, destroyed);
// V columns = eigenvectors, d = eigenvalues (unsorted)
__device__ void jacobi_eigen(double* A, double* V, double* d, int n) {
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) V[i * n + j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 40; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < n; ++i) {
      diag += A[i * n + i] * A[i * n + i];
      for (int j = i + 1; j < n; ++j) off += A[i * n + j] * A[i * n + j];
    }
    if (off <= 1e-30 * diag || off == 0.0) break;
    for (int p = 0; p < n - 1; ++p)
      for (int q = p + 1; q < n; ++q) {
        const double apq = A[p * n + q];
        if (fabs(apq) < 1e-300) continue;
        const double theta = (A[q * n + q] - A[p * n + p]) / (2.0 * apq);
        const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(1.0 + theta * theta));
        const double c = rsqrt(1.0 + t * t), s = c * t;
        for (int k = 0; k < n; ++k) {  // columns p,q
          const double akp = A[k * n + p], akq = A[k * n + q];
          A[k * n + p] = c * akp - s * akq;
          A[k * n + q] = s * akp + c * akq;
        }
        for (int k = 0; k < n; ++k) {  // rows p,q
          const double apk = A[p * n + k], aqk = A[q * n + k];
          A[p * n + k] = c * apk - s * aqk;
          A[q * n + k] = s * apk + c * aqk;
        }
        for (int k = 0; k < n; ++k) {
          const double vkp = V[k * n + p], vkq = V[k * n + q];
          V[k * n + p] = c * vkp - s * vkq;
          V[k * n + q] = s * vkp + c * vkq;
        }
      }
  }
  for (int i = 0; i < n; ++i) d[i] = A[i * n + i];
}


// Warp-cooperative form of jacobi_eigen for the 12x12 EPnP matrix.  Every lane of the warp runs
// the same hypothesis (identical local copies); the matrix lives in shared memory and lane k < 12
// owns element k of the two columns / rows a rotation touches.  The (p, q) order, the formulas
// and the column-then-row update order are those of the serial routine, so the result is
// bit-identical to it; only the 3 x 12-element inner loops run in parallel.
struct Jac12Smem {
  double A[144], V[144];
};
__device__ void jacobi_eigen12_warp(double* Aloc, double* Vloc, double* dloc, Jac12Smem& S) {
  const int lane = threadIdx.x & 31;
  const int n = 12;
  __syncwarp();
  for (int e = lane; e < 144; e += 32) {
    S.A[e] = Aloc[e];
    S.V[e] = ((e / n) == (e % n)) ? 1.0 : 0.0;
  }
  __syncwarp();
  for (int sweep = 0; sweep < 40; ++sweep) {
    // serial summation order of the reference loop (cheap: 78 adds), done redundantly per lane
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < n; ++i) {
      diag += S.A[i * n + i] * S.A[i * n + i];
      for (int j = i + 1; j < n; ++j) off += S.A[i * n + j] * S.A[i * n + j];
    }
    if (off <= 1e-30 * diag || off == 0.0) break;
    for (int p = 0; p < n - 1; ++p)
      for (int q = p + 1; q < n; ++q) {
        const double apq = S.A[p * n + q];
        if (fabs(apq) < 1e-300) continue;
        const double theta = (S.A[q * n + q] - S.A[p * n + p]) / (2.0 * apq);
        const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(1.0 + theta * theta));
        const double c = rsqrt(1.0 + t * t), s = c * t;
        __syncwarp();
        if (lane < n) {  // columns p,q
          const double akp = S.A[lane * n + p], akq = S.A[lane * n + q];
          S.A[lane * n + p] = c * akp - s * akq;
          S.A[lane * n + q] = s * akp + c * akq;
          const double vkp = S.V[lane * n + p], vkq = S.V[lane * n + q];
          S.V[lane * n + p] = c * vkp - s * vkq;
          S.V[lane * n + q] = s * vkp + c * vkq;
        }
        __syncwarp();
        if (lane < n) {  // rows p,q
          const double apk = S.A[p * n + lane], aqk = S.A[q * n + lane];
          S.A[p * n + lane] = c * apk - s * aqk;
          S.A[q * n + lane] = s * apk + c * aqk;
        }
        __syncwarp();
      }
  }
  __syncwarp();
  for (int e = 0; e < 144; ++e) Vloc[e] = S.V[e];
  for (int i = 0; i < n; ++i) dloc[i] = S.A[i * n + i];
  __syncwarp();
}

// least squares min |A x - b| for m x n (m >= n, n <= 5) by Householder QR (A, b destroyed)
__device__ bool ls_solve(double* A, double* b, double* x, int m, int n) {
  for (int k = 0; k < n; ++k) {
    double nrm = 0.0;
    for (int i = k; i < m; ++i) nrm += A[i * n + k] * A[i * n + k];
    nrm = sqrt(nrm);
    if (!(nrm > 0.0)) return false;
    const double alpha = A[k * n + k] >= 0 ? -nrm : nrm;
    const double v0 = A[k * n + k] - alpha;
    double vtv = v0 * v0;
    for (int i = k + 1; i < m; ++i) vtv += A[i * n + k] * A[i * n + k];
    if (vtv > 0.0) {
      for (int j = k + 1; j < n; ++j) {
        double dot = v0 * A[k * n + j];
        for (int i = k + 1; i < m; ++i) dot += A[i * n + k] * A[i * n + j];
        const double f = 2.0 * dot / vtv;
        A[k * n + j] -= f * v0;
        for (int i = k + 1; i < m; ++i) A[i * n + j] -= f * A[i * n + k];
      }
      double dot = v0 * b[k];
      for (int i = k + 1; i < m; ++i) dot += A[i * n + k] * b[i];
      const double f = 2.0 * dot / vtv;
      b[k] -= f * v0;
      for (int i = k + 1; i < m; ++i) b[i] -= f * A[i * n + k];
    }
    A[k * n + k] = alpha;
  }
  for (int k = n - 1; k >= 0; --k) {
    double s = b[k];
    for (int j = k + 1; j < n; ++j) s -= A[k * n + j] * x[j];
    x[k] = s / A[k * n + k];
  }
  return true;
}

// ------------------------------------------------------------------ EPnP on a 5-point sample
struct Epnp5 {
  double X[5][3], u[5][2];
  double cws[4][3], alphas[5][4];
  double MtM[144], V[144], ev[12];
  double L[6][10], rho[6];
  int vi[4];   // eigenvector columns of the 4 smallest eigenvalues, ascending
};

__device__ double epnp_R_and_t(const Epnp5& E, const double* betas, double fu, double fv, double uc,
                               double vc, double* R, double* t) {
  double ccs[4][3], pcs[5][3];
  for (int i = 0; i < 4; ++i)
    for (int k = 0; k < 3; ++k) {
      double s = 0.0;
      for (int j = 0; j < 4; ++j) s += betas[j] * E.V[(3 * i + k) * 12 + E.vi[j]];
      ccs[i][k] = s;
    }
  for (int p = 0; p < 5; ++p)
    for (int k = 0; k < 3; ++k)
      pcs[p][k] = E.alphas[p][0] * ccs[0][k] + E.alphas[p][1] * ccs[1][k] +
                  E.alphas[p][2] * ccs[2][k] + E.alphas[p][3] * ccs[3][k];
  if (pcs[0][2] < 0.0)
    for (int p = 0; p < 5; ++p)
      for (int k = 0; k < 3; ++k) pcs[p][k] = -pcs[p][k];
  double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
  for (int p = 0; p < 5; ++p)
    for (int k = 0; k < 3; ++k) { pc0[k] += pcs[p][k]; pw0[k] += E.X[p][k]; }
  for (int k = 0; k < 3; ++k) { pc0[k] /= 5.0; pw0[k] /= 5.0; }
  double A[3][3], Vv[3][3];
  for (int j = 0; j < 3; ++j)
    for (int k = 0; k < 3; ++k) {
      double s = 0.0;
      for (int p = 0; p < 5; ++p) s += (pcs[p][j] - pc0[j]) * (E.X[p][k] - pw0[k]);
      A[j][k] = s;
    }
  // R = U V^T from ABt = U S V^T (one-sided Jacobi: A <- U S, Vv = V)
  jacobi_svd<3>(A, Vv);
  double sg[3];
  int o[3] = {0, 1, 2};
  for (int j = 0; j < 3; ++j) sg[j] = sqrt(A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j]);
  if (sg[o[0]] < sg[o[1]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  if (sg[o[1]] < sg[o[2]]) { int x = o[1]; o[1] = o[2]; o[2] = x; }
  if (sg[o[0]] < sg[o[1]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  double U[3][3];
  for (int k = 0; k < 3; ++k) {
    U[k][0] = A[k][o[0]] / sg[o[0]];
    U[k][1] = A[k][o[1]] / sg[o[1]];
  }
  U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
  U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
  U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
  double W[3][3];
  for (int k = 0; k < 3; ++k) { W[k][0] = Vv[k][o[0]]; W[k][1] = Vv[k][o[1]]; }
  W[0][2] = W[1][0] * W[2][1] - W[2][0] * W[1][1];
  W[1][2] = W[2][0] * W[0][1] - W[0][0] * W[2][1];
  W[2][2] = W[0][0] * W[1][1] - W[1][0] * W[0][1];
  // both completed bases are right-handed, so det(U W^T) = +1 (OpenCV flips the last row of R
  // when det < 0, which yields the same proper rotation)
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) R[3 * i + j] = U[i][0] * W[j][0] + U[i][1] * W[j][1] + U[i][2] * W[j][2];
  for (int i = 0; i < 3; ++i) t[i] = pc0[i] - (R[3 * i] * pw0[0] + R[3 * i + 1] * pw0[1] + R[3 * i + 2] * pw0[2]);
  double err = 0.0;
  for (int p = 0; p < 5; ++p) {
    const double xc = R[0] * E.X[p][0] + R[1] * E.X[p][1] + R[2] * E.X[p][2] + t[0];
    const double yc = R[3] * E.X[p][0] + R[4] * E.X[p][1] + R[5] * E.X[p][2] + t[1];
    const double inv = 1.0 / (R[6] * E.X[p][0] + R[7] * E.X[p][1] + R[8] * E.X[p][2] + t[2]);
    const double du = E.u[p][0] - (uc + fu * xc * inv), dv = E.u[p][1] - (vc + fv * yc * inv);
    err += sqrt(du * du + dv * dv);
  }
  return err / 5.0;
}

__device__ void epnp_gauss_newton(const Epnp5& E, double* b) {
  for (int it = 0; it < 5; ++it) {
    double A[24], r[6], x[4];
    for (int i = 0; i < 6; ++i) {
      const double* Lr = E.L[i];
      A[4 * i + 0] = 2 * Lr[0] * b[0] + Lr[1] * b[1] + Lr[3] * b[2] + Lr[6] * b[3];
      A[4 * i + 1] = Lr[1] * b[0] + 2 * Lr[2] * b[1] + Lr[4] * b[2] + Lr[7] * b[3];
      A[4 * i + 2] = Lr[3] * b[0] + Lr[4] * b[1] + 2 * Lr[5] * b[2] + Lr[8] * b[3];
      A[4 * i + 3] = Lr[6] * b[0] + Lr[7] * b[1] + Lr[8] * b[2] + 2 * Lr[9] * b[3];
      r[i] = E.rho[i] - (Lr[0] * b[0] * b[0] + Lr[1] * b[0] * b[1] + Lr[2] * b[1] * b[1] +
                         Lr[3] * b[0] * b[2] + Lr[4] * b[1] * b[2] + Lr[5] * b[2] * b[2] +
                         Lr[6] * b[0] * b[3] + Lr[7] * b[1] * b[3] + Lr[8] * b[2] * b[3] +
                         Lr[9] * b[3] * b[3]);
    }
    if (!ls_solve(A, r, x, 6, 4)) return;
    for (int k = 0; k < 4; ++k) b[k] += x[k];
  }
}

// full EPnP (epnp.cpp compute_pose) for 5 points; returns false on numerical failure
__device__ bool epnp5(Epnp5& E, const double* K, double* R, double* t, Jac12Smem& S) {
  const double fu = K[0], fv = K[4], uc = K[2], vc = K[5];
  // control points: centroid + principal axes scaled by sqrt(lambda / n)
  double c0[3] = {0, 0, 0};
  for (int p = 0; p < 5; ++p)
    for (int k = 0; k < 3; ++k) c0[k] += E.X[p][k];
  for (int k = 0; k < 3; ++k) c0[k] /= 5.0;
  double C[9], CV[9], cd[3];
  for (int j = 0; j < 3; ++j)
    for (int k = 0; k < 3; ++k) {
      double s = 0.0;
      for (int p = 0; p < 5; ++p) s += (E.X[p][j] - c0[j]) * (E.X[p][k] - c0[k]);
      C[3 * j + k] = s;
    }
  jacobi_eigen(C, CV, cd, 3);
  int o[3] = {0, 1, 2};
  if (cd[o[0]] < cd[o[1]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  if (cd[o[1]] < cd[o[2]]) { int x = o[1]; o[1] = o[2]; o[2] = x; }
  if (cd[o[0]] < cd[o[1]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  for (int k = 0; k < 3; ++k) E.cws[0][k] = c0[k];
  for (int i = 1; i < 4; ++i) {
    const double kk = sqrt(fmax(cd[o[i - 1]], 0.0) / 5.0);
    for (int k = 0; k < 3; ++k) E.cws[i][k] = c0[k] + kk * CV[k * 3 + o[i - 1]];
  }
  // barycentric coordinates: CC = [c1-c0 c2-c0 c3-c0] (columns); alphas[1..3] = CC^-1 (p - c0)
  double CC[9], CI[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) CC[3 * i + j] = E.cws[j + 1][i] - E.cws[0][i];
  const double det = det3(CC);
  if (!(fabs(det) > 1e-300) || !isfinite(det)) return false;
  CI[0] = (CC[4] * CC[8] - CC[5] * CC[7]) / det; CI[1] = (CC[2] * CC[7] - CC[1] * CC[8]) / det;
  CI[2] = (CC[1] * CC[5] - CC[2] * CC[4]) / det; CI[3] = (CC[5] * CC[6] - CC[3] * CC[8]) / det;
  CI[4] = (CC[0] * CC[8] - CC[2] * CC[6]) / det; CI[5] = (CC[2] * CC[3] - CC[0] * CC[5]) / det;
  CI[6] = (CC[3] * CC[7] - CC[4] * CC[6]) / det; CI[7] = (CC[1] * CC[6] - CC[0] * CC[7]) / det;
  CI[8] = (CC[0] * CC[4] - CC[1] * CC[3]) / det;
  for (int p = 0; p < 5; ++p) {
    const double d0 = E.X[p][0] - c0[0], d1 = E.X[p][1] - c0[1], d2 = E.X[p][2] - c0[2];
    for (int j = 0; j < 3; ++j) E.alphas[p][1 + j] = CI[3 * j] * d0 + CI[3 * j + 1] * d1 + CI[3 * j + 2] * d2;
    E.alphas[p][0] = 1.0 - E.alphas[p][1] - E.alphas[p][2] - E.alphas[p][3];
  }
  // M^T M (12x12) accumulated row by row
  for (int e = 0; e < 144; ++e) E.MtM[e] = 0.0;
  for (int p = 0; p < 5; ++p) {
    double m1[12], m2[12];
    for (int j = 0; j < 4; ++j) {
      const double a = E.alphas[p][j];
      m1[3 * j] = a * fu; m1[3 * j + 1] = 0.0; m1[3 * j + 2] = a * (uc - E.u[p][0]);
      m2[3 * j] = 0.0; m2[3 * j + 1] = a * fv; m2[3 * j + 2] = a * (vc - E.u[p][1]);
    }
    for (int i = 0; i < 12; ++i)
      for (int j = 0; j < 12; ++j) E.MtM[12 * i + j] += m1[i] * m1[j] + m2[i] * m2[j];
  }
  jacobi_eigen12_warp(E.MtM, E.V, E.ev, S);
  {  // four smallest eigenvalues, ascending (selection)
    bool used[12];
    for (int i = 0; i < 12; ++i) used[i] = false;
    for (int k = 0; k < 4; ++k) {
      int bi = -1;
      for (int i = 0; i < 12; ++i)
        if (!used[i] && (bi < 0 || E.ev[i] < E.ev[bi])) bi = i;
      used[bi] = true;
      E.vi[k] = bi;
    }
  }
  // L (6x10) and rho
  const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
  for (int i = 0; i < 6; ++i) {
    double dv[4][3];
    for (int k = 0; k < 4; ++k)
      for (int c = 0; c < 3; ++c)
        dv[k][c] = E.V[(3 * pa[i] + c) * 12 + E.vi[k]] - E.V[(3 * pb[i] + c) * 12 + E.vi[k]];
    auto dot = [&](int a, int b) { return dv[a][0] * dv[b][0] + dv[a][1] * dv[b][1] + dv[a][2] * dv[b][2]; };
    double* Lr = E.L[i];
    Lr[0] = dot(0, 0); Lr[1] = 2 * dot(0, 1); Lr[2] = dot(1, 1); Lr[3] = 2 * dot(0, 2);
    Lr[4] = 2 * dot(1, 2); Lr[5] = dot(2, 2); Lr[6] = 2 * dot(0, 3); Lr[7] = 2 * dot(1, 3);
    Lr[8] = 2 * dot(2, 3); Lr[9] = dot(3, 3);
    double s = 0.0;
    for (int c = 0; c < 3; ++c) { const double d = E.cws[pa[i]][c] - E.cws[pb[i]][c]; s += d * d; }
    E.rho[i] = s;
  }
  // three linearised starts, Gauss-Newton, keep the smallest reprojection error (N = 1,2,3)
  double bestR[9], bestT[3], best_err = 0.0;
  bool have = false;
  for (int N = 1; N <= 3; ++N) {
    const int ncol = (N == 1) ? 4 : (N == 2 ? 3 : 5);
    const int cols1[4] = {0, 1, 3, 6}, cols2[3] = {0, 1, 2}, cols3[5] = {0, 1, 2, 3, 4};
    const int* cols = (N == 1) ? cols1 : (N == 2 ? cols2 : cols3);
    double A[30], r[6], x[5], b[4];
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j < ncol; ++j) A[i * ncol + j] = E.L[i][cols[j]];
      r[i] = E.rho[i];
    }
    if (!ls_solve(A, r, x, 6, ncol)) continue;
    if (N == 1) {
      if (x[0] < 0) { b[0] = sqrt(-x[0]); b[1] = -x[1] / b[0]; b[2] = -x[2] / b[0]; b[3] = -x[3] / b[0]; }
      else { b[0] = sqrt(x[0]); b[1] = x[1] / b[0]; b[2] = x[2] / b[0]; b[3] = x[3] / b[0]; }
    } else {
      if (x[0] < 0) { b[0] = sqrt(-x[0]); b[1] = (x[2] < 0) ? sqrt(-x[2]) : 0.0; }
      else { b[0] = sqrt(x[0]); b[1] = (x[2] > 0) ? sqrt(x[2]) : 0.0; }
      if (x[1] < 0) b[0] = -b[0];
      b[2] = (N == 3) ? x[3] / b[0] : 0.0;
      b[3] = 0.0;
    }
    epnp_gauss_newton(E, b);
    double R[9], t[3];
    const double err = epnp_R_and_t(E, b, fu, fv, uc, vc, R, t);
    if (!isfinite(err)) continue;
    if (!have || err < best_err) {
      have = true; best_err = err;
      for (int e = 0; e < 9; ++e) bestR[e] = R[e];
      for (int e = 0; e < 3; ++e) bestT[e] = t[e];
    }
  }
  if (!have) return false;
  for (int e = 0; e < 9; ++e) R[e] = bestR[e];
  for (int e = 0; e < 3; ++e) t[e] = bestT[e];
  return true;
}

// ------------------------------------------------------------------ P3P (the npoints == 4 case)
struct Cplx { double re, im; };
__device__ __forceinline__ Cplx cmul(Cplx a, Cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
__device__ __forceinline__ Cplx cdiv(Cplx a, Cplx b) {
  const double d = b.re * b.re + b.im * b.im;
  return {(a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d};
}
// the four roots of c[4] v^4 + ... + c[0] (c[4] != 0): Durand-Kerner from a circle of the Cauchy
// bound radius
__device__ void quartic_roots(const double* c, Cplx* z) {
  double a[4], bound = 0.0;
  for (int k = 0; k < 4; ++k) { a[k] = c[k] / c[4]; bound = fmax(bound, fabs(a[k])); }
  bound += 1.0;
  for (int k = 0; k < 4; ++k) {
    const double ang = 0.4 + 1.5707963267948966 * k;
    z[k] = {0.5 * bound * cos(ang), 0.5 * bound * sin(ang)};
  }
  for (int it = 0; it < 200; ++it) {
    double moved = 0.0;
    for (int k = 0; k < 4; ++k) {
      Cplx pv = {1.0, 0.0};
      for (int d = 3; d >= 0; --d) { pv = cmul(pv, z[k]); pv.re += a[d]; }
      Cplx den = {1.0, 0.0};
      for (int m = 0; m < 4; ++m)
        if (m != k) den = cmul(den, Cplx{z[k].re - z[m].re, z[k].im - z[m].im});
      if (den.re == 0.0 && den.im == 0.0) den.re = 1e-300;
      const Cplx st = cdiv(pv, den);
      z[k].re -= st.re; z[k].im -= st.im;
      moved = fmax(moved, fabs(st.re) + fabs(st.im));
    }
    if (moved < 1e-15 * bound) break;
  }
}

// rotation taking the triad of X to the triad of C (frames from P0->P1 and the plane normal)
__device__ void triad_frame(const double (*P)[3], double* F /*3x3 columns e1,e2,e3*/) {
  double e1[3], d2[3], e3[3], e2[3];
  for (int k = 0; k < 3; ++k) { e1[k] = P[1][k] - P[0][k]; d2[k] = P[2][k] - P[0][k]; }
  double n = rsqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
  for (int k = 0; k < 3; ++k) e1[k] *= n;
  e3[0] = e1[1] * d2[2] - e1[2] * d2[1]; e3[1] = e1[2] * d2[0] - e1[0] * d2[2]; e3[2] = e1[0] * d2[1] - e1[1] * d2[0];
  n = rsqrt(e3[0] * e3[0] + e3[1] * e3[1] + e3[2] * e3[2]);
  for (int k = 0; k < 3; ++k) e3[k] *= n;
  e2[0] = e3[1] * e1[2] - e3[2] * e1[1]; e2[1] = e3[2] * e1[0] - e3[0] * e1[2]; e2[2] = e3[0] * e1[1] - e3[1] * e1[0];
  for (int k = 0; k < 3; ++k) { F[3 * k] = e1[k]; F[3 * k + 1] = e2[k]; F[3 * k + 2] = e3[k]; }
}

// cv2.solvePnP(SOLVEPNP_P3P) on exactly four points: Grunert's quartic in v = s3 / s1 on the first
// three (any exact P3P has the same solution set as OpenCV's p3p.cpp), depths polished by Newton
// on the cosine-law system, pose from the two triads; the solution with the smallest squared
// reprojection error of the FOURTH point wins.  false = no real solution with positive depths.
__device__ bool pnp_p3p4(const double* __restrict__ Xf, const double* __restrict__ xf,
                         const double* __restrict__ K, double* R, double* t) {
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  double y[3][3], X3[3][3];
  for (int i = 0; i < 3; ++i) {
    const double a = (xf[2 * i] - cx) / fx, b = (xf[2 * i + 1] - cy) / fy;
    const double n = rsqrt(a * a + b * b + 1.0);
    y[i][0] = a * n; y[i][1] = b * n; y[i][2] = n;
    for (int k = 0; k < 3; ++k) X3[i][k] = Xf[3 * i + k];
  }
  auto d2 = [&](int i, int j) {
    double s = 0.0;
    for (int k = 0; k < 3; ++k) s += (X3[i][k] - X3[j][k]) * (X3[i][k] - X3[j][k]);
    return s;
  };
  auto dt = [&](int i, int j) { return y[i][0] * y[j][0] + y[i][1] * y[j][1] + y[i][2] * y[j][2]; };
  const double a2 = d2(1, 2), b2 = d2(0, 2), c2 = d2(0, 1);
  const double ca = dt(1, 2), cb = dt(0, 2), cg = dt(0, 1);
  if (!(a2 > 0.0) || !(b2 > 0.0) || !(c2 > 0.0)) return false;
  const double q = (a2 - c2) / b2, pp = (a2 + c2) / b2;
  double c[5];
  c[4] = (q - 1) * (q - 1) - 4 * c2 / b2 * ca * ca;
  c[3] = 4 * (q * (1 - q) * cb - (1 - pp) * ca * cg + 2 * c2 / b2 * ca * ca * cb);
  c[2] = 2 * (q * q - 1 + 2 * q * q * cb * cb + 2 * (b2 - c2) / b2 * ca * ca - 4 * pp * ca * cb * cg +
              2 * (b2 - a2) / b2 * cg * cg);
  c[1] = 4 * (-q * (1 + q) * cb + 2 * a2 / b2 * cg * cg * cb - (1 - pp) * ca * cg);
  c[0] = (1 + q) * (1 + q) - 4 * a2 / b2 * cg * cg;
  const double cmax = fmax(fmax(fabs(c[0]), fabs(c[1])), fmax(fabs(c[2]), fmax(fabs(c[3]), fabs(c[4]))));
  if (!(fabs(c[4]) > 1e-14 * cmax)) return false;
  Cplx z[4];
  quartic_roots(c, z);
  bool have = false;
  double best_e = 0.0;
  for (int r = 0; r < 4; ++r) {
    const double v = z[r].re;
    if (fabs(z[r].im) > 1e-7 * fmax(1.0, fabs(v)) || !(v > 0.0)) continue;
    const double den = 2 * (cg - v * ca);
    if (fabs(den) < 1e-14) continue;
    const double u = ((q - 1) * v * v - 2 * q * cb * v + 1 + q) / den;
    const double d1 = 1 + v * v - 2 * v * cb;
    if (!(u > 0.0) || !(d1 > 0.0)) continue;
    double sd[3];
    sd[0] = sqrt(b2 / d1); sd[1] = u * sd[0]; sd[2] = v * sd[0];
    double f[3] = {0, 0, 0};
    for (int it = 0; it < 10; ++it) {
      f[0] = sd[1] * sd[1] + sd[2] * sd[2] - 2 * sd[1] * sd[2] * ca - a2;
      f[1] = sd[0] * sd[0] + sd[2] * sd[2] - 2 * sd[0] * sd[2] * cb - b2;
      f[2] = sd[0] * sd[0] + sd[1] * sd[1] - 2 * sd[0] * sd[1] * cg - c2;
      const double J[9] = {0, 2 * sd[1] - 2 * sd[2] * ca, 2 * sd[2] - 2 * sd[1] * ca,
                           2 * sd[0] - 2 * sd[2] * cb, 0, 2 * sd[2] - 2 * sd[0] * cb,
                           2 * sd[0] - 2 * sd[1] * cg, 2 * sd[1] - 2 * sd[0] * cg, 0};
      const double det = J[0] * (J[4] * J[8] - J[5] * J[7]) - J[1] * (J[3] * J[8] - J[5] * J[6]) +
                         J[2] * (J[3] * J[7] - J[4] * J[6]);
      if (fabs(det) < 1e-300) break;
      // Cramer
      const double dx = (f[0] * (J[4] * J[8] - J[5] * J[7]) - J[1] * (f[1] * J[8] - J[5] * f[2]) +
                         J[2] * (f[1] * J[7] - J[4] * f[2])) / det;
      const double dy = (J[0] * (f[1] * J[8] - J[5] * f[2]) - f[0] * (J[3] * J[8] - J[5] * J[6]) +
                         J[2] * (J[3] * f[2] - f[1] * J[6])) / det;
      const double dz = (J[0] * (J[4] * f[2] - f[1] * J[7]) - J[1] * (J[3] * f[2] - f[1] * J[6]) +
                         f[0] * (J[3] * J[7] - J[4] * J[6])) / det;
      sd[0] -= dx; sd[1] -= dy; sd[2] -= dz;
    }
    f[0] = sd[1] * sd[1] + sd[2] * sd[2] - 2 * sd[1] * sd[2] * ca - a2;
    f[1] = sd[0] * sd[0] + sd[2] * sd[2] - 2 * sd[0] * sd[2] * cb - b2;
    f[2] = sd[0] * sd[0] + sd[1] * sd[1] - 2 * sd[0] * sd[1] * cg - c2;
    const double fmx = fmax(fabs(f[0]), fmax(fabs(f[1]), fabs(f[2])));
    if (!(sd[0] > 0 && sd[1] > 0 && sd[2] > 0) || !(fmx <= 1e-6 * fmax(a2, fmax(b2, c2)))) continue;
    double C[3][3], FC[9], FX[9], Rc[9], tc[3];
    for (int i = 0; i < 3; ++i)
      for (int k = 0; k < 3; ++k) C[i][k] = y[i][k] * sd[i];
    triad_frame(C, FC);
    triad_frame(X3, FX);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        Rc[3 * i + j] = FC[3 * i] * FX[3 * j] + FC[3 * i + 1] * FX[3 * j + 1] + FC[3 * i + 2] * FX[3 * j + 2];
    for (int i = 0; i < 3; ++i)
      tc[i] = C[0][i] - (Rc[3 * i] * X3[0][0] + Rc[3 * i + 1] * X3[0][1] + Rc[3 * i + 2] * X3[0][2]);
    // fourth point
    const double* X4 = Xf + 9;
    const double px = Rc[0] * X4[0] + Rc[1] * X4[1] + Rc[2] * X4[2] + tc[0];
    const double py = Rc[3] * X4[0] + Rc[4] * X4[1] + Rc[5] * X4[2] + tc[1];
    const double pz = Rc[6] * X4[0] + Rc[7] * X4[1] + Rc[8] * X4[2] + tc[2];
    if (pz == 0.0) continue;
    const double eu = fx * px / pz + cx - xf[6], ev = fy * py / pz + cy - xf[7];
    const double e = eu * eu + ev * ev;
    if (!have || e < best_e) {
      have = true; best_e = e;
      for (int k = 0; k < 9; ++k) R[k] = Rc[k];
      for (int k = 0; k < 3; ++k) t[k] = tc[k];
    }
  }
  return have;
}

// ------------------------------------------------------------------ kernels (grid.y = frame)
__global__ void pnp_prepare_kernel(const double* __restrict__ X, const double* __restrict__ x,
                                   long n_total, PnpWs w) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_total) return;
  // solvePnPRansac converts its inputs to CV_32F before anything else
  for (int k = 0; k < 3; ++k) w.Xf[3 * i + k] = (double)__double2float_rn(X[3 * i + k]);
  for (int k = 0; k < 2; ++k) w.xf[2 * i + k] = (double)__double2float_rn(x[2 * i + k]);
}

__device__ __forceinline__ uint32_t pnp_rng_next(uint64_t& s) {
  s = (uint64_t)(uint32_t)s * 4164903690ull + (uint32_t)(s >> 32);
  return (uint32_t)s;
}

// one thread per frame replays cv::RNG(2^64-1) for its point count
__global__ void pnp_sample_kernel(const int32_t* __restrict__ off, int n_single, int n_frames,
                                  PnpWs w) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_frames) return;
  const PnpFrame fr = pnp_frame(off, n_single, f);
  if (!pnp_is_ransac(fr.n)) return;
  const int iters = w.iters;
  int32_t* out = w.samples + (size_t)f * iters * 5;
  uint64_t s = 0xffffffffffffffffull;
  for (int it = 0; it < iters; ++it) {
    int32_t idx[5];
    for (int i = 0; i < 5;) {
      const int v = (int)(pnp_rng_next(s) % (uint32_t)fr.n);
      bool dup = false;
      for (int j = 0; j < i; ++j) dup |= (idx[j] == v);
      idx[i] = v;
      if (!dup) ++i;
    }
    for (int k = 0; k < 5; ++k) out[5 * it + k] = idx[k];
  }
}

// one WARP per RANSAC iteration: EPnP on its 5-point sample.  All lanes carry the hypothesis
// (identical scalar work, same latency as one thread); the 12x12 eigen-decomposition, which is
// ~90 % of the serial solve, is spread over the lanes in shared memory.
// Frames with 4 or 5 observations have ONE hypothesis (iteration 0): P3P / EPnP on all points.
constexpr int PNP_SOLVE_WARPS = 4;
__global__ void __launch_bounds__(PNP_SOLVE_WARPS * 32)
pnp_solve_kernel(const double* __restrict__ K, const int32_t* __restrict__ off, int n_single, PnpWs w) {
  __shared__ Jac12Smem sm[PNP_SOLVE_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int it = blockIdx.x * PNP_SOLVE_WARPS + warp, f = blockIdx.y;
  const int iters = w.iters;
  if (it >= iters) return;
  const PnpFrame fr = pnp_frame(off, n_single, f);
  const size_t hyp = (size_t)f * iters + it;
  if (!pnp_is_ransac(fr.n) && (it > 0 || fr.n < 4)) {
    if (lane == 0) w.valid[hyp] = 0;
    return;
  }
  double Kl[9], R[9], t[3];
  for (int e = 0; e < 9; ++e) Kl[e] = K[e];
  bool ok;
  if (fr.n == 4) {
    ok = pnp_p3p4(w.Xf + 3 * (size_t)fr.base, w.xf + 2 * (size_t)fr.base, Kl, R, t);
  } else {
    Epnp5 E;
    for (int p = 0; p < 5; ++p) {
      const int i = fr.base + (fr.n == 5 ? p : w.samples[hyp * 5 + p]);
      for (int k = 0; k < 3; ++k) E.X[p][k] = w.Xf[3 * (size_t)i + k];
      for (int k = 0; k < 2; ++k) E.u[p][k] = w.xf[2 * (size_t)i + k];
    }
    ok = epnp5(E, Kl, R, t, sm[warp]);
  }
  if (lane == 0) {
    w.valid[hyp] = ok ? 1 : 0;
    if (ok) {
      for (int e = 0; e < 9; ++e) w.models[12 * hyp + e] = R[e];
      for (int e = 0; e < 3; ++e) w.models[12 * hyp + 9 + e] = t[e];
    }
  }
}

// PnPRansacCallback::computeError: projectPoints in double, result stored as float32, error =
// float32 of the squared pixel distance between float32 points
__device__ __forceinline__ float pnp_error(const double* __restrict__ Rt, const double* __restrict__ K,
                                           const double* __restrict__ Xp, const double* __restrict__ up) {
  const double X = Xp[0], Y = Xp[1], Z = Xp[2];
  double xc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[0], X), __dmul_rn(Rt[1], Y)), __dmul_rn(Rt[2], Z)), Rt[9]);
  double yc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[3], X), __dmul_rn(Rt[4], Y)), __dmul_rn(Rt[5], Z)), Rt[10]);
  double zc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[6], X), __dmul_rn(Rt[7], Y)), __dmul_rn(Rt[8], Z)), Rt[11]);
  zc = zc != 0.0 ? __ddiv_rn(1.0, zc) : 1.0;
  xc = __dmul_rn(xc, zc);
  yc = __dmul_rn(yc, zc);
  const float pu = __double2float_rn(__dadd_rn(__dmul_rn(xc, K[0]), K[2]));
  const float pv = __double2float_rn(__dadd_rn(__dmul_rn(yc, K[4]), K[5]));
  const float du = __fsub_rn((float)up[0], pu), dv = __fsub_rn((float)up[1], pv);
  return __double2float_rn(__dadd_rn(__dmul_rn((double)du, (double)du), __dmul_rn((double)dv, (double)dv)));
}

__global__ void __launch_bounds__(256)
pnp_score_kernel(const double* __restrict__ K, const int32_t* __restrict__ off, int n_single,
                 float thr2, PnpWs w) {
  const int it = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31, f = blockIdx.y;
  const int iters = w.iters;
  if (it >= iters) return;
  const PnpFrame fr = pnp_frame(off, n_single, f);
  const size_t hyp = (size_t)f * iters + it;
  if (!pnp_is_ransac(fr.n) || !w.valid[hyp]) { if (lane == 0) w.counts[hyp] = 0; return; }
  double Rt[12], Kl[9];
  for (int e = 0; e < 12; ++e) Rt[e] = w.models[12 * hyp + e];
  for (int e = 0; e < 9; ++e) Kl[e] = K[e];
  const double* Xb = w.Xf + 3 * (size_t)fr.base;
  const double* xb = w.xf + 2 * (size_t)fr.base;
  int cnt = 0;
  for (int i = lane; i < fr.n; i += 32) cnt += pnp_error(Rt, Kl, Xb + 3 * i, xb + 2 * i) <= thr2;
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (lane == 0) w.counts[hyp] = cnt;
}

__device__ int pnp_update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = fmin(fmax(p, 0.0), 1.0);
  ep = fmin(fmax(ep, 0.0), 1.0);
  double num = fmax(1.0 - p, DBL_MIN);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < DBL_MIN) return 0;
  num = log(num);
  denom = log(denom);
  return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : (int)rint(num / denom);
}

// ordered scan (OpenCV's sequential rule) + inlier mask of the winner + ascending index list.
// One block per frame; inliers of frame f are written at inliers[base ...] (frame-local indices).
__global__ void __launch_bounds__(1024)
pnp_select_kernel(const double* __restrict__ K, const int32_t* __restrict__ off, int n_single,
                  double conf, float thr2, PnpWs w, int32_t* __restrict__ inliers,
                  int32_t* __restrict__ n_inl, int32_t* __restrict__ ok) {
  __shared__ int s_best, s_base, warp_tot[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, f = blockIdx.x;
  const PnpFrame fr = pnp_frame(off, n_single, f);
  const int n = fr.n, iters = w.iters;
  const size_t h0 = (size_t)f * iters;
  int32_t* best4 = w.best + 4 * f;
  int32_t* inl = inliers + fr.base;
  if (!pnp_is_ransac(n)) {
    // 4 or 5 points: the single solvePnP model, every point reported as an inlier, no refit
    const bool good = n >= 4 && w.valid[h0];
    if (tid == 0) {
      best4[0] = good ? 0 : -1; best4[1] = 0; best4[2] = good ? n : 0; best4[3] = -1;   // -1: skip the refit
      n_inl[f] = good ? n : 0;
      ok[f] = good ? 1 : 0;
    }
    if (good && tid < n) inl[tid] = tid;
    return;
  }
  if (tid == 0) {
    int niters = iters, maxgood = 0, best = -1;
    for (int it = 0; it < iters && it < niters; ++it) {
      const int c = w.counts[h0 + it];
      if (w.valid[h0 + it] && c > max(maxgood, 4)) {
        maxgood = c; best = it;
        niters = pnp_update_num_iters(conf, (double)(n - c) / n, 5, niters);
      }
    }
    s_best = best; s_base = 0;
    best4[0] = best; best4[1] = niters; best4[2] = maxgood;
  }
  __syncthreads();
  const int best = s_best;
  if (best < 0) {
    if (tid == 0) { n_inl[f] = 0; ok[f] = 0; best4[3] = 0; }
    return;
  }
  double Rt[12], Kl[9];
  for (int e = 0; e < 12; ++e) Rt[e] = w.models[12 * (h0 + best) + e];
  for (int e = 0; e < 9; ++e) Kl[e] = K[e];
  const double* Xb = w.Xf + 3 * (size_t)fr.base;
  const double* xb = w.xf + 2 * (size_t)fr.base;
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int i = i0 + tid;
    const bool in = i < n && pnp_error(Rt, Kl, Xb + 3 * i, xb + 2 * i) <= thr2;
    if (i < n) w.mask[fr.base + i] = in;
    const unsigned bal = __ballot_sync(0xffffffffu, in);
    if (lane == 0) warp_tot[wid] = __popc(bal);
    __syncthreads();
    int o = s_base;
    for (int k = 0; k < wid; ++k) o += warp_tot[k];
    o += __popc(bal & ((1u << lane) - 1u));
    if (in) inl[o] = i;
    __syncthreads();
    if (tid == 0) { int tot = 0; for (int k = 0; k < 32; ++k) tot += warp_tot[k]; s_base += tot; }
    __syncthreads();
  }
  if (tid == 0) { n_inl[f] = s_base; ok[f] = 1; best4[3] = s_base; }
}

// ------------------------------------------------------------------ iterative refit (LM)
// minimise sum |proj(R X + t) - x|^2 over the inliers, updates R <- exp(w) R, t <- t + d.
// The optimum does not depend on the parametrisation, so it equals the one cv2's
// SOLVEPNP_ITERATIVE Levenberg-Marquardt (over rvec,tvec) converges to.
constexpr int REFINE_THREADS = 256;

__device__ void block_reduce(double* vals, int nvals, double* smem /*[8][nvals]*/) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int k = 0; k < nvals; ++k) {
    double v = vals[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) smem[wid * nvals + k] = v;
  }
  __syncthreads();
  for (int k = 0; k < nvals; ++k) {
    double s = 0.0;
    for (int wv = 0; wv < REFINE_THREADS / 32; ++wv) s += smem[wv * nvals + k];
    vals[k] = s;
  }
  __syncthreads();
}

__device__ void exp_so3(const double* w, double* R /*3x3*/) {
  const double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  double a, b;
  if (th < 1e-8) { a = 1.0 - th * th / 6.0; b = 0.5 - th * th / 24.0; }
  else { a = sin(th) / th; b = (1.0 - cos(th)) / (th * th); }
  const double Kx[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double k2 = 0.0;
      for (int k = 0; k < 3; ++k) k2 += Kx[3 * i + k] * Kx[3 * k + j];
      R[3 * i + j] = (i == j ? 1.0 : 0.0) + a * Kx[3 * i + j] + b * k2;
    }
}

// cv2.Rodrigues(R) -> rvec (matrix to vector)
__device__ void log_so3(const double* R, double* r) {
  double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
  const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  const double th = acos(c);
  if (s < 1e-5) {
    if (c > 0) { r[0] = r[1] = r[2] = 0.0; return; }
    double t0 = sqrt(fmax((R[0] + 1) * 0.5, 0.0)), t1 = sqrt(fmax((R[4] + 1) * 0.5, 0.0)) * (R[1] < 0 ? -1.0 : 1.0),
           t2 = sqrt(fmax((R[8] + 1) * 0.5, 0.0)) * (R[2] < 0 ? -1.0 : 1.0);
    if (fabs(t0) < fabs(t1) && fabs(t0) < fabs(t2) && ((R[5] > 0) != (t1 * t2 > 0))) t2 = -t2;
    const double sc = th / sqrt(t0 * t0 + t1 * t1 + t2 * t2);
    r[0] = t0 * sc; r[1] = t1 * sc; r[2] = t2 * sc;
    return;
  }
  const double vth = 0.5 / s * th;
  r[0] = rx * vth; r[1] = ry * vth; r[2] = rz * vth;
}

__device__ bool chol_solve6(const double* H, const double* g, double lambda, double* x) {
  double L[36];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = H[6 * i + j] + (i == j ? lambda * H[6 * i + i] : 0.0);
      for (int k = 0; k < j; ++k) s -= L[6 * i + k] * L[6 * j + k];
      if (i == j) { if (!(s > 0.0)) return false; L[6 * i + i] = sqrt(s); }
      else L[6 * i + j] = s / L[6 * j + j];
    }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double s = -g[i];
    for (int k = 0; k < i; ++k) s -= L[6 * i + k] * y[k];
    y[i] = s / L[6 * i + i];
  }
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < 6; ++k) s -= L[6 * k + i] * x[k];
    x[i] = s / L[6 * i + i];
  }
  return true;
}

__global__ void __launch_bounds__(REFINE_THREADS)
pnp_refine_kernel(const double* __restrict__ K, const int32_t* __restrict__ off, int n_single, PnpWs w,
                  const int32_t* __restrict__ inliers_all, double* __restrict__ rvec_all,
                  double* __restrict__ tvec_all) {
  __shared__ double red[8 * 28];
  __shared__ double sR[9], st[3], cR[9], ct[3];
  __shared__ double s_cost, s_lambda;
  __shared__ int s_state;   // 0 continue, 1 converged/stop
  const int f = blockIdx.x;
  const PnpFrame fr = pnp_frame(off, n_single, f);
  const int best = w.best[4 * f], ni = w.best[4 * f + 3];
  double* rvec = rvec_all + 3 * f;
  double* tvec = tvec_all + 3 * f;
  const int tid = threadIdx.x;
  if (best < 0) {
    if (tid < 3) { rvec[tid] = 0.0; tvec[tid] = 0.0; }
    return;
  }
  const int32_t* inliers = inliers_all + fr.base;
  const size_t hyp = (size_t)f * w.iters + best;
  if (tid < 9) sR[tid] = w.models[12 * hyp + tid];
  if (tid < 3) st[tid] = w.models[12 * hyp + 9 + tid];
  if (tid == 0) { s_lambda = 1e-3; s_state = 0; s_cost = -1.0; }
  __syncthreads();
  if (ni < 0) {   // 4 / 5 points: solvePnP's own answer, no iterative refit
    if (tid == 0) {
      log_so3(sR, rvec);
      for (int k = 0; k < 3; ++k) tvec[k] = st[k];
    }
    return;
  }
  const double* Xb = w.Xf + 3 * (size_t)fr.base;
  const double* xb = w.xf + 2 * (size_t)fr.base;
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  for (int iter = 0; iter < 60; ++iter) {
    // normal equations at (sR, st)
    double acc[28];
    for (int k = 0; k < 28; ++k) acc[k] = 0.0;
    for (int q = tid; q < ni; q += REFINE_THREADS) {
      const int i = inliers[q];
      const double X = Xb[3 * i], Y = Xb[3 * i + 1], Z = Xb[3 * i + 2];
      const double rx = sR[0] * X + sR[1] * Y + sR[2] * Z, ry = sR[3] * X + sR[4] * Y + sR[5] * Z,
                   rz = sR[6] * X + sR[7] * Y + sR[8] * Z;
      const double xc = rx + st[0], yc = ry + st[1], zc = rz + st[2];
      const double iz = 1.0 / zc;
      const double eu = fx * xc * iz + cx - xb[2 * i], ev = fy * yc * iz + cy - xb[2 * i + 1];
      // d(u,v)/dXc
      const double a0 = fx * iz, a2 = -fx * xc * iz * iz, b1 = fy * iz, b2 = -fy * yc * iz * iz;
      // dXc/dw = -[RX]x ; dXc/dt = I
      double Ju[6], Jv[6];
      Ju[0] = a2 * ry;            Ju[1] = a0 * rz - a2 * rx;  Ju[2] = -a0 * ry;
      Jv[0] = -b1 * rz + b2 * ry; Jv[1] = -b2 * rx;           Jv[2] = b1 * rx;
      Ju[3] = a0; Ju[4] = 0.0; Ju[5] = a2;
      Jv[3] = 0.0; Jv[4] = b1; Jv[5] = b2;
      int e = 0;
      for (int r = 0; r < 6; ++r)
        for (int c = 0; c <= r; ++c) acc[e++] += Ju[r] * Ju[c] + Jv[r] * Jv[c];
      for (int r = 0; r < 6; ++r) acc[21 + r] += Ju[r] * eu + Jv[r] * ev;
      acc[27] += eu * eu + ev * ev;
    }
    block_reduce(acc, 28, red);
    if (tid == 0) {
      double H[36], g[6], d[6];
      int e = 0;
      for (int r = 0; r < 6; ++r)
        for (int c = 0; c <= r; ++c) { H[6 * r + c] = acc[e]; H[6 * c + r] = acc[e]; ++e; }
      for (int r = 0; r < 6; ++r) g[r] = acc[21 + r];
      s_cost = acc[27];
      if (!chol_solve6(H, g, s_lambda, d)) { s_lambda *= 10.0; for (int k = 0; k < 6; ++k) d[k] = 0.0; }
      double dR[9];
      exp_so3(d, dR);
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) cR[3 * i + j] = dR[3 * i] * sR[j] + dR[3 * i + 1] * sR[3 + j] + dR[3 * i + 2] * sR[6 + j];
      // t' = exp(w) t_rot-part is NOT applied to t: X_c' = exp(w) R X + t + d
      for (int i = 0; i < 3; ++i) ct[i] = st[i] + d[3 + i];
      const double dn = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2] + d[3] * d[3] + d[4] * d[4] + d[5] * d[5]);
      const double pn = sqrt(st[0] * st[0] + st[1] * st[1] + st[2] * st[2]) + 1.0;
      if (dn < 1e-13 * pn) s_state = 1;
    }
    __syncthreads();
    if (s_state) break;
    // cost at the candidate
    double c1[1] = {0.0};
    for (int q = tid; q < ni; q += REFINE_THREADS) {
      const int i = inliers[q];
      const double X = Xb[3 * i], Y = Xb[3 * i + 1], Z = Xb[3 * i + 2];
      const double xc = cR[0] * X + cR[1] * Y + cR[2] * Z + ct[0], yc = cR[3] * X + cR[4] * Y + cR[5] * Z + ct[1],
                   zc = cR[6] * X + cR[7] * Y + cR[8] * Z + ct[2];
      const double eu = fx * xc / zc + cx - xb[2 * i], ev = fy * yc / zc + cy - xb[2 * i + 1];
      c1[0] += eu * eu + ev * ev;
    }
    block_reduce(c1, 1, red);
    if (tid == 0) {
      if (c1[0] <= s_cost) {
        for (int k = 0; k < 9; ++k) sR[k] = cR[k];
        for (int k = 0; k < 3; ++k) st[k] = ct[k];
        s_lambda = fmax(s_lambda * 0.1, 1e-15);
      } else {
        s_lambda *= 10.0;
        if (s_lambda > 1e15) s_state = 1;
      }
    }
    __syncthreads();
    if (s_state) break;
  }
  if (tid == 0) {
    log_so3(sR, rvec);
    for (int k = 0; k < 3; ++k) tvec[k] = st[k];
  }
}

}  // namespace vo

using namespace vo;

static int pnp_launch(const double* X, const double* x, const int32_t* off, int n_single,
                      long n_total, int n_frames, const double* K, int iters, double reproj_px,
                      double conf, double* rvec, double* tvec, int32_t* inliers, int32_t* n_inl,
                      int32_t* ok, const PnpWs& w, cudaStream_t st) {
  const float thr2 = (float)(reproj_px * reproj_px);
  if (n_total > 0)
    pnp_prepare_kernel<<<(unsigned)((n_total + 255) / 256), 256, 0, st>>>(X, x, n_total, w);
  pnp_sample_kernel<<<(n_frames + 63) / 64, 64, 0, st>>>(off, n_single, n_frames, w);
  pnp_solve_kernel<<<dim3((iters + PNP_SOLVE_WARPS - 1) / PNP_SOLVE_WARPS, n_frames), PNP_SOLVE_WARPS * 32, 0, st>>>(
      K, off, n_single, w);
  pnp_score_kernel<<<dim3((iters + 7) / 8, n_frames), 256, 0, st>>>(K, off, n_single, thr2, w);
  pnp_select_kernel<<<n_frames, 1024, 0, st>>>(K, off, n_single, conf, thr2, w, inliers, n_inl, ok);
  pnp_refine_kernel<<<n_frames, REFINE_THREADS, 0, st>>>(K, off, n_single, w, inliers, rvec, tvec);
  return check_launch("vo_pnp_ransac");
}

extern "C" size_t vo_pnp_workspace_bytes(int n, int iters) {
  if (n <= 0 || iters <= 0) return 0;
  return pnp_carve(nullptr, n, 1, iters).bytes;
}

extern "C" size_t vo_pnp_batch_workspace_bytes(long n_total, int n_frames, int iters) {
  if (n_total <= 0 || n_frames <= 0 || iters <= 0) return 0;
  return pnp_carve(nullptr, n_total, n_frames, iters).bytes;
}

extern "C" int vo_pnp_ransac(const double* X, const double* x, int n, const double* K, int iters,
                             double reproj_px, double conf, double* rvec, double* tvec,
                             int32_t* inliers, int32_t* n_inl, int32_t* ok, void* ws, size_t ws_bytes,
                             void* stream) {
  VO_REQUIRE(X && x && K && rvec && tvec && inliers && n_inl && ok && ws, VO_ERR_INVALID_ARG,
             "vo_pnp_ransac: null pointer");
  VO_REQUIRE(n >= 4, VO_ERR_TOO_FEW_POINTS, "vo_pnp_ransac: needs >= 4 points (got %d)", n);
  VO_REQUIRE(iters > 0 && iters <= PNP_MAX_ITERS, VO_ERR_INVALID_ARG, "vo_pnp_ransac: iters must be in 1..1024");
  PnpWs w = pnp_carve(ws, n, 1, iters);
  VO_REQUIRE(ws_bytes >= w.bytes, VO_ERR_WORKSPACE, "vo_pnp_ransac: workspace too small");
  return pnp_launch(X, x, nullptr, n, n, 1, K, iters, reproj_px, conf, rvec, tvec, inliers, n_inl, ok, w,
                    (cudaStream_t)stream);
}

extern "C" int vo_pnp_ransac_batch(const double* X, const double* x, const int32_t* offsets,
                                   long n_total, int n_frames, const double* K, int iters,
                                   double reproj_px, double conf, double* rvec, double* tvec,
                                   int32_t* inliers, int32_t* n_inl, int32_t* ok, void* ws,
                                   size_t ws_bytes, void* stream) {
  VO_REQUIRE(X && x && offsets && K && rvec && tvec && inliers && n_inl && ok && ws, VO_ERR_INVALID_ARG,
             "vo_pnp_ransac_batch: null pointer");
  VO_REQUIRE(n_total >= 0 && n_frames >= 0 && n_frames <= 65535, VO_ERR_INVALID_ARG,
             "vo_pnp_ransac_batch: bad sizes");
  VO_REQUIRE(iters > 0 && iters <= PNP_MAX_ITERS, VO_ERR_INVALID_ARG, "vo_pnp_ransac_batch: iters must be in 1..1024");
  if (n_frames == 0) return VO_OK;
  PnpWs w = pnp_carve(ws, n_total > 0 ? n_total : 1, n_frames, iters);
  VO_REQUIRE(ws_bytes >= w.bytes, VO_ERR_WORKSPACE, "vo_pnp_ransac_batch: workspace too small");
  return pnp_launch(X, x, offsets, 0, n_total, n_frames, K, iters, reproj_px, conf, rvec, tvec, inliers,
                    n_inl, ok, w, (cudaStream_t)stream);
}
