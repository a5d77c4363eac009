// linalg.cuh — tiny fp64 dense algebra kept entirely in registers (fully unrolled).
// One-sided (Hestenes) Jacobi SVD, the same family OpenCV's cv::SVD uses for small matrices
// (core/src/lapack.cpp JacobiSVDImpl_), so null vectors agree with cv2 to rounding.
#pragma once
#include "common.cuh"

namespace vo {

__device__ __forceinline__ double rcp_nr(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r * (2.0 - x * r);
}
__device__ __forceinline__ double sqrt_nr(double x) {   // x >= 1 here
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  y = y * (1.5 - 0.5 * x * y * y);
  return x * y;
}

// A (N x N, row-major, modified in place to U*diag(s)); V accumulates the right singular vectors
// as columns. Column norms of the result are the singular values (unsorted).
template <int N>
__device__ __forceinline__ void jacobi_svd(double (&A)[N][N], double (&V)[N][N]) {
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    bool changed = false;
#pragma unroll
    for (int p = 0; p < N - 1; ++p) {
#pragma unroll
      for (int q = p + 1; q < N; ++q) {
        double alpha = 0.0, beta = 0.0, gamma = 0.0;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          alpha += A[k][p] * A[k][p];
          beta += A[k][q] * A[k][q];
          gamma += A[k][p] * A[k][q];
        }
        // |gamma| <= eps * sqrt(alpha * beta), without the square root
        if (gamma * gamma <= (2.2e-15 * 2.2e-15) * (alpha * beta)) continue;
        changed = true;
        // The rotation ANGLE only steers convergence, so zeta and t use fast reciprocals /
        // square roots (MUFU seed + one Newton step, ~2^-45); c = 1/sqrt(1+t^2) is computed to
        // full precision so that (c, s) stays an exact rotation and V stays orthonormal.
        const double zeta = (beta - alpha) * 0.5 * rcp_nr(gamma);
        const double az = fabs(zeta);
        const double t = copysign(rcp_nr(az + sqrt_nr(fma(zeta, zeta, 1.0))), zeta);
        const double c = rsqrt(fma(t, t, 1.0));
        const double s = c * t;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const double ap = A[k][p], aq = A[k][q];
          A[k][p] = c * ap - s * aq;
          A[k][q] = s * ap + c * aq;
          const double vp = V[k][p], vq = V[k][q];
          V[k][p] = c * vp - s * vq;
          V[k][q] = s * vp + c * vq;
        }
      }
    }
    if (!changed) break;
  }
}

// Right singular vector of the smallest singular value of the 4x4 DLT matrix (unit norm).
__device__ __forceinline__ void null_vector4(double (&A)[4][4], double (&x)[4]) {
  double V[4][4];
  jacobi_svd<4>(A, V);
  double best = 1e300;
  int bj = 3;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double n2 = A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j] + A[3][j] * A[3][j];
    if (n2 < best) { best = n2; bj = j; }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
    x[k] = (bj == 0) ? V[k][0] : (bj == 1) ? V[k][1] : (bj == 2) ? V[k][2] : V[k][3];
}

// full-precision reciprocal: MUFU seed + two Newton steps
__device__ __forceinline__ double rcp_full(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = r * (2.0 - x * r);
  return fma(r, fma(-x, r, 1.0), r);
}

// The same vector ~8x cheaper: the smallest right singular vector of A is the eigenvector of the
// smallest eigenvalue of B = A^T A, and for a DLT matrix that eigenvalue is separated from the
// next by the square of (noise / parallax), so inverse iteration on an LDL^T factorisation of B
// converges in 2-4 steps (error ratio lambda_4 / lambda_3 per step).  Accuracy: the factorisation
// is backward stable for a symmetric positive definite B, the fixed point is the eigenvector of
// B + dB with |dB| <= c eps |B|, i.e. within eps sigma_1^2 / (sigma_3^2 - sigma_4^2) of the true
// one (~1e-13 for the geometries of this path; Jacobi on A itself: ~1e-15).  Points that have not
// settled to 1e-14 after 16 steps (sigma_4 ~ sigma_3: no parallax at all) return false and take
// the Jacobi route, whose answer they then share with OpenCV to rounding.
__device__ __forceinline__ bool null_vector4_invit(const double (&A)[4][4], double (&x)[4]) {
  double b[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = i; j < 4; ++j)
      b[i][j] = fma(A[0][i], A[0][j], fma(A[1][i], A[1][j], fma(A[2][i], A[2][j], A[3][i] * A[3][j])));
  const double tiny = 1e-300 + 1e-34 * (b[0][0] + b[1][1] + b[2][2] + b[3][3]);
  // B = L D L^T
  const double d0 = fmax(b[0][0], tiny), i0 = rcp_full(d0);
  const double l10 = b[0][1] * i0, l20 = b[0][2] * i0, l30 = b[0][3] * i0;
  const double d1 = fmax(fma(-l10, b[0][1], b[1][1]), tiny), i1 = rcp_full(d1);
  const double t21 = fma(-l20, b[0][1], b[1][2]), t31 = fma(-l30, b[0][1], b[1][3]);
  const double l21 = t21 * i1, l31 = t31 * i1;
  const double d2 = fmax(fma(-l21, t21, fma(-l20, b[0][2], b[2][2])), tiny), i2 = rcp_full(d2);
  const double t32 = fma(-l31, t21, fma(-l30, b[0][2], b[2][3]));
  const double l32 = t32 * i2;
  const double d3r = fma(-l32, t32, fma(-l31, t31, fma(-l30, b[0][3], b[3][3])));
  const double d3 = fmax(d3r, tiny), i3 = rcp_full(d3);
  double v0 = 0.5, v1 = 0.5, v2 = 0.5, v3 = 0.5;
  bool settled = false;
  for (int it = 0; it < 16; ++it) {
    // solve L D L^T y = v
    const double z0 = v0;
    const double z1 = fma(-l10, z0, v1);
    const double z2 = fma(-l21, z1, fma(-l20, z0, v2));
    const double z3 = fma(-l32, z2, fma(-l31, z1, fma(-l30, z0, v3)));
    double y3 = z3 * i3;
    double y2 = fma(-l32, y3, z2 * i2);
    double y1 = fma(-l31, y3, fma(-l21, y2, z1 * i1));
    double y0 = fma(-l30, y3, fma(-l20, y2, fma(-l10, y1, z0 * i0)));
    const double n2 = fma(y0, y0, fma(y1, y1, fma(y2, y2, y3 * y3)));
    const double dot = fma(y0, v0, fma(y1, v1, fma(y2, v2, y3 * v3)));
    const double inv = copysign(rsqrt(n2), dot);    // keep the sign of the previous iterate
    y0 *= inv; y1 *= inv; y2 *= inv; y3 *= inv;
    const double e0 = y0 - v0, e1 = y1 - v1, e2 = y2 - v2, e3 = y3 - v3;
    const double dd = fma(e0, e0, fma(e1, e1, fma(e2, e2, e3 * e3)));
    v0 = y0; v1 = y1; v2 = y2; v3 = y3;
    if (settled) break;                 // one more step after the iterates agree to 1e-14
    if (dd < 1e-28) settled = true;
    if (!(n2 < 1e300)) break;           // overflow guard (exactly singular B): direction is exact
  }
  x[0] = v0; x[1] = v1; x[2] = v2; x[3] = v3;
  return settled;
}

// cv::triangulatePoints for one point: rows x*P[2]-P[0], y*P[2]-P[1] for both cameras.
__device__ __forceinline__ void dlt_point(const double* __restrict__ P1,
                                          const double* __restrict__ P2, double x1, double y1,
                                          double x2, double y2, double (&X)[4]) {
  double A[4][4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    A[0][k] = x1 * P1[8 + k] - P1[k];
    A[1][k] = y1 * P1[8 + k] - P1[4 + k];
    A[2][k] = x2 * P2[8 + k] - P2[k];
    A[3][k] = y2 * P2[8 + k] - P2[4 + k];
  }
  if (null_vector4_invit(A, X)) return;
  null_vector4(A, X);
}

__device__ __forceinline__ double det3(const double* M) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) +
         M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// decomposeEssentialMat (five-point.cpp): E = U diag(1,1,0) V^T, det(U),det(V^T) > 0,
// R1 = U W V^T, R2 = U W^T V^T, t = U[:,2].  Row-major outputs.
__device__ inline void decompose_essential(const double* __restrict__ E, double* R1, double* R2,
                                           double* t) {
  double A[3][3], V[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) A[i][j] = E[3 * i + j];
  jacobi_svd<3>(A, V);
  double s2[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) s2[j] = A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j];
  // order columns by descending singular value
  int o0 = 0, o1 = 1, o2 = 2;
  if (s2[o0] < s2[o1]) { int x = o0; o0 = o1; o1 = x; }
  if (s2[o1] < s2[o2]) { int x = o1; o1 = o2; o2 = x; }
  if (s2[o0] < s2[o1]) { int x = o0; o0 = o1; o1 = x; }
  double U[3][3], W[3][3];  // W here = V with ordered columns
  const double i0 = rsqrt(s2[o0]), i1 = rsqrt(s2[o1]);
  for (int k = 0; k < 3; ++k) {
    U[k][0] = A[k][o0] * i0;
    U[k][1] = A[k][o1] * i1;
    W[k][0] = V[k][o0];
    W[k][1] = V[k][o1];
    W[k][2] = V[k][o2];
  }
  // third left vector completes a right-handed basis (sigma_3 = 0 leaves it free): det(U) = +1
  U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
  U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
  U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
  double Vt[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Vt[3 * i + j] = W[j][i];
  if (det3(Vt) < 0)
    for (int e = 0; e < 9; ++e) Vt[e] = -Vt[e];
  // U*W = [-u2, u1, u3] ; U*W^T = [u2, -u1, u3]
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const double a = -U[i][1] * Vt[j] + U[i][0] * Vt[3 + j];
      const double b = U[i][2] * Vt[6 + j];
      R1[3 * i + j] = a + b;
      R2[3 * i + j] = -a + b;
    }
  t[0] = U[0][2]; t[1] = U[1][2]; t[2] = U[2][2];
}

}  // namespace vo
