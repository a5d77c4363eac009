// fivept.cuh — K3: five-point essential-matrix minimal solver, ONE WARP PER SAMPLE.
//
// Replaces EMEstimatorCallback::runKernel of OpenCV's five-point.cpp, reached from the reference
// through cv2.findEssentialMat (src/image_pair.py:280-286).  Same algebra (Nister 2004:
// 4-dim null space -> ten cubic constraints -> Gauss-Jordan -> 3x3 polynomial matrix B(z) ->
// degree-10 polynomial -> real roots -> back-substitution) but B200-shaped numerics:
//   * null space by Householder QR of Q^T (orthonormal, no 9x9 SVD);
//   * all ten constraint rows built in parallel by lanes 0..9 from table-driven polynomial
//     products, elimination spread over the 32 lanes in shared memory;
//   * the ten complex roots found simultaneously by Aberth-Ehrlich iteration, one root per lane
//     (OpenCV: 300 sequential Durand-Kerner sweeps), `|Im| <= 1e-10` kept as in OpenCV;
//   * every root is polished by Gauss-Newton on the ten constraints, so solutions satisfy
//     2EE'E - tr(EE')E = 0 to ~1e-12 even for ill-conditioned samples.
// Output: up to 10 matrices, Frobenius norm 1, ordered by the hidden variable z.
#pragma once
#include "common.cuh"

namespace vo {

struct FiveptSmem {
  double M[9][5];      // Q^T (9x5), overwritten by the QR
  double V[5][9];      // Householder vectors
  double basis[4][9];  // X, Y, Z, W  (row-major 3x3 each)
  double T2[9][10];    // degree-2 temporaries (EE^T upper triangle, cofactors of E)
  double P2[12][10];   // Lambda = EE^T - tr/2 I (9 entries) + 3 cofactors
  double A[10][21];    // constraint matrix, 20 monomials (+1 pad)
  double B[9][5];      // B(z) entries, ascending powers of z
  double poly[11];     // det B(z), ascending
  double2 zs[10];      // current root estimates (Aberth), broadcast to all lanes
};

// index of the degree-2 monomial v_a*v_b among the 10 of (x,y,z,w); compile-time when unrolled
__host__ __device__ constexpr int idx2(int a, int b) {
  return a <= b ? (a == 0 ? b : a == 1 ? 3 + b : a == 2 ? 5 + b : 9)
                : (b == 0 ? a : b == 1 ? 3 + a : b == 2 ? 5 + a : 9);
}
static_assert(idx2(0, 0) == 0 && idx2(0, 3) == 3 && idx2(1, 1) == 4 && idx2(1, 3) == 6 &&
              idx2(2, 2) == 7 && idx2(3, 2) == 8 && idx2(3, 3) == 9, "idx2 table");
// (degree-2 monomial m) x (variable c) -> column of the degree-3 monomial in elimination order
// [x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1]
__constant__ int c_COL3[10][4] = {{0, 2, 4, 5},     {2, 3, 8, 9},     {4, 8, 10, 11},
                                  {5, 9, 11, 12},   {3, 1, 6, 7},     {8, 6, 13, 14},
                                  {9, 7, 14, 15},   {10, 13, 16, 17}, {11, 14, 17, 18},
                                  {12, 15, 18, 19}};

__host__ __device__ constexpr int col3(int m, int c) {
  constexpr int T[10][4] = {{0, 2, 4, 5},     {2, 3, 8, 9},     {4, 8, 10, 11},
                            {5, 9, 11, 12},   {3, 1, 6, 7},     {8, 6, 13, 14},
                            {9, 7, 14, 15},   {10, 13, 16, 17}, {11, 14, 17, 18},
                            {12, 15, 18, 19}};
  return T[m][c];
}

struct cplx {
  double re, im;
};
__device__ __forceinline__ cplx cmul(cplx a, cplx b) {
  return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re};
}
__device__ __forceinline__ cplx cdiv(cplx a, cplx b) {
  const double d = b.re * b.re + b.im * b.im;
  return {(a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d};
}

// fast fp64 reciprocal (MUFU.RCP64H, ~2^-23 relative) and the same refined by one Newton step
// (~2^-46).  Only used inside the Aberth iteration, whose fixed points (p(z) = 0) do not depend on
// how accurately the correction is evaluated; every root is re-polished in full fp64 afterwards.
__device__ __forceinline__ double rcp_fast(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r;
}
__device__ __forceinline__ double rcp_mid(double x) {
  const double r = rcp_fast(x);
  return r * (2.0 - x * r);
}
__device__ __forceinline__ cplx cdiv_mid(cplx a, cplx b) {
  const double inv = rcp_mid(b.re * b.re + b.im * b.im);
  return {(a.re * b.re + a.im * b.im) * inv, (a.im * b.re - a.re * b.im) * inv};
}

// degree-1 polynomial of E[r][c] in (x,y,z,w): coefficient v is basis[v][3r+c]
#define E1(S, v, r, c) ((S).basis[(v)][3 * (r) + (c)])

// p5: the sample, 5 x (x1,y1,x2,y2) normalised coordinates (any address space).
// Eout: [10][9] doubles (global). Returns the number of models (uniform across the warp).
__device__ int fivept_solve_warp(const double* __restrict__ p5, FiveptSmem& S,
                                 double* __restrict__ Eout) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;

  // ---- 1. epipolar rows  x2' E x1 = 0  ->  M = Q^T
  if (lane < 5) {
    const double x1 = p5[4 * lane], y1 = p5[4 * lane + 1];
    const double x2 = p5[4 * lane + 2], y2 = p5[4 * lane + 3];
    S.M[0][lane] = x2 * x1; S.M[1][lane] = x2 * y1; S.M[2][lane] = x2;
    S.M[3][lane] = y2 * x1; S.M[4][lane] = y2 * y1; S.M[5][lane] = y2;
    S.M[6][lane] = x1;      S.M[7][lane] = y1;      S.M[8][lane] = 1.0;
  }
  __syncwarp();

  // ---- 2. Householder QR of M (9x5); the last 4 columns of the orthogonal factor span null(Q)
  for (int k = 0; k < 5; ++k) {
    double nrm2 = 0.0;
    for (int r = k; r < 9; ++r) nrm2 += S.M[r][k] * S.M[r][k];
    const double x0 = S.M[k][k];
    const double alpha = (x0 >= 0.0) ? -sqrt(nrm2) : sqrt(nrm2);
    const double v0 = x0 - alpha;
    const double vtv = nrm2 - x0 * x0 + v0 * v0;
    __syncwarp();
    if (lane < 9) S.V[k][lane] = (lane < k) ? 0.0 : (lane == k ? v0 : S.M[lane][k]);
    __syncwarp();
    if (lane > k && lane < 5 && vtv > 0.0) {
      double dot = 0.0;
      for (int r = k; r < 9; ++r) dot += S.V[k][r] * S.M[r][lane];
      const double f = 2.0 * dot / vtv;
      for (int r = k; r < 9; ++r) S.M[r][lane] -= f * S.V[k][r];
    }
    __syncwarp();
  }
  if (lane < 4) {
    double e[9];
#pragma unroll
    for (int r = 0; r < 9; ++r) e[r] = (r == 5 + lane) ? 1.0 : 0.0;
    for (int k = 4; k >= 0; --k) {
      double dot = 0.0, vtv = 0.0;
#pragma unroll
      for (int r = 0; r < 9; ++r) {
        dot += S.V[k][r] * e[r];
        vtv += S.V[k][r] * S.V[k][r];
      }
      const double f = vtv > 0.0 ? 2.0 * dot / vtv : 0.0;
#pragma unroll
      for (int r = 0; r < 9; ++r) e[r] -= f * S.V[k][r];
    }
#pragma unroll
    for (int r = 0; r < 9; ++r) S.basis[lane][r] = e[r];
  }
  __syncwarp();

  // ---- 3a. degree-2 pieces: EE^T (6 unique) on lanes 0..5, cofactors of row 0 on lanes 6..8
  if (lane < 9) {
    double acc[10];
#pragma unroll
    for (int m = 0; m < 10; ++m) acc[m] = 0.0;
    if (lane < 6) {
      const int i = (lane < 3) ? 0 : (lane < 5 ? 1 : 2);
      const int j = (lane < 3) ? lane : (lane < 5 ? lane - 2 : 2);
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[idx2(a, b)] += E1(S, a, i, k) * E1(S, b, j, k);
    } else {
      // cof0 = E11E22 - E12E21 ; cof1 = E12E20 - E10E22 ; cof2 = E10E21 - E11E20
      const int j = lane - 6;
      const int c1 = (j + 1) % 3, c2 = (j + 2) % 3;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
          acc[idx2(a, b)] +=
              E1(S, a, 1, c1) * E1(S, b, 2, c2) - E1(S, a, 1, c2) * E1(S, b, 2, c1);
    }
#pragma unroll
    for (int m = 0; m < 10; ++m) S.T2[lane][m] = acc[m];
  }
  __syncwarp();
  // Lambda[i][k] (9 entries) and cofactors -> P2
  for (int e = lane; e < 120; e += 32) {
    const int p = e / 10, m = e % 10;
    double v;
    if (p < 9) {
      const int i = p / 3, k = p % 3;
      const int lo = i < k ? i : k, hi = i < k ? k : i;
      const int t = (lo == 0) ? hi : (lo == 1 ? 2 + hi : 5);  // (0,0)(0,1)(0,2)(1,1)(1,2)(2,2)
      v = S.T2[t][m];
      if (i == k) v -= 0.5 * (S.T2[0][m] + S.T2[3][m] + S.T2[5][m]);
    } else {
      v = S.T2[6 + (p - 9)][m];
    }
    S.P2[p][m] = v;
  }
  __syncwarp();

  // ---- 3b. ten cubic rows: rows 0..8 = (Lambda E)[i][j], row 9 = det E.  The row lives in
  // registers (column indices are compile-time after unrolling) and is stored once.
  if (lane < 10) {
    const int i = lane / 3, j = lane % 3;
    double row[20];
#pragma unroll
    for (int c = 0; c < 20; ++c) row[c] = 0.0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double* d2 = (lane < 9) ? S.P2[3 * i + k] : S.P2[9 + k];
      double d1[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) d1[c] = (lane < 9) ? E1(S, c, k, j) : E1(S, c, 0, k);
#pragma unroll
      for (int m = 0; m < 10; ++m) {
        const double a = d2[m];
#pragma unroll
        for (int c = 0; c < 4; ++c) row[col3(m, c)] += a * d1[c];
      }
    }
#pragma unroll
    for (int c = 0; c < 20; ++c) S.A[lane][c] = row[c];
  }
  __syncwarp();

  // ---- 4. Gauss-Jordan on the first 10 columns, partial pivoting, 32 lanes on 10x20
  bool singular = false;
  for (int k = 0; k < 10; ++k) {
    double best = (lane >= k && lane < 10) ? fabs(S.A[lane][k]) : -1.0;
    int brow = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(FULL, best, o);
      const int orow = __shfl_xor_sync(FULL, brow, o);
      if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
    }
    if (!(best > 1e-300) || !isfinite(best)) { singular = true; break; }
    if (brow != k && lane < 20) {
      const double t = S.A[k][lane]; S.A[k][lane] = S.A[brow][lane]; S.A[brow][lane] = t;
    }
    __syncwarp();
    const double inv = 1.0 / S.A[k][k];
    __syncwarp();
    if (lane < 20) S.A[k][lane] *= inv;
    __syncwarp();
    // eliminate column k from the other 9 rows: 9 x 20 entries over 32 lanes
    double fac[6];
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int e = lane + 32 * u;
      int r = e / 20; r += (r >= k);
      fac[u] = (e < 180) ? S.A[r][k] : 0.0;
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int e = lane + 32 * u;
      if (e < 180) {
        int r = e / 20; r += (r >= k);
        const int c = e % 20;
        S.A[r][c] -= fac[u] * S.A[k][c];
      }
    }
    __syncwarp();
  }
  if (singular) return 0;

  // ---- 5. B(z): rows <k>=<e>-z<f>, <l>=<g>-z<h>, <m>=<i>-z<j> of the reduced system.
  // reduced columns 10..19 = [xz2 xz x | yz2 yz y | z3 z2 z 1]
  if (lane < 9) {
    const int bi = lane / 3, c = lane % 3;
    const double* p = &S.A[4 + 2 * bi][10];
    const double* q = &S.A[5 + 2 * bi][10];
    double* o = S.B[lane];
    if (c < 2) {
      const int s = 3 * c;
      o[0] = p[s + 2];
      o[1] = p[s + 1] - q[s + 2];
      o[2] = p[s] - q[s + 1];
      o[3] = -q[s];
      o[4] = 0.0;
    } else {
      o[0] = p[9];
      o[1] = p[8] - q[9];
      o[2] = p[7] - q[8];
      o[3] = p[6] - q[7];
      o[4] = -q[6];
    }
  }
  __syncwarp();
  // det B(z): coefficient `lane` of the degree-10 polynomial (triple convolution over the 6 perms)
  if (lane < 11) {
    const int perm[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {0, 2, 1}, {1, 0, 2}, {2, 1, 0}};
    double acc = 0.0;
    for (int pi = 0; pi < 6; ++pi) {
      const double* b0 = S.B[perm[pi][0]];
      const double* b1 = S.B[3 + perm[pi][1]];
      const double* b2 = S.B[6 + perm[pi][2]];
      double t = 0.0;
      for (int a = 0; a < 5; ++a)
        for (int b = 0; b < 5; ++b) {
          const int c = lane - a - b;
          if (c >= 0 && c < 5) t += b0[a] * b1[b] * b2[c];
        }
      acc += (pi < 3) ? t : -t;
    }
    S.poly[lane] = acc;
  }
  __syncwarp();

  // ---- 6. all ten complex roots at once: Aberth-Ehrlich, lane l owns root l % 10
  const double lead = S.poly[10];
  if (!(fabs(lead) > 0.0) || !isfinite(lead)) return 0;
  double mc[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) mc[k] = S.poly[k] / lead;
  double r0 = exp(log(fabs(mc[0])) * 0.1);
  if (!isfinite(r0) || r0 == 0.0) r0 = 1.0;
  const int slot = lane % 10;
  cplx z;
  {
    double sn, cs;
    sincos(0.62831853071795864769 * slot + 0.4, &sn, &cs);
    z = {r0 * cs, r0 * sn};
  }
  int settled = 0;
  for (int it = 0; it < 40 && settled < 2; ++it) {
    cplx p = {1.0, 0.0}, dp = {0.0, 0.0};
#pragma unroll
    for (int k = 9; k >= 0; --k) {
      dp = cmul(dp, z);
      dp.re += p.re; dp.im += p.im;
      p = cmul(p, z);
      p.re += mc[k];
    }
    const cplx nt = cdiv_mid(p, dp);
    __syncwarp();
    if (lane < 10) S.zs[lane] = make_double2(z.re, z.im);
    __syncwarp();
    cplx s = {0.0, 0.0};
#pragma unroll
    for (int j = 0; j < 10; ++j) {
      const double2 o = S.zs[j];
      const double dre = z.re - o.x, dim = z.im - o.y;
      // own slot: d = 0 -> contributes nothing (inv is forced to 0)
      const double dd = dre * dre + dim * dim;
      const double inv = (j != slot) ? rcp_fast(dd) : 0.0;
      s.re += dre * inv;
      s.im -= dim * inv;
    }
    cplx den = cmul(nt, s);
    den.re = 1.0 - den.re; den.im = -den.im;
    const cplx w = cdiv_mid(nt, den);
    const bool fin = isfinite(w.re) && isfinite(w.im);
    if (fin) { z.re -= w.re; z.im -= w.im; }
    const bool ok = !fin || (w.re * w.re + w.im * w.im <= 1e-24 * (z.re * z.re + z.im * z.im));
    settled = __all_sync(FULL, ok) ? settled + 1 : 0;
  }

  // ---- 7. real roots -> (x, y, z) -> E ; one root per lane (lanes 0..9)
  bool valid = lane < 10 && fabs(z.im) <= 1e-10 && isfinite(z.re);
  double zz = z.re, xx = 0.0, yy = 0.0;
  double Em[9];
  if (valid) {
    for (int itn = 0; itn < 2; ++itn) {  // Newton polish on the real polynomial
      double p = S.poly[10], dp = 0.0;
      for (int k = 9; k >= 0; --k) { dp = dp * zz + p; p = p * zz + S.poly[k]; }
      const double st = p / dp;
      if (isfinite(st)) zz -= st;
    }
    double b[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) {
      const double* c = S.B[e];
      b[e] = (((c[4] * zz + c[3]) * zz + c[2]) * zz + c[1]) * zz + c[0];
    }
    // null vector of the 3x3 B(z): best-conditioned cross product of two rows
    double best = -1.0, v0 = 0, v1 = 0, v2 = 0;
#pragma unroll
    for (int pr = 0; pr < 3; ++pr) {
      const int r1 = (pr == 2) ? 1 : 0, r2 = (pr == 0) ? 1 : 2;
      const double c0 = b[3 * r1 + 1] * b[3 * r2 + 2] - b[3 * r1 + 2] * b[3 * r2 + 1];
      const double c1 = b[3 * r1 + 2] * b[3 * r2] - b[3 * r1] * b[3 * r2 + 2];
      const double c2 = b[3 * r1] * b[3 * r2 + 1] - b[3 * r1 + 1] * b[3 * r2];
      const double n2 = c0 * c0 + c1 * c1 + c2 * c2;
      if (n2 > best) { best = n2; v0 = c0; v1 = c1; v2 = c2; }
    }
    const double inv = rsqrt(best);
    v0 *= inv; v1 *= inv; v2 *= inv;
    if (!(fabs(v2) >= 1e-10) || !isfinite(v2)) valid = false;
    else { xx = v0 / v2; yy = v1 / v2; }
  }
  if (valid) {
    // Gauss-Newton polish of (x,y,z) on the ten reduced constraints  m_r + sum_c Rm[r][c] m_{10+c}
    for (int itn = 0; itn < 2; ++itn) {
      const double x = xx, y = yy, zv = zz;
      const double mlo[10] = {x * x * x, y * y * y, x * x * y, x * y * y, x * x * zv,
                              x * x,     y * y * zv, y * y,     x * y * zv, x * y};
      const double dxl[10] = {3 * x * x, 0, 2 * x * y, y * y, 2 * x * zv, 2 * x, 0, 0, y * zv, y};
      const double dyl[10] = {0, 3 * y * y, x * x, 2 * x * y, 0, 0, 2 * y * zv, 2 * y, x * zv, x};
      const double dzl[10] = {0, 0, 0, 0, x * x, 0, y * y, 0, x * y, 0};
      const double mhi[10] = {x * zv * zv, x * zv, x, y * zv * zv, y * zv, y, zv * zv * zv,
                              zv * zv,     zv,     1.0};
      const double dxh[10] = {zv * zv, zv, 1, 0, 0, 0, 0, 0, 0, 0};
      const double dyh[10] = {0, 0, 0, zv * zv, zv, 1, 0, 0, 0, 0};
      const double dzh[10] = {2 * x * zv, x, 0, 2 * y * zv, y, 0, 3 * zv * zv, 2 * zv, 1, 0};
      double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0, g0 = 0, g1 = 0, g2 = 0;
#pragma unroll
      for (int r = 0; r < 10; ++r) {
        double f = mlo[r], jx = dxl[r], jy = dyl[r], jz = dzl[r];
#pragma unroll
        for (int c = 0; c < 10; ++c) {
          const double a = S.A[r][10 + c];
          f += a * mhi[c]; jx += a * dxh[c]; jy += a * dyh[c]; jz += a * dzh[c];
        }
        a00 += jx * jx; a01 += jx * jy; a02 += jx * jz; a11 += jy * jy; a12 += jy * jz;
        a22 += jz * jz; g0 += jx * f; g1 += jy * f; g2 += jz * f;
      }
      const double c00 = a11 * a22 - a12 * a12, c01 = a02 * a12 - a01 * a22,
                   c02 = a01 * a12 - a02 * a11;
      const double det = a00 * c00 + a01 * c01 + a02 * c02;
      const double c11 = a00 * a22 - a02 * a02, c12 = a01 * a02 - a00 * a12,
                   c22 = a00 * a11 - a01 * a01;
      const double sx = -(c00 * g0 + c01 * g1 + c02 * g2) / det;
      const double sy = -(c01 * g0 + c11 * g1 + c12 * g2) / det;
      const double sz = -(c02 * g0 + c12 * g1 + c22 * g2) / det;
      if (isfinite(sx) && isfinite(sy) && isfinite(sz)) { xx += sx; yy += sy; zz += sz; }
    }
    double n2 = 0.0;
#pragma unroll
    for (int e = 0; e < 9; ++e) {
      Em[e] = xx * S.basis[0][e] + yy * S.basis[1][e] + zz * S.basis[2][e] + S.basis[3][e];
      n2 += Em[e] * Em[e];
    }
    const double inv = 1.0 / sqrt(n2);
    if (!isfinite(inv)) valid = false;
#pragma unroll
    for (int e = 0; e < 9; ++e) Em[e] *= inv;
  }
  // order by z (ties by lane), compact
  const unsigned vmask = __ballot_sync(FULL, valid);
  int rank = 0;
#pragma unroll
  for (int j = 0; j < 10; ++j) {
    const double oz = __shfl_sync(FULL, zz, j);
    if (((vmask >> j) & 1u) && (oz < zz || (oz == zz && j < lane))) ++rank;
  }
  if (valid) {
#pragma unroll
    for (int e = 0; e < 9; ++e) Eout[9 * rank + e] = Em[e];
  }
  return __popc(vmask);
}

}  // namespace vo
