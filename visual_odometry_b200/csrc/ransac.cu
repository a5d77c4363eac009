// ransac.cu — K2/K3 + control: essential-matrix 5-point RANSAC with OpenCV's exact semantics.
//
// Replaces cv2.findEssentialMat(p1, p2, K, FM_RANSAC, 0.99, 1) (reference
// src/image_pair.py:280-286; run twice per pair, src/visual_slam.py:296,165).
// OpenCV's loop (ptsetreg.cpp RANSACPointSetRegistrator::run) is sequential:
//     sample(cv::RNG(2^64-1)) -> <=10 models -> count inliers -> strict-greater best update ->
//     shrink niters (RANSACUpdateNumIters)
// but every iteration's sample depends only on the RNG stream and the point count, so the GPU
// replays the stream, solves and scores a whole CHUNK of iterations in parallel (one warp per
// 5-point solve, one warp per sample for Sampson scoring with coalesced 32-byte correspondence
// reads) and then applies the sequential best/adaptive-stop rule as a cheap ordered scan.
// Iterations at or beyond the current bound are skipped, exactly as OpenCV never runs them.
#include <float.h>

#include "fivept.cuh"

namespace vo {

// ------------------------------------------------------------------ workspace carving
struct RansacWs {
  double* ptsn;       // [P][n_cap][4] normalised correspondences
  double* models;     // [P][C][10][9]
  uint64_t* rng;      // [P]
  int32_t* niters;    // [P]
  int32_t* maxgood;   // [P]
  int32_t* done;      // [P]
  int32_t* samples;   // [P][C][5]
  int32_t* nmodels;   // [P][C]
  int32_t* counts;    // [P][C][10]
  int32_t* active;    // [2][P] pairs still iterating (ping-pong between rounds)
  int32_t* n_active;  // [max_rounds + 1]
  int max_rounds;
  size_t bytes;
};

static inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

static RansacWs carve(void* base, int P, int n_cap, int C, int max_iters) {
  RansacWs w;
  w.max_rounds = max_iters / C + 8;   // chunks of C, plus the short warm-up rounds
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += align_up(bytes);
    return p;
  };
  w.ptsn = (double*)take((size_t)P * n_cap * 4 * sizeof(double));
  w.models = (double*)take((size_t)P * C * 90 * sizeof(double));
  w.rng = (uint64_t*)take((size_t)P * sizeof(uint64_t));
  w.niters = (int32_t*)take((size_t)P * sizeof(int32_t));
  w.maxgood = (int32_t*)take((size_t)P * sizeof(int32_t));
  w.done = (int32_t*)take((size_t)P * sizeof(int32_t));
  w.samples = (int32_t*)take((size_t)P * C * 5 * sizeof(int32_t));
  w.nmodels = (int32_t*)take((size_t)P * C * sizeof(int32_t));
  w.counts = (int32_t*)take((size_t)P * C * 10 * sizeof(int32_t));
  w.active = (int32_t*)take((size_t)2 * P * sizeof(int32_t));
  w.n_active = (int32_t*)take((size_t)(w.max_rounds + 1) * sizeof(int32_t));
  w.bytes = off;
  return w;
}

constexpr int RANSAC_CHUNK0 = 16;    // first round (adaptive runs usually stop inside it)
constexpr int RANSAC_CHUNK_MAX = 256;

// Iterations solved + scored per round.  256 for big batches; for few pairs the cap grows (up to
// the whole budget) so that a fixed-hypothesis run (config C3: 4096 hypotheses, adaptive stop
// off) needs one or two rounds instead of sixteen latency-bound ones.  Bounded so that the
// per-round model store (720 B per iteration per pair) stays below ~760 MB.
static inline int chunk_cap(int max_iters, int n_pairs) {
  long cap = 1048576L / (n_pairs > 0 ? n_pairs : 1);
  if (cap < RANSAC_CHUNK_MAX) cap = RANSAC_CHUNK_MAX;
  if (cap > max_iters) cap = max_iters;
  return cap < 1 ? 1 : (int)cap;
}

__device__ __forceinline__ double4 ldg_d4(const double4* p) {
  const double2 a = __ldg(reinterpret_cast<const double2*>(p));
  const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  return make_double4(a.x, a.y, b.x, b.y);
}

// ------------------------------------------------------------------ cv::RNG replay
__device__ __forceinline__ uint32_t cvrng_next(uint64_t& s) {
  s = (uint64_t)(uint32_t)s * 4164903690ull + (uint32_t)(s >> 32);
  return (uint32_t)s;
}
// RANSACPointSetRegistrator::getSubset: 5 distinct indices, redraw on duplicate
// x % n for 32-bit x, n without a division (Lemire's fastmod: M = floor(2^64 / n) + 1, exact)
__device__ __forceinline__ uint32_t fastmod_u32(uint32_t x, uint64_t M, uint32_t n) {
  return (uint32_t)__umul64hi(M * (uint64_t)x, (uint64_t)n);
}
__device__ __forceinline__ void draw_sample(uint64_t& s, int count, uint64_t M, int32_t* idx) {
  for (int i = 0; i < 5;) {
    const int v = (int)fastmod_u32(cvrng_next(s), M, (uint32_t)count);
    bool dup = false;
    for (int j = 0; j < i; ++j) dup |= (idx[j] == v);
    idx[i] = v;
    if (!dup) ++i;
  }
}

__global__ void ransac_init_kernel(const double* __restrict__ pts, const int32_t* __restrict__ m,
                                   int n_cap, const double* __restrict__ K, int max_iters,
                                   RansacWs w, int32_t* __restrict__ n_inl,
                                   int32_t* __restrict__ status) {
  const int p = blockIdx.y;
  const int mp = m[p];
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < mp; i += gridDim.x * blockDim.x) {
    const double4 v = reinterpret_cast<const double4*>(pts)[(size_t)p * n_cap + i];
    double4 o;
    o.x = (v.x - cx) / fx; o.y = (v.y - cy) / fy;
    o.z = (v.z - cx) / fx; o.w = (v.w - cy) / fy;
    reinterpret_cast<double4*>(w.ptsn)[(size_t)p * n_cap + i] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    w.rng[p] = 0xffffffffffffffffull;
    w.niters[p] = max_iters;
    w.maxgood[p] = 0;
    w.done[p] = (mp < 5) ? 1 : 0;
    if (mp >= 5) w.active[atomicAdd(&w.n_active[0], 1)] = p;
    n_inl[p] = 0;
    status[p] = (mp < 5) ? VO_ERR_TOO_FEW_POINTS : VO_OK;
  }
}

// Every kernel of a round walks the list of pairs that are still iterating (built by the init
// kernel, rebuilt by every scan), with a fixed small grid: a round in which all pairs have
// already stopped costs a few microseconds instead of tens of thousands of empty blocks.
__global__ void ransac_sample_kernel(const int32_t* __restrict__ m, int round, int n_pairs, int it0,
                                     int chunk, int chunk_stride, RansacWs w) {
  const int n_act = w.n_active[round];
  const int32_t* list = w.active + (size_t)(round & 1) * n_pairs;
  for (int a = blockIdx.x * blockDim.x + threadIdx.x; a < n_act; a += gridDim.x * blockDim.x) {
    const int p = list[a];
    uint64_t s = w.rng[p];
    const int count = m[p];
    const uint64_t M = 0xffffffffffffffffull / (uint32_t)count + 1ull;
    int32_t* out = w.samples + (size_t)p * chunk_stride * 5;
    // only the samples the pair's bound still allows: a pair whose bound ends inside this round
    // stops after it (the bound never grows), so its generator state is not needed again
    const int left = w.niters[p] - it0;
    const int need = left < chunk ? left : chunk;
    for (int it = 0; it < need; ++it) {
      int32_t idx[5];
      draw_sample(s, count, M, idx);
#pragma unroll
      for (int k = 0; k < 5; ++k) out[5 * it + k] = idx[k];
    }
    w.rng[p] = s;
  }
}

// ------------------------------------------------------------------ K3 launch: warp per sample
constexpr int SOLVE_WARPS = 4;
// MINB = resident blocks per SM the register budget is sized for: the solver is a chain of
// dependent fp64 operations, so throughput comes from warps in flight, not from ILP.
static int g_solve_minb = 4;
template <int MINB>
__global__ void __launch_bounds__(SOLVE_WARPS * 32, MINB)
ransac_solve_kernel(int n_cap, int round, int n_pairs, int it0, int chunk, int chunk_stride,
                    RansacWs w) {
  __shared__ FiveptSmem sm[SOLVE_WARPS];
  __shared__ double p5[SOLVE_WARPS][20];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_act = w.n_active[round];
  const int32_t* list = w.active + (size_t)(round & 1) * n_pairs;
  const int per_pair = (chunk + SOLVE_WARPS - 1) / SOLVE_WARPS;
  const long n_items = (long)n_act * per_pair;
  for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int p = list[item / per_pair];
    const int s = (int)(item % per_pair) * SOLVE_WARPS + warp;
    if (s >= chunk) continue;
    int32_t* nm = w.nmodels + (size_t)p * chunk_stride + s;
    if (it0 + s >= w.niters[p]) {
      if (lane == 0) *nm = 0;
      continue;
    }
    const int32_t* idx = w.samples + ((size_t)p * chunk_stride + s) * 5;
    __syncwarp();
    if (lane < 20) {
      const int i = idx[lane >> 2];
      p5[warp][lane] = w.ptsn[((size_t)p * n_cap + i) * 4 + (lane & 3)];
    }
    __syncwarp();
    const int n = fivept_solve_warp(p5[warp], sm[warp],
                                    w.models + ((size_t)p * chunk_stride + s) * 90);
    if (lane == 0) *nm = n;
  }
}

// ------------------------------------------------------------------ K2: Sampson scoring
// EMEstimatorCallback::computeError restated with OpenCV's operation order and no FMA
// contraction: err = float((x2'Ex1)^2 / (a+b+c+d)); inlier iff err <= float(thr^2).
__device__ __forceinline__ float sampson_exact(const double* __restrict__ E, double x1, double y1,
                                               double x2, double y2) {
  const double ex0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x1), __dmul_rn(E[1], y1)), E[2]);
  const double ex1 = __dadd_rn(__dadd_rn(__dmul_rn(E[3], x1), __dmul_rn(E[4], y1)), E[5]);
  const double ex2 = __dadd_rn(__dadd_rn(__dmul_rn(E[6], x1), __dmul_rn(E[7], y1)), E[8]);
  const double et0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x2), __dmul_rn(E[3], y2)), E[6]);
  const double et1 = __dadd_rn(__dadd_rn(__dmul_rn(E[1], x2), __dmul_rn(E[4], y2)), E[7]);
  const double xe = __dadd_rn(__dadd_rn(__dmul_rn(x2, ex0), __dmul_rn(y2, ex1)), ex2);
  const double a = __dmul_rn(ex0, ex0), b = __dmul_rn(ex1, ex1);
  const double c = __dmul_rn(et0, et0), d = __dmul_rn(et1, et1);
  const double den = __dadd_rn(__dadd_rn(__dadd_rn(a, b), c), d);
  return __double2float_rn(__ddiv_rn(__dmul_rn(xe, xe), den));
}

// Division-free fast decision (FMA allowed): only correspondences within 1e-6 relative of the
// threshold (where float rounding of err could matter) take the exact path above.
__device__ __forceinline__ bool sampson_inlier(const double* __restrict__ E, double x1, double y1,
                                               double x2, double y2, float thr2f, double t_lo,
                                               double t_hi) {
  const double ex0 = fma(E[0], x1, fma(E[1], y1, E[2]));
  const double ex1 = fma(E[3], x1, fma(E[4], y1, E[5]));
  const double ex2 = fma(E[6], x1, fma(E[7], y1, E[8]));
  const double et0 = fma(E[0], x2, fma(E[3], y2, E[6]));
  const double et1 = fma(E[1], x2, fma(E[4], y2, E[7]));
  const double xe = fma(x2, ex0, fma(y2, ex1, ex2));
  const double den = fma(ex0, ex0, fma(ex1, ex1, fma(et0, et0, et1 * et1)));
  const double num = xe * xe;
  bool inl = num < t_lo * den;
  if (!inl && !(num > t_hi * den)) inl = sampson_exact(E, x1, y1, x2, y2) <= thr2f;
  return inl;
}

// First sieve in fp32 (twice the fp64 rate): normalised coordinates and E entries are O(1), so
// the residual x2'Ex1 carries an absolute error of a few 1e-7 against ~7e-4 at the threshold,
// i.e. < 1.5e-3 relative on its square; only correspondences within +-5e-3 of the threshold go on
// to the fp64 decision above (and from there, within 1e-6, to the exact OpenCV-ordered one).
// returns 1 inlier, 0 outlier, -1 undecided
__device__ __forceinline__ int sampson_sieve_f32(const float* __restrict__ E, float x1, float y1,
                                                 float x2, float y2, float t_lo, float t_hi) {
  const float ex0 = fmaf(E[0], x1, fmaf(E[1], y1, E[2]));
  const float ex1 = fmaf(E[3], x1, fmaf(E[4], y1, E[5]));
  const float ex2 = fmaf(E[6], x1, fmaf(E[7], y1, E[8]));
  const float et0 = fmaf(E[0], x2, fmaf(E[3], y2, E[6]));
  const float et1 = fmaf(E[1], x2, fmaf(E[4], y2, E[7]));
  const float xe = fmaf(x2, ex0, fmaf(y2, ex1, ex2));
  const float den = fmaf(ex0, ex0, fmaf(ex1, ex1, fmaf(et0, et0, et1 * et1)));
  const float num = xe * xe;
  if (num < t_lo * den) return 1;
  if (num > t_hi * den) return 0;
  return -1;
}

// The same sieve for TWO correspondences at once in packed fp32 (FFMA2 / FMUL2, sm_100): every
// lane of the pair is rounded exactly like the scalar form above (same nesting of the fused
// operations), so the decisions are identical; it halves the issue slots of the arithmetic.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ void sampson_sieve_f32x2(const f32x2* __restrict__ E, f32x2 x1, f32x2 y1,
                                                    f32x2 x2, f32x2 y2, f32x2 t_lo, f32x2 t_hi,
                                                    int& dec_lo, int& dec_hi) {
  const f32x2 ex0 = fma2(E[0], x1, fma2(E[1], y1, E[2]));
  const f32x2 ex1 = fma2(E[3], x1, fma2(E[4], y1, E[5]));
  const f32x2 ex2 = fma2(E[6], x1, fma2(E[7], y1, E[8]));
  const f32x2 et0 = fma2(E[0], x2, fma2(E[3], y2, E[6]));
  const f32x2 et1 = fma2(E[1], x2, fma2(E[4], y2, E[7]));
  const f32x2 xe = fma2(x2, ex0, fma2(y2, ex1, ex2));
  const f32x2 den = fma2(ex0, ex0, fma2(ex1, ex1, fma2(et0, et0, mul2(et1, et1))));
  const f32x2 num = mul2(xe, xe);
  float n0, n1, l0, l1, h0, h1;
  unpack2(num, n0, n1);
  unpack2(mul2(t_lo, den), l0, l1);
  unpack2(mul2(t_hi, den), h0, h1);
  dec_lo = n0 < l0 ? 1 : (n0 > h0 ? 0 : -1);
  dec_hi = n1 < l1 ? 1 : (n1 > h1 ? 0 : -1);
}

// findEssentialMat: threshold /= (fx+fy)/2 ; findInliers compares with float(thr*thr)
__device__ __forceinline__ float thr2_from_K(const double* __restrict__ K, double thr_px) {
  const double thr = thr_px / ((K[0] + K[4]) / 2.0);
  return (float)(thr * thr);
}

constexpr int SCORE_WARPS = 8;
constexpr int SCORE_TILE = 1024;       // correspondences staged in shared memory at a time
// One block = one pair and 8 of its samples (warp per sample).  The pair's correspondences go
// through shared memory as float4 tiles shared by the 8 warps; per tile and model the nine
// coefficients sit in registers, so a test is one 16-byte shared-memory load and ~20 fp32
// operations.
__global__ void __launch_bounds__(SCORE_WARPS * 32)
ransac_score_kernel(const int32_t* __restrict__ m, int n_cap, int round, int n_pairs, int chunk,
                    int chunk_stride, const double* __restrict__ K, double thr_px, RansacWs w) {
  __shared__ double Es[SCORE_WARPS][90];
  __shared__ float Ef[SCORE_WARPS][90];
  // correspondences i and i + 32 of every group of 64 side by side: one 8-byte load per coordinate
  // yields the packed pair the FFMA2 sieve works on
  __shared__ float2 tx1[SCORE_TILE / 2], ty1[SCORE_TILE / 2], tx2[SCORE_TILE / 2], ty2[SCORE_TILE / 2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_act = w.n_active[round];
  const int32_t* list = w.active + (size_t)(round & 1) * n_pairs;
  const int per_pair = (chunk + SCORE_WARPS - 1) / SCORE_WARPS;
  const long n_items = (long)n_act * per_pair;
  const float thr2f = thr2_from_K(K, thr_px);
  const double t_lo = (double)thr2f * (1.0 - 1e-6), t_hi = (double)thr2f * (1.0 + 1e-6);
  const float tf_lo = thr2f * (1.f - 5e-3f), tf_hi = thr2f * (1.f + 5e-3f);
  for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int p = list[item / per_pair];
    const int s = (int)(item % per_pair) * SCORE_WARPS + warp;
    const int nm = s < chunk ? w.nmodels[(size_t)p * chunk_stride + s] : 0;
    if (!__syncthreads_or(nm > 0)) continue;      // samples beyond the pair's bound: nothing to score
    const double* src = w.models + ((size_t)p * chunk_stride + (s < chunk ? s : 0)) * 90;
    for (int e = lane; e < nm * 9; e += 32) { const double v = src[e]; Es[warp][e] = v; Ef[warp][e] = (float)v; }
    const int mp = m[p];
    const double4* pts = reinterpret_cast<const double4*>(w.ptsn) + (size_t)p * n_cap;
    int cnt = 0;
    for (int t0 = 0; t0 < mp; t0 += SCORE_TILE) {
      const int tn = mp - t0 < SCORE_TILE ? mp - t0 : SCORE_TILE;
      __syncthreads();                       // the previous tile (or item) is no longer read
      for (int i = threadIdx.x; i < ((tn + 63) & ~63); i += SCORE_WARPS * 32) {
        double4 v = make_double4(0, 0, 0, 0);
        if (i < tn) v = ldg_d4(pts + t0 + i);
        const int j = (i >> 6) * 32 + (i & 31), half = (i >> 5) & 1;
        reinterpret_cast<float*>(&tx1[j])[half] = (float)v.x;
        reinterpret_cast<float*>(&ty1[j])[half] = (float)v.y;
        reinterpret_cast<float*>(&tx2[j])[half] = (float)v.z;
        reinterpret_cast<float*>(&ty2[j])[half] = (float)v.w;
      }
      __syncthreads();
      const f32x2 tlo2 = pack2(tf_lo, tf_lo), thi2 = pack2(tf_hi, tf_hi);
      for (int k = 0; k < nm; ++k) {
        f32x2 e[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) { const float c = Ef[warp][9 * k + q]; e[q] = pack2(c, c); }
        int local = 0;
        // two correspondences per trip in packed fp32; the fp64 decision of the rare undecided
        // ones sits behind one branch per trip
        for (int i = lane; i < tn; i += 64) {
          const int i2 = i + 32;
          const bool has2 = i2 < tn;
          const int j = (i >> 6) * 32 + lane;
          int dec, dec2;
          sampson_sieve_f32x2(e, *reinterpret_cast<const f32x2*>(&tx1[j]), *reinterpret_cast<const f32x2*>(&ty1[j]),
                              *reinterpret_cast<const f32x2*>(&tx2[j]), *reinterpret_cast<const f32x2*>(&ty2[j]),
                              tlo2, thi2, dec, dec2);
          if (!has2) dec2 = 0;
          if ((dec | dec2) < 0) {
            if (dec < 0) {
              const double4 v = ldg_d4(pts + t0 + i);
              dec = sampson_inlier(&Es[warp][9 * k], v.x, v.y, v.z, v.w, thr2f, t_lo, t_hi);
            }
            if (dec2 < 0) {
              const double4 v = ldg_d4(pts + t0 + i2);
              dec2 = sampson_inlier(&Es[warp][9 * k], v.x, v.y, v.z, v.w, thr2f, t_lo, t_hi);
            }
          }
          local += dec + dec2;
        }
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if (lane == k) cnt += local;
      }
    }
    if (lane < nm) w.counts[((size_t)p * chunk_stride + s) * 10 + lane] = cnt;
  }
}

// standalone scoring of arbitrary hypotheses (C-ABI vo_sampson_score): warp per hypothesis
__global__ void __launch_bounds__(256)
sampson_score_kernel(const double* __restrict__ E, int H, const double* __restrict__ ptsn, int m,
                     float thr2f, int32_t* __restrict__ counts, uint8_t* __restrict__ masks) {
  const int h = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (h >= H) return;
  double e[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) e[k] = E[(size_t)h * 9 + k];
  const double4* pts = reinterpret_cast<const double4*>(ptsn);
  int cnt = 0;
  for (int i = lane; i < m; i += 32) {
    const double4 v = ldg_d4(pts + i);
    const bool inl = sampson_exact(e, v.x, v.y, v.z, v.w) <= thr2f;
    cnt += inl;
    if (masks) masks[(size_t)h * m + i] = inl;
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (lane == 0) counts[h] = cnt;
}

// ------------------------------------------------------------------ ordered scan (the sequential rule)
__device__ int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = fmax(p, 0.0); p = fmin(p, 1.0);
  ep = fmax(ep, 0.0); ep = fmin(ep, 1.0);
  double num = fmax(1.0 - p, DBL_MIN);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < DBL_MIN) return 0;
  num = log(num);
  denom = log(denom);
  return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : (int)rint(num / denom);
}

// One WARP per active pair.  OpenCV's rule is sequential (strict-greater best, bound shrinks after
// every improvement, iterations at or beyond the bound never run), but it reduces to prefix
// maxima: a sample improves the best iff its top count exceeds the running maximum before it, the
// bound after an improvement is min(previous, f(count)) with f decreasing in the count, so the
// bound in force at sample s is min(carry, f(running max before s)).  Per block of 32 samples:
// lane = sample, warp prefix max, f evaluated only by lanes that follow an improvement, first
// dead sample by ballot, winner = last improving live lane.
constexpr int SCAN_WARPS = 4;
__global__ void __launch_bounds__(SCAN_WARPS * 32)
ransac_scan_kernel(const int32_t* __restrict__ m, int round, int n_pairs, int it0, int chunk,
                   int chunk_stride, double conf, int adaptive, RansacWs w,
                   double* __restrict__ E_out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n_act = w.n_active[round];
  const int32_t* list = w.active + (size_t)(round & 1) * n_pairs;
  int32_t* next = w.active + (size_t)((round + 1) & 1) * n_pairs;
  for (int a = blockIdx.x * SCAN_WARPS + (threadIdx.x >> 5); a < n_act; a += gridDim.x * SCAN_WARPS) {
    const int p = list[a];
    int niters = w.niters[p], maxgood = w.maxgood[p];
    const int count = m[p];
    int bs = -1, bk = -1;
    bool stop = false;
    for (int s0 = 0; s0 < chunk && !stop; s0 += 32) {
      const int s = s0 + lane;
      // this sample's best count (first k on ties = the one the sequential rule keeps)
      int c = -1, k_best = 0;
      if (s < chunk) {
        const int nm = w.nmodels[(size_t)p * chunk_stride + s];
        for (int k = 0; k < nm; ++k) {
          const int ck = w.counts[((size_t)p * chunk_stride + s) * 10 + k];
          if (ck > c) { c = ck; k_best = k; }
        }
      }
      // inclusive prefix max over the lanes, then the running max BEFORE this lane
      int incl = c;
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl = max(incl, v);
      }
      int before = __shfl_up_sync(FULL, incl, 1);
      before = (lane == 0) ? maxgood : max(before, maxgood);
      // bound in force at this sample
      int bound = niters;
      if (adaptive && before > max(maxgood, 4))
        bound = ransac_update_num_iters(conf, (double)(count - before) / count, 5, niters);
      const bool dead = (s >= chunk) || (it0 + s >= bound);
      const unsigned dead_mask = __ballot_sync(FULL, dead);
      const int first_dead = dead_mask ? __ffs(dead_mask) - 1 : 32;
      const bool improves = (lane < first_dead) && (c > max(before, 4));
      const unsigned imp = __ballot_sync(FULL, improves);
      if (imp) {
        const int win = 31 - __clz(imp);          // last improvement = highest count so far
        maxgood = __shfl_sync(FULL, c, win);
        bs = s0 + win;
        bk = __shfl_sync(FULL, k_best, win);
        if (adaptive) niters = ransac_update_num_iters(conf, (double)(count - maxgood) / count, 5, niters);
      }
      if (first_dead < 32) stop = true;
    }
    if (bs >= 0 && lane < 9) {
      const double* src = w.models + ((size_t)p * chunk_stride + bs) * 90 + 9 * bk;
      E_out[(size_t)p * 9 + lane] = src[lane];
    }
    if (lane == 0) {
      w.niters[p] = niters;
      w.maxgood[p] = maxgood;
      if (it0 + chunk >= niters) w.done[p] = 1;
      else next[atomicAdd(&w.n_active[round + 1], 1)] = p;
    }
  }
}

__global__ void ransac_mask_kernel(const int32_t* __restrict__ m, int n_cap,
                                   const double* __restrict__ K, double thr_px, RansacWs w, double* __restrict__ E_out,
                                   uint8_t* __restrict__ mask, int32_t* __restrict__ n_inl,
                                   int32_t* __restrict__ n_iters, int32_t* __restrict__ status) {
  const int p = blockIdx.y;
  const int mp = m[p];
  const float thr2f = thr2_from_K(K, thr_px);
  const int good = w.maxgood[p];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    n_inl[p] = good;
    n_iters[p] = w.niters[p];
    if (good == 0) {
      if (status[p] == VO_OK) status[p] = VO_ERR_TOO_FEW_POINTS;  // no model found
      for (int e = 0; e < 9; ++e) E_out[(size_t)p * 9 + e] = 0.0;
    }
  }
  double e[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) e[k] = good ? E_out[(size_t)p * 9 + k] : 0.0;
  const double4* pts = reinterpret_cast<const double4*>(w.ptsn) + (size_t)p * n_cap;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_cap; i += gridDim.x * blockDim.x) {
    uint8_t v = 0;
    if (i < mp && good) {
      const double4 q = pts[i];
      v = sampson_exact(e, q.x, q.y, q.z, q.w) <= thr2f;
    }
    mask[(size_t)p * n_cap + i] = v;
  }
}

}  // namespace vo

using namespace vo;

// Iterations per round: 16, then x4 per round up to the cap (adaptive; iterations at or beyond a
// pair's current bound are skipped inside the kernels, so a long late round costs only the pairs
// that still need it), or the cap straight away (fixed hypothesis count).  One definition, also
// exported so that the host side (launch counts, bench accounting) cannot drift from it.
static int round_schedule(int max_iters, int C, int adaptive, int* chunks, int max_rounds) {
  int it0 = 0, n = 0;
  int chunk = RANSAC_CHUNK0 < C ? RANSAC_CHUNK0 : C;
  if (!adaptive) chunk = C;
  while (it0 < max_iters && n < max_rounds) {
    if (it0 + chunk > max_iters) chunk = max_iters - it0;
    if (n == max_rounds - 1) chunk = max_iters - it0 < C ? max_iters - it0 : C;
    if (chunks) chunks[n] = chunk;
    it0 += chunk;
    ++n;
    if (adaptive) chunk = (chunk * 4 < C) ? chunk * 4 : C;
    else chunk = C;
  }
  return n;
}

extern "C" int vo_ransac_round_schedule(int max_iters, int n_pairs, int adaptive, int* chunks_h,
                                        int max_rounds) {
  if (max_iters <= 0 || max_rounds <= 0) return 0;
  return round_schedule(max_iters, chunk_cap(max_iters, n_pairs), adaptive, chunks_h, max_rounds);
}

extern "C" int vo_debug_ransac(int solve_minb) {
  g_solve_minb = solve_minb;
  return VO_OK;
}

extern "C" size_t vo_essential_workspace_bytes(int n_pairs, int n_cap, int max_iters) {
  if (n_pairs <= 0 || n_cap <= 0) return 0;
  return carve(nullptr, n_pairs, n_cap, chunk_cap(max_iters, n_pairs), max_iters).bytes;
}

extern "C" int vo_essential_ransac_batch(const double* pts, const int32_t* m, int n_cap,
                                         int n_pairs, const double* K, double conf, double thr_px,
                                         int max_iters, int adaptive, double* E, uint8_t* mask,
                                         int32_t* n_inl, int32_t* n_iters, int32_t* status,
                                         void* ws, size_t ws_bytes, void* stream) {
  VO_REQUIRE(pts && m && K && E && mask && n_inl && n_iters && status && ws, VO_ERR_INVALID_ARG,
             "vo_essential_ransac_batch: null pointer");
  VO_REQUIRE(n_cap > 0 && n_pairs >= 0 && max_iters > 0, VO_ERR_INVALID_ARG,
             "vo_essential_ransac_batch: bad sizes");
  VO_REQUIRE(n_pairs <= 65535, VO_ERR_INVALID_ARG,
             "vo_essential_ransac_batch: at most 65535 pairs per call");
  // the caller may have sized the workspace for a larger batch (smaller per-round cap): shrink
  // the cap until the carve fits
  int C = chunk_cap(max_iters, n_pairs);
  RansacWs w = carve(ws, n_pairs, n_cap, C, max_iters);
  while (w.bytes > ws_bytes && C > RANSAC_CHUNK_MAX) {
    C = C / 2 > RANSAC_CHUNK_MAX ? C / 2 : RANSAC_CHUNK_MAX;
    w = carve(ws, n_pairs, n_cap, C, max_iters);
  }
  VO_REQUIRE(ws_bytes >= w.bytes, VO_ERR_WORKSPACE, "vo_essential_ransac_batch: workspace too small");
  if (n_pairs == 0) return VO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = (n_cap + 255) / 256;
  cudaMemsetAsync(w.n_active, 0, (size_t)(w.max_rounds + 1) * sizeof(int32_t), st);
  {
    ProfScope prof(9, st);
    ransac_init_kernel<<<dim3(nb < 8 ? nb : 8, n_pairs), 256, 0, st>>>(pts, m, n_cap, K, max_iters, w,
                                                                      n_inl, status);
  }
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  int sched[128];
  const int n_rounds = round_schedule(max_iters, C, adaptive, sched, w.max_rounds < 128 ? w.max_rounds : 128);
  const int list_blocks = (n_pairs + 63) / 64 < sms ? (n_pairs + 63) / 64 : sms;
  int it0 = 0;
  for (int round = 0; round < n_rounds; ++round) {
    const int chunk = sched[round];
    {
      ProfScope prof(9, st);
      ransac_sample_kernel<<<list_blocks, 64, 0, st>>>(m, round, n_pairs, it0, chunk, C, w);
    }
    {
      // grid sized for the first (full) round; later rounds loop over what is left
      const long want = (long)n_pairs * ((chunk + SOLVE_WARPS - 1) / SOLVE_WARPS);
      const long cap = (long)sms * g_solve_minb * 4;
      const int sg = (int)(want < cap ? want : cap);
      ProfScope prof(3, st);
      switch (g_solve_minb) {
        case 2: ransac_solve_kernel<2><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
        case 3: ransac_solve_kernel<3><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
        case 5: ransac_solve_kernel<5><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
        case 6: ransac_solve_kernel<6><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
        case 8: ransac_solve_kernel<8><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
        default: ransac_solve_kernel<4><<<sg, SOLVE_WARPS * 32, 0, st>>>(n_cap, round, n_pairs, it0, chunk, C, w); break;
      }
    }
    {
      const long want = (long)n_pairs * ((chunk + SCORE_WARPS - 1) / SCORE_WARPS);
      const long cap = (long)sms * 8 * 4;
      const int sg = (int)(want < cap ? want : cap);
      ProfScope prof(1, st);
      ransac_score_kernel<<<sg, SCORE_WARPS * 32, 0, st>>>(m, n_cap, round, n_pairs, chunk, C, K,
                                                          thr_px, w);
    }
    {
      const int want = (n_pairs + SCAN_WARPS - 1) / SCAN_WARPS;
      ProfScope prof(9, st);
      ransac_scan_kernel<<<want < sms * 8 ? want : sms * 8, SCAN_WARPS * 32, 0, st>>>(
          m, round, n_pairs, it0, chunk, C, conf, adaptive, w, E);
    }
    it0 += chunk;
  }
  {
    ProfScope prof(9, st);
    ransac_mask_kernel<<<dim3(nb < 8 ? nb : 8, n_pairs), 256, 0, st>>>(m, n_cap, K, thr_px, w, E, mask,
                                                                      n_inl, n_iters, status);
  }
  return check_launch("vo_essential_ransac_batch");
}

__global__ void __launch_bounds__(SOLVE_WARPS * 32, 4)
fivept_standalone_kernel(const double* __restrict__ ptsn, const int32_t* __restrict__ samples,
                         int n_samples, double* __restrict__ E, int32_t* __restrict__ n_models) {
  __shared__ FiveptSmem sm[SOLVE_WARPS];
  __shared__ double p5[SOLVE_WARPS][20];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s = blockIdx.x * SOLVE_WARPS + warp;
  if (s >= n_samples) return;
  if (lane < 20) p5[warp][lane] = ptsn[(size_t)samples[5 * s + (lane >> 2)] * 4 + (lane & 3)];
  __syncwarp();
  const int n = fivept_solve_warp(p5[warp], sm[warp], E + (size_t)s * 90);
  if (lane == 0) n_models[s] = n;
}

extern "C" int vo_fivept_solve(const double* pts_n, int m, const int32_t* samples, int n_samples,
                               double* E, int32_t* n_models, void* stream) {
  VO_REQUIRE(pts_n && samples && E && n_models, VO_ERR_INVALID_ARG, "vo_fivept_solve: null pointer");
  VO_REQUIRE(m >= 5, VO_ERR_TOO_FEW_POINTS, "vo_fivept_solve: needs >= 5 points (got %d)", m);
  if (n_samples <= 0) return VO_OK;
  fivept_standalone_kernel<<<(n_samples + SOLVE_WARPS - 1) / SOLVE_WARPS, SOLVE_WARPS * 32, 0,
                             (cudaStream_t)stream>>>(pts_n, samples, n_samples, E, n_models);
  return check_launch("vo_fivept_solve");
}

extern "C" int vo_sampson_score(const double* E, int H, const double* pts_n, int m, float thr2,
                                int32_t* counts, uint8_t* masks, void* stream) {
  VO_REQUIRE(E && pts_n && counts, VO_ERR_INVALID_ARG, "vo_sampson_score: null pointer");
  if (H <= 0) return VO_OK;
  sampson_score_kernel<<<(H + 7) / 8, 256, 0, (cudaStream_t)stream>>>(E, H, pts_n, m, thr2, counts,
                                                                     masks);
  return check_launch("vo_sampson_score");
}

__global__ void samples_standalone_kernel(int count, int n_iters, int32_t* out) {
  if (blockIdx.x || threadIdx.x) return;
  uint64_t s = 0xffffffffffffffffull;
  const uint64_t M = 0xffffffffffffffffull / (uint32_t)count + 1ull;
  for (int it = 0; it < n_iters; ++it) {
    int32_t idx[5];
    draw_sample(s, count, M, idx);
    for (int k = 0; k < 5; ++k) out[5 * it + k] = idx[k];
  }
}

extern "C" int vo_ransac_samples(int count, int n_iters, int32_t* samples, void* stream) {
  VO_REQUIRE(samples, VO_ERR_INVALID_ARG, "vo_ransac_samples: null pointer");
  VO_REQUIRE(count >= 5, VO_ERR_TOO_FEW_POINTS, "vo_ransac_samples: needs >= 5 points");
  if (n_iters <= 0) return VO_OK;
  samples_standalone_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(count, n_iters, samples);
  return check_launch("vo_ransac_samples");
}
