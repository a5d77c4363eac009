// orb.cu — N4: ORB feature extraction, the detector the reference hands to
// FrameGenerator (src/frame_generator.py:25 `self.detector.detectAndCompute(frame.image, None)`,
// src/visual_slam.py:16 `cv2.ORB_create()`, src/image_and_keypoints.py:8).
//
// Restates cv2.ORB_create(nfeatures, scaleFactor, nlevels, edgeThreshold, 0, 2, HARRIS_SCORE, 31,
// fastThreshold).detectAndCompute for 8-bit gray images, bit for bit (oracle/cv_orb.py pins every
// stage against the cv2 binary):
//   pyramid     INTER_LINEAR_EXACT: 8-bit fixed-point taps, 8.8 row pass, 16.16 column pass
//   FAST-9/16   score = max over the 16 arcs of 9 circle pixels of min |v - p| - 1, 3x3 strict
//               non-maximum suppression, image-border filter (edgeThreshold)
//   retainBest  2 n_level best by FAST score, Harris response (7x7 block, integer gradients,
//               un-fused float formula), n_level best by Harris; ties at the cut are kept
//   orientation intensity centroid over the radius-15 disc, cv::fastAtan2 polynomial
//   rBRIEF      7x7 sigma-2 float Gaussian (OpenCV's separable-filter order, fused multiply-adds
//               where OpenCV's AVX2 code has them), 256 rotated point pairs, round half to even
// Keypoints of a level come out in (y, x) order (cv2's order inside a level is whatever
// std::nth_element leaves; the SET is the same).
//
// Everything is byte / integer work on a 2 MB image: HBM / L2 bound.  Batched over the frames of a
// sequence (grid.z / grid.y = frame), one launch per stage and pyramid level:
//   orb_coef_kernel    interpolation tables of all levels (shared by the frames)
//   orb_resize_kernel  level l from level l-1
//   orb_fast_kernel    FAST score + NMS + border filter fused over a shared-memory tile (scores
//                      never reach HBM): packed corners (y | x | score) + score histogram
//   orb_cut1_kernel    histogram -> score cut, compaction of the survivors
//   orb_harris_kernel  one warp per candidate
//   orb_keep / orb_place  Harris cut and (y, x) ranking by counting, key-point records
//   orb_describe_kernel  one 256-thread block per key point: 47x47 source patch in shared memory
//                      -> orientation, blurred 41x41 patch, 256 comparisons, 8 ballots
#include "common.cuh"

namespace vo {

constexpr int ORB_MAX_LEVELS = 16;
constexpr int ORB_HALF_PATCH = 15;
constexpr int ORB_STAGE = 4096;                 // candidates of one level staged in shared memory
constexpr int ORB_BR = 20;                      // radius of the blurred patch (pattern reach 18.4)
constexpr int ORB_SR = ORB_BR + 3;              // radius of the source patch
constexpr int ORB_SW = 2 * ORB_SR + 1;          // 47
constexpr int ORB_BW = 2 * ORB_BR + 1;          // 41
constexpr int FAST_TW = 126, FAST_TH = 30;      // NMS tile; scores for a 128x32 ring, source 134x38

struct OrbLevel {
  int w, h;             // level size
  int pitch;            // bytes per row: width for the caller's base level, 16-byte multiple above
  int nfeat;            // features wanted on this level
  int cap;              // capacity of its corner lists
  float scale;          // layerScale
  size_t pix_off;       // offset of the level inside one frame's pyramid block (levels >= 1)
  size_t list_off;      // offset of its corner list inside one frame's list block
  size_t coef_off;      // offset of (x table, y table) inside the coefficient block
};
struct OrbGeom {
  int nlevels, n_frames, width, height;
  int edge, fast_thr, cap_out;
  size_t pyr_stride;    // bytes of levels >= 1 per frame
  size_t list_stride;   // corner-list entries per frame
  OrbLevel lv[ORB_MAX_LEVELS];
};

__constant__ int8_t c_orb_pattern[256][4] = {
#include "orb_pattern.inc"
};
__constant__ int c_orb_umax[ORB_HALF_PATCH + 1] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
// float32 taps of cv::getGaussianKernel(7, 2, CV_32F)
__constant__ float c_orb_taps[7] = {0x1.1f5f62p-4f, 0x1.0c70fcp-3f, 0x1.869472p-3f, 0x1.ba95c0p-3f,
                                    0x1.869472p-3f, 0x1.0c70fcp-3f, 0x1.1f5f62p-4f};

// ------------------------------------------------------------------ pyramid
// cv::resize INTER_LINEAR_EXACT, uint8: src position (d + 0.5) * src/dst - 0.5 in double, tap
// weights rounded to 8 fractional bits; entry = (offset of the left/top tap, weight of the
// right/bottom tap); the other weight is 256 - it.  Clamped taps read the same pixel twice.
__global__ void orb_coef_kernel(OrbGeom g, int2* __restrict__ coef) {
  const int l = blockIdx.y + 1;
  if (l >= g.nlevels) return;
  const OrbLevel L = g.lv[l], P = g.lv[l - 1];
  int2* tab = coef + L.coef_off;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < L.w + L.h; i += gridDim.x * blockDim.x) {
    const bool isx = i < L.w;
    const int d = isx ? i : i - L.w, src = isx ? P.w : P.h, dst = isx ? L.w : L.h;
    const double scale = (double)src / (double)dst;
    const double x = __dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
    int fx = (int)floor(x);
    double fr = x - (double)fx;
    if (fx < 0) { fx = 0; fr = 0.0; }
    if (fx >= src - 1) { fx = src - 1; fr = 0.0; }
    tab[i] = make_int2(fx, (int)rint(fr * 256.0));
  }
}

// four consecutive output pixels per thread, one aligned 32-bit store (the row pitch of levels >= 1
// is a multiple of 16; the padding columns receive the clamped edge taps and are never read)
__global__ void __launch_bounds__(128)
orb_resize_kernel(OrbGeom g, int l, const uint8_t* __restrict__ images, uint8_t* __restrict__ pyr,
                  const int2* __restrict__ coef) {
  const OrbLevel L = g.lv[l], P = g.lv[l - 1];
  const int f = blockIdx.z, y = blockIdx.y;
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (x4 >= L.w) return;
  const uint8_t* src = (l == 1) ? images + (size_t)f * g.width * g.height
                                : pyr + (size_t)f * g.pyr_stride + P.pix_off;
  uint8_t* dst = pyr + (size_t)f * g.pyr_stride + L.pix_off;
  const int2 cy = coef[L.coef_off + L.w + y];
  const int y0 = cy.x, y1 = min(cy.x + 1, P.h - 1);
  const uint32_t ay1 = cy.y, ay0 = 256 - cy.y;
  const uint8_t* r0 = src + (size_t)y0 * P.pitch;
  const uint8_t* r1 = src + (size_t)y1 * P.pitch;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int2 cx = coef[L.coef_off + min(x4 + k, L.w - 1)];
    const int x0 = cx.x, x1 = min(cx.x + 1, P.w - 1);
    const uint32_t ax1 = cx.y, ax0 = 256 - cx.y;
    const uint32_t h0 = __ldg(r0 + x0) * ax0 + __ldg(r0 + x1) * ax1;
    const uint32_t h1 = __ldg(r1 + x0) * ax0 + __ldg(r1 + x1) * ax1;
    out |= (((h0 * ay0 + h1 * ay1) + (1u << 15)) >> 16) << (8 * k);
  }
  *reinterpret_cast<uint32_t*>(dst + (size_t)y * L.pitch + x4) = out;
}

__device__ __forceinline__ const uint8_t* orb_level_ptr(const OrbGeom& g, int l, int f,
                                                        const uint8_t* images, const uint8_t* pyr) {
  return l == 0 ? images + (size_t)f * g.width * g.height
                : pyr + (size_t)f * g.pyr_stride + g.lv[l].pix_off;
}

// ------------------------------------------------------------------ FAST-9/16 + NMS
// Circle of radius 3, clockwise from (0, 3); (dx, dy)
__device__ __forceinline__ void fast_circle(const uint8_t* c, int pitch, int (&d)[16]) {
  const int v = c[0];
  d[0] = v - c[3 * pitch];          d[1] = v - c[3 * pitch + 1];   d[2] = v - c[2 * pitch + 2];
  d[3] = v - c[pitch + 3];          d[4] = v - c[3];               d[5] = v - c[-pitch + 3];
  d[6] = v - c[-2 * pitch + 2];     d[7] = v - c[-3 * pitch + 1];  d[8] = v - c[-3 * pitch];
  d[9] = v - c[-3 * pitch - 1];     d[10] = v - c[-2 * pitch - 2]; d[11] = v - c[-pitch - 3];
  d[12] = v - c[-3];                d[13] = v - c[pitch - 3];      d[14] = v - c[2 * pitch - 2];
  d[15] = v - c[3 * pitch - 1];
}
// corner at threshold t: 9 contiguous circle pixels all darker than v - t or all brighter than v + t
__device__ __forceinline__ bool fast_corner(const int (&d)[16], int t) {
  uint32_t mb = 0, md = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    mb |= (uint32_t)(d[k] > t) << k;
    md |= (uint32_t)(d[k] < -t) << k;
  }
  // a circular run of >= 9 set bits in a 16-bit mask
  auto run9 = [](uint32_t m) {
    const uint32_t x = m | (m << 16);
    const uint32_t r2 = x & (x >> 1), r4 = r2 & (r2 >> 2), r8 = r4 & (r4 >> 4);
    return (r8 & (x >> 8) & 0xffffu) != 0u;
  };
  return run9(mb) || run9(md);
}
// score + 1 = max over the 16 arcs of min |d| = the largest t + 1 at which the pixel still is a
// corner (the test is monotone in t): bisection over [thr, 255].  (A direct min / max tree over
// the arcs compiled to code that returned wrong values on the B200 although its PTX read
// correctly; the run test is what the detection itself uses, so score and detection cannot
// disagree.)
__device__ __forceinline__ int fast_raw_score(const int (&d)[16], int thr) {
  int lo_t = thr, hi_t = 255;
  while (hi_t - lo_t > 1) {
    const int mid = (lo_t + hi_t) >> 1;
    if (fast_corner(d, mid)) lo_t = mid; else hi_t = mid;
  }
  return lo_t + 1;   // > thr for a corner; cv2's response is this - 1
}

// corner word: y << 20 | x << 8 | raw score (so unsigned order = (y, x) order).
// Tile = 126 x 30 pixels: the score ring (tile + 1) is 128 x 32 = 4 lane strides x 4 rows per
// warp, the source tile (ring + 3) 134 x 38 bytes.  Loops are (row per warp, column per lane):
// no index divisions, bounds tests hoisted per row.
__global__ void __launch_bounds__(256)
orb_fast_kernel(OrbGeom g, int l, const uint8_t* __restrict__ images, const uint8_t* __restrict__ pyr,
                uint32_t* __restrict__ list1, int32_t* __restrict__ cnt1, int32_t* __restrict__ hist,
                uint8_t* __restrict__ score_map) {
  constexpr int SP = FAST_TW + 8 + 6, SH = FAST_TH + 8;   // source tile 134 (+ alignment shift, pad) x 38
  constexpr int CP = FAST_TW + 2, CH = FAST_TH + 2;       // score ring 128 x 32
  __shared__ __align__(16) uint8_t s_src[SH][SP];
  __shared__ uint8_t s_sc[CH][CP + 4];
  __shared__ uint16_t s_cand[CH * CP], s_corner[CH * CP];
  __shared__ uint32_t s_out[(FAST_TW / 2 + 1) * (FAST_TH / 2 + 1)];
  __shared__ int s_n, s_base, s_nc, s_nk;
  const OrbLevel L = g.lv[l];
  const int f = blockIdx.z;
  const uint8_t* img = orb_level_ptr(g, l, f, images, pyr);
  const int tx0 = blockIdx.x * FAST_TW, ty0 = blockIdx.y * FAST_TH;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { s_n = 0; s_nc = 0; s_nk = 0; }
  // source tile: aligned 32-bit loads when rows start on word boundaries (every level the library
  // allocates; the caller's base level when its width is a multiple of 4), the tile then sits
  // `al` bytes to the right in shared memory; byte loads otherwise
  const bool words = (L.pitch & 3) == 0 && ((uintptr_t)img & 3) == 0;
  const int al = words ? ((tx0 - 4) & 3) : 0;
  if (words) {
    constexpr int NW = SP / 4;                               // 35 words per row
    const int gx = tx0 - 4 - al;                             // multiple of 4 (may be -4)
    constexpr int NIT = (SH * NW + 255) / 256;               // 6: all loads in flight, then the stores
    uint32_t v[NIT];
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int i = threadIdx.x + 256 * k;
      const int sy = i / NW, wi = i - sy * NW;
      const int y = ty0 - 4 + sy, x = gx + 4 * wi;
      v[k] = 0;
      if (i < SH * NW && y >= 0 && y < L.h && x >= 0 && x < L.pitch)
        v[k] = __ldg(reinterpret_cast<const uint32_t*>(img + (size_t)y * L.pitch + x));
    }
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int i = threadIdx.x + 256 * k;
      if (i < SH * NW) reinterpret_cast<uint32_t*>(&s_src[0][0])[i] = v[k];
    }
  } else {
    for (int sy = warp; sy < SH; sy += 8) {
      const int y = ty0 - 4 + sy;
      const bool yok = y >= 0 && y < L.h;
      const uint8_t* row = img + (size_t)(yok ? y : 0) * L.pitch;
#pragma unroll
      for (int sx = lane; sx < FAST_TW + 8; sx += 32) {
        const int x = tx0 - 4 + sx;
        s_src[sy][sx] = (yok && x >= 0 && x < L.w) ? __ldg(row + x) : (uint8_t)0;
      }
    }
  }
  __syncthreads();
  // pass 1a: cheap necessary condition over the ring; survivors go to a list so that the full
  // test below runs with all lanes busy.  Nine contiguous circle pixels contain four cyclically
  // consecutive ones of the eight even-numbered pixels: test that on |d| > t masks of each sign.
  const int t = g.fast_thr;
  for (int cy = warp; cy < CH; cy += 8) {
    const int y = ty0 - 1 + cy;
    const bool yok = y >= 3 && y < L.h - 3;
#pragma unroll
    for (int cx = lane; cx < CP; cx += 32) {
      const int x = tx0 - 1 + cx;
      s_sc[cy][cx] = 0;
      if (!(yok && x >= 3 && x < L.w - 3)) continue;
      const uint8_t* c = &s_src[cy + 3][cx + 3 + al];
      const int hi = c[0] + t, lo = c[0] - t;
      // opposite pairs first: flat regions (whole warps) leave after 5 loads
      const int p0 = c[3 * SP], p8 = c[-3 * SP], p4 = c[3], p12 = c[-3];
      const bool v_any = p0 > hi || p0 < lo || p8 > hi || p8 < lo;
      const bool h_any = p4 > hi || p4 < lo || p12 > hi || p12 < lo;
      if (!(v_any && h_any)) continue;
      const int p2 = c[2 * SP + 2], p6 = c[-2 * SP + 2], p10 = c[-2 * SP - 2], p14 = c[2 * SP - 2];
      const uint32_t mb = (uint32_t)(p0 < lo) | (uint32_t)(p2 < lo) << 1 | (uint32_t)(p4 < lo) << 2 |
                          (uint32_t)(p6 < lo) << 3 | (uint32_t)(p8 < lo) << 4 | (uint32_t)(p10 < lo) << 5 |
                          (uint32_t)(p12 < lo) << 6 | (uint32_t)(p14 < lo) << 7;
      const uint32_t md = (uint32_t)(p0 > hi) | (uint32_t)(p2 > hi) << 1 | (uint32_t)(p4 > hi) << 2 |
                          (uint32_t)(p6 > hi) << 3 | (uint32_t)(p8 > hi) << 4 | (uint32_t)(p10 > hi) << 5 |
                          (uint32_t)(p12 > hi) << 6 | (uint32_t)(p14 > hi) << 7;
      auto run4 = [](uint32_t m) {
        const uint32_t x8 = m | (m << 8);
        const uint32_t r2 = x8 & (x8 >> 1);
        return (r2 & (r2 >> 2) & 0xffu) != 0u;
      };
      if (run4(mb) || run4(md)) s_cand[atomicAdd(&s_nc, 1)] = (uint16_t)(cy * CP + cx);
    }
  }
  __syncthreads();
  // pass 1b: the full run-of-9 test on the listed pixels; corners go to a second list
  for (int k = threadIdx.x; k < s_nc; k += 256) {
    const int i = s_cand[k];
    const int cy = i / CP, cx = i - cy * CP;
    int d[16];
    fast_circle(&s_src[cy + 3][cx + 3 + al], SP, d);
    if (fast_corner(d, t)) s_corner[atomicAdd(&s_nk, 1)] = (uint16_t)i;
  }
  __syncthreads();
  // pass 2: scores of the corners only (about 1.5 % of the pixels)
  for (int k = threadIdx.x; k < s_nk; k += 256) {
    const int i = s_corner[k];
    const int cy = i / CP, cx = i - cy * CP;
    int d[16];
    fast_circle(&s_src[cy + 3][cx + 3 + al], SP, d);
    s_sc[cy][cx] = (uint8_t)fast_raw_score(d, t);
  }
  __syncthreads();
  // test hook (vo_orb_fast_score_map): cv2's response is raw - 1, 0 = not a corner
  if (score_map) {
    for (int py = warp; py < FAST_TH; py += 8)
      for (int px = lane; px < FAST_TW; px += 32)
        if (tx0 + px < L.w && ty0 + py < L.h)
          score_map[((size_t)f * L.h + ty0 + py) * L.w + tx0 + px] = s_sc[py + 1][px + 1];
  }
  // pass 3: strict 3x3 maxima inside the border
  const bool border_ok = !(g.edge > 0 && (L.h <= 2 * g.edge || L.w <= 2 * g.edge));
  for (int py = warp; py < FAST_TH; py += 8) {
    const int y = ty0 + py;
    if (!border_ok || y < g.edge || y >= L.h - g.edge) continue;
#pragma unroll
    for (int px = lane; px < FAST_TW; px += 32) {
      const int x = tx0 + px;
      const int s = s_sc[py + 1][px + 1];
      if (s == 0) continue;
      if (x < g.edge || x >= L.w - g.edge) continue;
      const bool is_max = s > s_sc[py][px] && s > s_sc[py][px + 1] && s > s_sc[py][px + 2] &&
                          s > s_sc[py + 1][px] && s > s_sc[py + 1][px + 2] && s > s_sc[py + 2][px] &&
                          s > s_sc[py + 2][px + 1] && s > s_sc[py + 2][px + 2];
      if (!is_max) continue;
      const int k = atomicAdd(&s_n, 1);
      s_out[k] = ((uint32_t)y << 20) | ((uint32_t)x << 8) | (uint32_t)s;
      atomicAdd(&hist[((size_t)f * g.nlevels + l) * 256 + s], 1);
    }
  }
  __syncthreads();
  if (s_n == 0) return;
  if (threadIdx.x == 0) s_base = atomicAdd(&cnt1[f * g.nlevels + l], s_n);
  __syncthreads();
  uint32_t* out = list1 + (size_t)f * g.list_stride + L.list_off;
  for (int i = threadIdx.x; i < s_n; i += 256)
    if (s_base + i < L.cap) out[s_base + i] = s_out[i];   // cap = (w/2+1)(h/2+1): cannot overflow
}

// ------------------------------------------------------------------ first cut: FAST score
// KeyPointsFilter::retainBest(2 n): everything >= the (2n)-th best score when there are more
__global__ void __launch_bounds__(256)
orb_cut1_kernel(OrbGeom g, const uint32_t* __restrict__ list1, const int32_t* __restrict__ cnt1,
                const int32_t* __restrict__ hist, uint32_t* __restrict__ list2,
                int32_t* __restrict__ cnt2) {
  __shared__ int s_thr, s_n;
  const int l = blockIdx.x, f = blockIdx.y;
  const OrbLevel L = g.lv[l];
  const int m = min(cnt1[f * g.nlevels + l], L.cap);
  const int want = 2 * L.nfeat;
  if (threadIdx.x == 0) {
    int thr = 0;
    if (want == 0) thr = 256;
    else if (m > want) {
      const int32_t* h = hist + ((size_t)f * g.nlevels + l) * 256;
      int cum = 0;
      for (thr = 255; thr > 0; --thr) {
        cum += h[thr];
        if (cum >= want) break;
      }
    }
    s_thr = thr;
    s_n = 0;
  }
  __syncthreads();
  const uint32_t* in = list1 + (size_t)f * g.list_stride + L.list_off;
  uint32_t* out = list2 + (size_t)f * g.list_stride + L.list_off;
  const int thr = s_thr;
  for (int i = threadIdx.x; i < m; i += 256) {
    const uint32_t c = in[i];
    if ((int)(c & 255u) >= thr) out[atomicAdd(&s_n, 1)] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) cnt2[f * g.nlevels + l] = s_n;
}

// ------------------------------------------------------------------ Harris response
// orb.cpp HarrisResponses: 7x7 block, Ix/Iy = 3x3 Sobel on the raw level, int sums, float formula
__global__ void __launch_bounds__(256)
orb_harris_kernel(OrbGeom g, const uint8_t* __restrict__ images, const uint8_t* __restrict__ pyr,
                  const uint32_t* __restrict__ list2, const int32_t* __restrict__ cnt2,
                  float* __restrict__ resp) {
  const int l = blockIdx.y, f = blockIdx.z;
  const OrbLevel L = g.lv[l];
  const int m = cnt2[f * g.nlevels + l];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint8_t* img = orb_level_ptr(g, l, f, images, pyr);
  const float scale = __fdiv_rn(1.f, __fmul_rn(28.f, 255.f));
  const float scale4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
  for (int i = blockIdx.x * 8 + warp; i < m; i += gridDim.x * 8) {
    const uint32_t c = list2[(size_t)f * g.list_stride + L.list_off + i];
    const int x0 = (c >> 8) & 0xfff, y0 = c >> 20;
    int a = 0, b = 0, cc = 0;
    for (int k = lane; k < 49; k += 32) {
      const int yy = y0 - 3 + k / 7, xx = x0 - 3 + k % 7;
      // clamped reads: only reached when edgeThreshold < 4 (cv2 reads its reflected border there)
      auto px = [&](int dx, int dy) {
        const int X = min(max(xx + dx, 0), L.w - 1), Y = min(max(yy + dy, 0), L.h - 1);
        return (int)__ldg(img + (size_t)Y * L.pitch + X);
      };
      const int ix = (px(1, 0) - px(-1, 0)) * 2 + (px(1, -1) - px(-1, -1)) + (px(1, 1) - px(-1, 1));
      const int iy = (px(0, 1) - px(0, -1)) * 2 + (px(-1, 1) - px(-1, -1)) + (px(1, 1) - px(1, -1));
      a += ix * ix; b += iy * iy; cc += ix * iy;
    }
    a = __reduce_add_sync(0xffffffffu, a);
    b = __reduce_add_sync(0xffffffffu, b);
    cc = __reduce_add_sync(0xffffffffu, cc);
    if (lane == 0) {
      const float fa = (float)a, fb = (float)b, fc = (float)cc;
      const float s = __fadd_rn(fa, fb);
      const float r = __fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)),
                                __fmul_rn(__fmul_rn(0.04f, s), s));
      resp[(size_t)f * g.list_stride + L.list_off + i] = __fmul_rn(r, scale4);
    }
  }
}

// ------------------------------------------------------------------ second cut + ordering
// One block per (level, frame).  keep: retainBest(n) on the Harris response (kept: fewer than n
// candidates are strictly better; ties at the cut survive, as in OpenCV).  place: rank of a
// survivor = number of survivors of its level with a smaller (y, x) word, after the survivors of
// the lower levels.  Record: x, y (scaled), size, angle (filled later), response, octave.
__global__ void __launch_bounds__(256)
orb_keep_kernel(OrbGeom g, const int32_t* __restrict__ cnt2, const float* __restrict__ resp,
                uint8_t* __restrict__ keepflag, int32_t* __restrict__ cnt3) {
  __shared__ int s_kept;
  const int l = blockIdx.x, f = blockIdx.y;
  const OrbLevel L = g.lv[l];
  const int m = cnt2[f * g.nlevels + l];
  const float* r = resp + (size_t)f * g.list_stride + L.list_off;
  uint8_t* keep = keepflag + (size_t)f * g.list_stride + L.list_off;
  __shared__ float s_r[ORB_STAGE];
  if (threadIdx.x == 0) s_kept = 0;
  const bool staged = m <= ORB_STAGE;
  if (staged)
    for (int i = threadIdx.x; i < m; i += 256) s_r[i] = r[i];
  __syncthreads();
  const float* rr = staged ? s_r : r;     // every lane reads the same j: shared-memory broadcast
  int mine = 0;
  for (int i = threadIdx.x; i < m; i += 256) {
    int better = 0;
    if (m > L.nfeat) {
      const float ri = rr[i];
      int j = 0;
      for (; j + 4 <= m; j += 4)
        better += (rr[j] > ri) + (rr[j + 1] > ri) + (rr[j + 2] > ri) + (rr[j + 3] > ri);
      for (; j < m; ++j) better += (rr[j] > ri);
    }
    const bool k = better < L.nfeat;
    keep[i] = k;
    mine += k;
  }
  if (mine) atomicAdd(&s_kept, mine);
  __syncthreads();
  if (threadIdx.x == 0) cnt3[f * g.nlevels + l] = s_kept;
}

__global__ void __launch_bounds__(256)
orb_place_kernel(OrbGeom g, const uint32_t* __restrict__ list2, const int32_t* __restrict__ cnt2,
                 const float* __restrict__ resp, const uint8_t* __restrict__ keepflag,
                 const int32_t* __restrict__ cnt3, float* __restrict__ kp,
                 int32_t* __restrict__ kp_level_xy, int32_t* __restrict__ n_out,
                 int32_t* __restrict__ status) {
  const int l = blockIdx.x, f = blockIdx.y;
  const OrbLevel L = g.lv[l];
  int base = 0, total = 0;
  for (int k = 0; k < g.nlevels; ++k) {
    const int c = cnt3[f * g.nlevels + k];
    if (k < l) base += c;
    total += c;
  }
  if (l == 0 && threadIdx.x == 0) {
    n_out[f] = min(total, g.cap_out);
    status[f] = total > g.cap_out ? 1 : 0;
  }
  const int m = cnt2[f * g.nlevels + l];
  const uint32_t* cand = list2 + (size_t)f * g.list_stride + L.list_off;
  const float* r = resp + (size_t)f * g.list_stride + L.list_off;
  const uint8_t* keep = keepflag + (size_t)f * g.list_stride + L.list_off;
  __shared__ uint32_t s_w[ORB_STAGE];
  __shared__ uint16_t s_i[ORB_STAGE];
  __shared__ int s_nk;
  const int kept = cnt3[f * g.nlevels + l];
  const bool staged = kept <= ORB_STAGE && m <= 65535;
  if (threadIdx.x == 0) s_nk = 0;
  __syncthreads();
  if (staged) {
    for (int i = threadIdx.x; i < m; i += 256)
      if (keep[i]) {
        const int k = atomicAdd(&s_nk, 1);
        s_w[k] = cand[i] >> 8;
        s_i[k] = (uint16_t)i;
      }
  }
  __syncthreads();
  const int n_items = staged ? kept : m;
  for (int k = threadIdx.x; k < n_items; k += 256) {
    int i, rank = 0;
    uint32_t ci;
    if (staged) {
      i = s_i[k];
      ci = s_w[k];
      int j = 0;
      for (; j + 4 <= kept; j += 4)
        rank += (s_w[j] < ci) + (s_w[j + 1] < ci) + (s_w[j + 2] < ci) + (s_w[j + 3] < ci);
      for (; j < kept; ++j) rank += (s_w[j] < ci);
    } else {
      i = k;
      if (!keep[i]) continue;
      ci = cand[i] >> 8;
      for (int j = 0; j < m; ++j) rank += (keep[j] && (cand[j] >> 8) < ci);
    }
    const int o = base + rank;
    if (o < g.cap_out) {
      const int x = ci & 0xfff, y = ci >> 12;
      float* rec = kp + ((size_t)f * g.cap_out + o) * 6;
      rec[0] = __fmul_rn((float)x, L.scale);
      rec[1] = __fmul_rn((float)y, L.scale);
      rec[2] = __fmul_rn(31.f, L.scale);
      rec[3] = -1.f;
      rec[4] = r[i];
      rec[5] = (float)l;
      int32_t* q = kp_level_xy + ((size_t)f * g.cap_out + o) * 2;
      q[0] = x; q[1] = y;
    }
  }
}

// ------------------------------------------------------------------ orientation + rBRIEF
__device__ __forceinline__ float orb_fast_atan2(float y, float x) {
  // cv::fastAtan2: degrees, plain float multiplies and adds (no contraction)
  const float p1 = 0x1.ca44dep+5f, p3 = -0x1.2aaddcp+4f, p5 = 0x1.1d3f7ep+3f, p7 = -0x1.4515b2p+1f;
  const float eps = 0x1p-52f;
  const float ax = fabsf(x), ay = fabsf(y);
  const float hi = fmaxf(ax, ay), lo = fminf(ax, ay);
  const float c = __fdiv_rn(lo, __fadd_rn(hi, eps));
  const float c2 = __fmul_rn(c, c);
  float a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  if (!(ax >= ay)) a = __fsub_rn(90.f, a);
  if (x < 0.f) a = __fsub_rn(180.f, a);
  if (y < 0.f) a = __fsub_rn(360.f, a);
  return a;
}

__device__ __forceinline__ int orb_reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return min(max(i, 0), n - 1);
}

__global__ void __launch_bounds__(256)
orb_describe_kernel(OrbGeom g, const uint8_t* __restrict__ images, const uint8_t* __restrict__ pyr,
                    float* __restrict__ kp, const int32_t* __restrict__ kp_level_xy,
                    const int32_t* __restrict__ n_out, uint8_t* __restrict__ desc) {
  __shared__ uint8_t s_src[ORB_SW][ORB_SW + 1];
  __shared__ float s_srcf[ORB_SW][ORB_SW + 2];
  __shared__ float s_row[ORB_SW][ORB_BW + 2];
  __shared__ uint8_t s_blur[ORB_BW][ORB_BW + 3];
  __shared__ float s_ab[2];
  const int f = blockIdx.y, i = blockIdx.x;
  if (i >= n_out[f]) return;
  float* rec = kp + ((size_t)f * g.cap_out + i) * 6;
  const int l = (int)rec[5];
  const OrbLevel L = g.lv[l];
  const int x0 = kp_level_xy[((size_t)f * g.cap_out + i) * 2], y0 = kp_level_xy[((size_t)f * g.cap_out + i) * 2 + 1];
  const uint8_t* img = orb_level_ptr(g, l, f, images, pyr);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool inside = x0 >= ORB_SR && x0 + ORB_SR < L.w && y0 >= ORB_SR && y0 + ORB_SR < L.h;
  for (int sy = warp; sy < ORB_SW; sy += 8) {
    const int y = inside ? y0 - ORB_SR + sy : orb_reflect101(y0 - ORB_SR + sy, L.h);
    const uint8_t* row = img + (size_t)y * L.pitch;
#pragma unroll
    for (int sx = lane; sx < ORB_SW; sx += 32) {
      const int x = inside ? x0 - ORB_SR + sx : orb_reflect101(x0 - ORB_SR + sx, L.w);
      const uint8_t p = __ldg(row + x);
      s_src[sy][sx] = p;
      s_srcf[sy][sx] = (float)p;
    }
  }
  __syncthreads();
  // --- orientation: warp 7 (idle during the row pass below, which occupies threads 0..187),
  // lane = row of the disc (31 rows)
  if (warp == 7) {
    int m01 = 0, m10 = 0;
    if (lane < 2 * ORB_HALF_PATCH + 1) {
      const int v = lane - ORB_HALF_PATCH;
      const int d = c_orb_umax[abs(v)];
      int sum = 0;
      for (int u = -d; u <= d; ++u) {
        const int p = s_src[ORB_SR + v][ORB_SR + u];
        sum += p;
        m10 += u * p;
      }
      m01 = v * sum;
    }
    m01 = __reduce_add_sync(0xffffffffu, m01);
    m10 = __reduce_add_sync(0xffffffffu, m10);
    if (lane == 0) {
      const float a = orb_fast_atan2((float)m01, (float)m10);
      rec[3] = a;
      // angle *= (float)(CV_PI / 180.f); a = (float)cos(angle), b = (float)sin(angle)  (orb.cpp)
      const float rad = __fmul_rn(a, (float)(3.14159265358979323846 / 180.0));
      s_ab[0] = (float)cos((double)rad);
      s_ab[1] = (float)sin((double)rad);
    }
  }
  // --- blur, row pass: k0 p0, then fused multiply-adds in tap order.  A thread owns 11
  // consecutive outputs of one row (47 rows x 4 runs = 188 tasks): 17 loads feed 77 multiply-adds
  if (threadIdx.x < ORB_SW * 4) {
    const int sy = threadIdx.x >> 2, b0 = (threadIdx.x & 3) * 11;
    float p[17];
#pragma unroll
    for (int k = 0; k < 17; ++k) p[k] = (b0 + k < ORB_SW) ? s_srcf[sy][b0 + k] : 0.f;
#pragma unroll
    for (int o = 0; o < 11; ++o) {
      float h = __fmul_rn(c_orb_taps[0], p[o]);
#pragma unroll
      for (int k = 1; k < 7; ++k) h = __fmaf_rn(p[o + k], c_orb_taps[k], h);
      if (b0 + o < ORB_BW) s_row[sy][b0 + o] = h;
    }
  }
  __syncthreads();
  // --- column pass, symmetric form, round half to even, saturate.  A thread owns 7 consecutive
  // rows of one column (41 columns x 6 runs = 246 tasks)
  if (threadIdx.x < ORB_BW * 6) {
    const int run = threadIdx.x / ORB_BW, bx = threadIdx.x - run * ORB_BW, r0 = run * 7;
    float hcol[13];
#pragma unroll
    for (int k = 0; k < 13; ++k) hcol[k] = (r0 + k < ORB_SW) ? s_row[r0 + k][bx] : 0.f;
#pragma unroll
    for (int o = 0; o < 7; ++o) {
      float v = __fmul_rn(c_orb_taps[3], hcol[o + 3]);
#pragma unroll
      for (int k = 1; k <= 3; ++k)
        v = __fmaf_rn(__fadd_rn(hcol[o + 3 + k], hcol[o + 3 - k]), c_orb_taps[3 + k], v);
      const int q = __float2int_rn(v);
      if (r0 + o < ORB_BW) s_blur[r0 + o][bx] = (uint8_t)min(max(q, 0), 255);
    }
  }
  __syncthreads();
  // --- 256 comparisons
  const float a = s_ab[0], b = s_ab[1];
  const int bit = threadIdx.x;
  const float pax = (float)c_orb_pattern[bit][0], pay = (float)c_orb_pattern[bit][1];
  const float pbx = (float)c_orb_pattern[bit][2], pby = (float)c_orb_pattern[bit][3];
  const int xa = __float2int_rn(__fsub_rn(__fmul_rn(pax, a), __fmul_rn(pay, b)));
  const int ya = __float2int_rn(__fadd_rn(__fmul_rn(pax, b), __fmul_rn(pay, a)));
  const int xb = __float2int_rn(__fsub_rn(__fmul_rn(pbx, a), __fmul_rn(pby, b)));
  const int yb = __float2int_rn(__fadd_rn(__fmul_rn(pbx, b), __fmul_rn(pby, a)));
  const bool lt = s_blur[ORB_BR + ya][ORB_BR + xa] < s_blur[ORB_BR + yb][ORB_BR + xb];
  const uint32_t word = __ballot_sync(0xffffffffu, lt);
  if ((threadIdx.x & 31) == 0)
    reinterpret_cast<uint32_t*>(desc + ((size_t)f * g.cap_out + i) * 32)[threadIdx.x >> 5] = word;
}

// ------------------------------------------------------------------ host side
struct OrbWs {
  int2* coef;
  uint8_t* pyr;
  uint32_t *list1, *list2;
  uint8_t* keep;
  float* resp;
  int32_t *cnt1, *cnt2, *cnt3, *hist, *level_xy;
  size_t bytes;
};

static bool orb_geometry(OrbGeom& g, int n_frames, int width, int height, int nfeatures,
                         float scale_factor, int nlevels, int edge, int fast_thr, int cap_out) {
  if (nlevels < 1 || nlevels > ORB_MAX_LEVELS || width < 1 || height < 1 || width > 4095 ||
      height > 4095 || nfeatures < 0 || !(scale_factor > 1.f) || n_frames < 0)
    return false;
  g.nlevels = nlevels; g.n_frames = n_frames; g.width = width; g.height = height;
  g.edge = edge; g.fast_thr = fast_thr < 0 ? 0 : (fast_thr > 255 ? 255 : fast_thr); g.cap_out = cap_out;
  // ORB::create takes scaleFactor as float and keeps it in a double (orb.cpp)
  const double sf = (double)scale_factor;
  const float factor = (float)(1.0 / sf);
  volatile float nd = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
  int sum = 0;
  size_t pix = 0, list = 0, coef = 0;
  for (int l = 0; l < nlevels; ++l) {
    OrbLevel& L = g.lv[l];
    L.scale = (float)pow(sf, (double)l);
    L.w = (int)lrintf((float)width / L.scale);
    L.h = (int)lrintf((float)height / L.scale);
    if (L.w < 1 || L.h < 1) return false;
    if (l < nlevels - 1) {
      L.nfeat = (int)lrintf(nd);
      sum += L.nfeat;
      nd = nd * factor;
    } else {
      L.nfeat = nfeatures - sum > 0 ? nfeatures - sum : 0;
    }
    L.cap = (L.w / 2 + 1) * (L.h / 2 + 1);   // strict 3x3 maxima: at most one per 2x2 cell
    L.pix_off = pix;
    L.pitch = l == 0 ? width : (L.w + 15) & ~15;
    if (l > 0) pix += ((size_t)L.pitch * L.h + 255) & ~(size_t)255;
    L.list_off = list;
    list += ((size_t)L.cap + 63) & ~(size_t)63;
    L.coef_off = coef;
    if (l > 0) coef += (size_t)L.w + L.h;
  }
  g.pyr_stride = pix;
  g.list_stride = list;
  return true;
}

static OrbWs orb_carve(void* base, const OrbGeom& g) {
  OrbWs w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += (bytes + 1023) & ~(size_t)1023;
    return p;
  };
  const OrbLevel& last = g.lv[g.nlevels - 1];
  const size_t ncoef = last.coef_off + (g.nlevels > 1 ? (size_t)last.w + last.h : 0);
  const size_t F = (size_t)g.n_frames;
  w.coef = (int2*)take((ncoef + 1) * sizeof(int2));
  w.pyr = (uint8_t*)take(F * g.pyr_stride + 16);
  w.list1 = (uint32_t*)take(F * g.list_stride * 4);
  w.list2 = (uint32_t*)take(F * g.list_stride * 4);
  w.keep = (uint8_t*)take(F * g.list_stride);
  // the Harris responses reuse the memory of the first corner list, which is dead after cut1
  w.resp = (float*)w.list1;
  // counters and histogram are contiguous: one memset
  w.cnt1 = (int32_t*)take(F * g.nlevels * (3 + 256) * sizeof(int32_t));
  w.cnt2 = w.cnt1 + F * g.nlevels;
  w.cnt3 = w.cnt2 + F * g.nlevels;
  w.hist = w.cnt3 + F * g.nlevels;
  w.level_xy = (int32_t*)take(F * (size_t)(g.cap_out > 0 ? g.cap_out : 0) * 2 * sizeof(int32_t));
  w.bytes = off;
  return w;
}

}  // namespace vo

using namespace vo;

extern "C" size_t vo_orb_workspace_bytes(int n_frames, int width, int height, int nfeatures,
                                         float scale_factor, int nlevels, int cap_out) {
  OrbGeom g;
  if (!orb_geometry(g, n_frames, width, height, nfeatures, scale_factor, nlevels, 31, 20, cap_out))
    return 0;
  return orb_carve(nullptr, g).bytes;
}

extern "C" int vo_orb_level_info(int width, int height, int nfeatures, float scale_factor,
                                 int nlevels, int32_t* w_h, int32_t* h_h, int32_t* nfeat_h,
                                 float* scale_h) {
  OrbGeom g;
  VO_REQUIRE(orb_geometry(g, 1, width, height, nfeatures, scale_factor, nlevels, 31, 20, 1),
             VO_ERR_INVALID_ARG, "vo_orb_level_info: bad geometry");
  for (int l = 0; l < nlevels; ++l) {
    if (w_h) w_h[l] = g.lv[l].w;
    if (h_h) h_h[l] = g.lv[l].h;
    if (nfeat_h) nfeat_h[l] = g.lv[l].nfeat;
    if (scale_h) scale_h[l] = g.lv[l].scale;
  }
  return VO_OK;
}

extern "C" int vo_orb_detect_batch(const uint8_t* images, int n_frames, int width, int height,
                                   int nfeatures, float scale_factor, int nlevels,
                                   int edge_threshold, int fast_threshold, int cap_out, float* kp,
                                   uint8_t* desc, int32_t* n_out, int32_t* status, void* ws,
                                   size_t ws_bytes, void* stream) {
  OrbGeom g;
  VO_REQUIRE(orb_geometry(g, n_frames, width, height, nfeatures, scale_factor, nlevels,
                          edge_threshold, fast_threshold, cap_out),
             VO_ERR_INVALID_ARG,
             "vo_orb_detect_batch: bad geometry (1 <= nlevels <= 16, size <= 4095, scale > 1)");
  VO_REQUIRE(edge_threshold >= 0 && cap_out >= 1 && n_frames <= 65535, VO_ERR_INVALID_ARG,
             "vo_orb_detect_batch: edge_threshold >= 0, cap_out >= 1 and at most 65535 frames per call");
  if (n_frames == 0) return VO_OK;
  VO_REQUIRE(images && kp && desc && n_out && status && ws, VO_ERR_INVALID_ARG,
             "vo_orb_detect_batch: null pointer");
  const OrbWs w = orb_carve(ws, g);
  VO_REQUIRE(ws_bytes >= w.bytes, VO_ERR_WORKSPACE, "vo_orb_detect_batch: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int F = n_frames;
  cudaMemsetAsync(w.cnt1, 0, (size_t)F * nlevels * (3 + 256) * sizeof(int32_t), st);
  if (nlevels > 1) orb_coef_kernel<<<dim3(8, nlevels - 1), 256, 0, st>>>(g, w.coef);
  for (int l = 0; l < nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    if (l > 0)
      orb_resize_kernel<<<dim3((L.w + 511) / 512, L.h, F), 128, 0, st>>>(g, l, images, w.pyr, w.coef);
    orb_fast_kernel<<<dim3((L.w + FAST_TW - 1) / FAST_TW, (L.h + FAST_TH - 1) / FAST_TH, F), 256, 0, st>>>(
        g, l, images, w.pyr, w.list1, w.cnt1, w.hist, nullptr);
  }
  orb_cut1_kernel<<<dim3(nlevels, F), 256, 0, st>>>(g, w.list1, w.cnt1, w.hist, w.list2, w.cnt2);
  orb_harris_kernel<<<dim3(64, nlevels, F), 256, 0, st>>>(g, images, w.pyr, w.list2, w.cnt2, w.resp);
  orb_keep_kernel<<<dim3(nlevels, F), 256, 0, st>>>(g, w.cnt2, w.resp, w.keep, w.cnt3);
  orb_place_kernel<<<dim3(nlevels, F), 256, 0, st>>>(g, w.list2, w.cnt2, w.resp, w.keep, w.cnt3, kp, w.level_xy, n_out, status);
  orb_describe_kernel<<<dim3(cap_out, F), 256, 0, st>>>(g, images, w.pyr, kp, w.level_xy, n_out, desc);
  return check_launch("vo_orb_detect_batch");
}

// Test hook: the FAST stage alone on level 0.  scores uint8 [n_frames, height, width] = score + 1
// of cv2.FastFeatureDetector (0 = not a corner), before non-maximum suppression; corners uint32
// [n_frames, (w/2+1)(h/2+1)] = packed (y << 20 | x << 8 | score + 1) survivors of the suppression
// and of the border filter, unordered; n_corners int32 [n_frames] (zeroed by the call).
extern "C" int vo_orb_fast_score_map(const uint8_t* images, int n_frames, int width, int height,
                                     int fast_threshold, int edge_threshold, uint8_t* scores,
                                     uint32_t* corners, int32_t* n_corners, int32_t* hist,
                                     void* stream) {
  OrbGeom g;
  VO_REQUIRE(orb_geometry(g, n_frames, width, height, 1, 1.2f, 1, edge_threshold, fast_threshold, 1),
             VO_ERR_INVALID_ARG, "vo_orb_fast_score_map: bad geometry");
  VO_REQUIRE(images && scores && corners && n_corners && hist, VO_ERR_INVALID_ARG,
             "vo_orb_fast_score_map: null pointer");
  if (n_frames == 0) return VO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(n_corners, 0, (size_t)n_frames * sizeof(int32_t), st);
  cudaMemsetAsync(hist, 0, (size_t)n_frames * 256 * sizeof(int32_t), st);
  cudaMemsetAsync(scores, 0, (size_t)n_frames * width * height, st);
  orb_fast_kernel<<<dim3((width + FAST_TW - 1) / FAST_TW, (height + FAST_TH - 1) / FAST_TH, n_frames), 256, 0, st>>>(
      g, 0, images, nullptr, corners, n_corners, hist, scores);
  return check_launch("vo_orb_fast_score_map");
}
