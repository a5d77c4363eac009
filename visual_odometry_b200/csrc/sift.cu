// sift.cu — N4: SIFT feature extraction for a batch of equally sized frames, the
// detector the reference builds at /root/reference/src/visual_slam.py:17
// (`cv2.SIFT_create()`) and calls at /root/reference/src/frame_generator.py:25
// (`detectAndCompute(frame.image, None)`).  The algorithm is OpenCV's (features2d sift.dispatch.cpp
// / sift.simd.hpp; restated on the CPU in oracle/cv_sift.py, which is pinned against the cv2
// binary): base image doubled, sigma 1.6, 3 layers per octave, contrast 0.04, edge 10,
// nfeatures 0 (all key points).
//
// Compiled with -fmad=false (Makefile EXTRA_sift): the Gaussian pyramid is BIT-exact against
// cv2.GaussianBlur on an AVX2/FMA host because the fused and the separately rounded operations
// are written out with intrinsics exactly where cv2's vector loops and scalar tails have them.
//
// Layout in HBM (workspace): per frame one Gaussian pyramid, octave after octave, 6 layers of
// h_o x w_o floats each; DoG values are differences of two layers computed where they are used
// (no DoG pyramid is stored).  Key points leave as float32 [F, cap, 6] = (x, y, size, angle,
// response, octave word) in cv2's order (KeyPoint12_LessThan, duplicates removed) and
// descriptors as uint8 [F, cap, 128] (cv2 stores the same integers as float32).
#include <math.h>

#include "common.cuh"

namespace vo {
namespace sift {

constexpr int NL = 3;                 // nOctaveLayers
constexpr int LAYERS = NL + 3;        // Gaussian layers per octave
constexpr int MAX_OCT = 16;
constexpr int BORDER = 5;             // SIFT_IMG_BORDER
constexpr int MAX_R = 15;             // largest Gaussian half width (sigma 3.09 -> 13)
constexpr int ORI_BINS = 36;
constexpr int CAND_WORDS = 10;

struct Geom {
  int n_oct;
  int h[MAX_OCT], w[MAX_OCT];
  long long off[MAX_OCT];             // float offset of octave o inside one frame's pyramid
  long long frame_stride;             // floats per frame
};

struct Kernels {                      // the six Gaussian kernels: [0] base blur, [1..5] layers
  int r[LAYERS];
  float f[LAYERS][2 * MAX_R + 1];     // full symmetric kernel, 2 r + 1 taps
};

static int make_geom(int width, int height, Geom* g) {
  // nOctaves = cvRound(log(min(cols, rows)) / log(2) - 2) - firstOctave, firstOctave = -1
  const double v = log((double)(width < height ? width : height)) / log(2.0) - 2.0;
  const int n_oct = (int)nearbyint(v) + 1;
  if (n_oct < 1 || n_oct > MAX_OCT) return -1;
  g->n_oct = n_oct;
  long long off = 0;
  int h = 2 * height, w = 2 * width;
  for (int o = 0; o < n_oct; ++o) {
    if (h < 1 || w < 1) return -1;
    g->h[o] = h; g->w[o] = w; g->off[o] = off;
    off += (long long)LAYERS * h * w;
    h /= 2; w /= 2;
  }
  g->frame_stride = off;
  return 0;
}

// getGaussianKernel(ksize = cvRound(sigma * 8 + 1) | 1, sigma, CV_32F): double arithmetic,
// normalised, stored as float
static int make_kernel(double sigma, int* r_out, float* f) {
  int ksize = (int)nearbyint(sigma * 8.0 + 1.0) | 1;
  const int r = ksize / 2;
  if (r > MAX_R) return -1;
  double k[2 * MAX_R + 1], sum = 0.0;
  for (int i = 0; i < ksize; ++i) {
    const double x = (double)(i - r);
    k[i] = exp(-(x * x) / (2.0 * sigma * sigma));
    sum += k[i];
  }
  for (int i = 0; i < ksize; ++i) f[i] = (float)(k[i] / sum);
  *r_out = r;
  return 0;
}

static int make_kernels(Kernels* K) {
  const float sigma = 1.6f, init = 0.5f;
  float d = sigma * sigma - init * init * 4.f;
  const float sig_diff = sqrtf(d > 0.01f ? d : 0.01f);
  if (make_kernel((double)sig_diff, &K->r[0], K->f[0])) return -1;
  const double kk = pow(2.0, 1.0 / NL);
  for (int i = 1; i < LAYERS; ++i) {
    const double prev = pow(kk, (double)(i - 1)) * 1.6;   // SIFT_Impl::sigma is a double here
    const double total = prev * kk;
    if (make_kernel(sqrt(total * total - prev * prev), &K->r[i], K->f[i])) return -1;
  }
  return 0;
}

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  const int period = 2 * (n - 1);
  i %= period;
  if (i < 0) i += period;
  return i >= n ? period - i : i;
}

// one fold covers every position a tile's halo can reach unless the image is smaller than the halo
__device__ __forceinline__ int reflect_once(int i, int n) {
  int j = i < 0 ? -i : i;
  j = j >= n ? 2 * (n - 1) - j : j;
  return (unsigned)j < (unsigned)n ? j : reflect101(i, n);
}

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src_gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem) : "memory");
}

// ------------------------------------------------------------------ S1: uint8 -> float, 2x linear
// resize(INTER_LINEAR) by exactly 2: taps (d + 0.5) / 2 - 0.5, weights 0.25 / 0.75, clamped; all
// products and sums are exact in float, so the order does not matter.
__global__ void upsample_kernel(const uint8_t* __restrict__ img, int h, int w, float* __restrict__ out,
                                long long out_stride) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  const int f = blockIdx.z;
  if (x >= 2 * w) return;
  const uint8_t* src = img + (size_t)f * h * w;
  auto taps = [](int d, int n, int& i0, int& i1, float& a) {
    const float fx = (d + 0.5f) * 0.5f - 0.5f;
    i0 = (int)floorf(fx);
    a = fx - i0;
    if (i0 < 0) { i0 = 0; a = 0.f; }
    if (i0 >= n - 1) { i1 = n - 1; i0 = i0 < n - 1 ? i0 : n - 1; a = 0.f; } else { i1 = i0 + 1; }
  };
  int x0, x1, y0, y1; float ax, ay;
  taps(x, w, x0, x1, ax);
  taps(y, h, y0, y1, ay);
  const float r0 = src[(size_t)y0 * w + x0] * (1.f - ax) + src[(size_t)y0 * w + x1] * ax;
  const float r1 = src[(size_t)y1 * w + x0] * (1.f - ax) + src[(size_t)y1 * w + x1] * ax;
  out[(size_t)f * out_stride + (size_t)y * 2 * w + x] = r0 * (1.f - ay) + r1 * ay;
}

// packed fp32 (sm_100: FFMA2 / FMUL2): two independent IEEE operations per instruction
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

// ------------------------------------------------------------------ S2: separable Gaussian blur
// One block = a 64 x 32 output tile: the row pass of the tile's rows plus r halo rows goes to
// shared memory, the column pass reads it.  Operation order = cv2's (see oracle/cv_sift.py blur):
// rows: k = 0 .. 2r, fused multiply-add, except the last w % 4 columns (multiply, then add);
// columns: f0 * x, then fma(x[-k] + x[+k], f[k], acc), except the last w % 8 columns.
constexpr int TW = 64, TH = 32;
// R = half width (5, 6, 8, 10, 13 for cv2's six sigmas).  The tile plus its halo is staged in
// shared memory once (coalesced, borders reflected there); the row pass keeps a window of 8 + 2R
// inputs in registers for 8 adjacent outputs of one row (lanes = rows, odd strides: no bank
// conflicts), the column pass does the same down a column (lanes = columns).  Blocks that contain
// columns of cv2's scalar tails take the per-output form of the same arithmetic.
template <int R>
__global__ void __launch_bounds__(256)
blur_kernel(const float* __restrict__ src, float* __restrict__ dst, int h, int w, long long stride,
            Kernels K, int ki) {
  constexpr int ROWS = TH + 2 * R, COLS = TW + 2 * R, IS = COLS | 1, RS = TW + 1;
  // rows 2p and 2p + 1 of the staged tile side by side: one 8-byte load is the operand pair of a
  // packed fp32 operation (FFMA2) that filters both rows at once
  __shared__ float2 in2[ROWS / 2][IS];
  __shared__ float rowp[ROWS][RS];
#define SIFT_IN(ry, cx) (reinterpret_cast<float*>(&in2[(ry) >> 1][cx])[(ry) & 1])
  __shared__ float fk[2 * R + 1];
  const int tid = threadIdx.x;
  if (tid <= 2 * R) fk[tid] = K.f[ki][tid];
  const float* s = src + (size_t)blockIdx.z * stride;
  float* d = dst + (size_t)blockIdx.z * stride;
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  const int row_tail = (w / 4) * 4, col_tail = (w / 8) * 8;
  const int x_end = x0 + TW < w ? x0 + TW : w;
  const bool tails = x_end > col_tail;            // col_tail <= row_tail
  {
    // staging: asynchronous 4-byte copies (no register round trip, all of a warp's rows in flight)
    const int lane = tid & 31, wrp = tid >> 5;
    const bool inside = x0 - R >= 0 && x0 + TW + R <= w;
    for (int ry = wrp; ry < ROWS; ry += 8) {
      const float* line = s + (size_t)reflect_once(y0 - R + ry, h) * w;
#pragma unroll
      for (int it = 0; it < (COLS + 31) / 32; ++it) {
        const int cx = lane + 32 * it;
        if (cx < COLS) {
          const int x = inside ? x0 - R + cx : reflect_once(x0 - R + cx, w);
          cp_async4(&SIFT_IN(ry, cx), line + x);
        }
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  __syncthreads();
  if (!tails) {
    {
      // row pass: warp g = columns 8g .. 8g + 7, lane = row pair; both rows of a pair go through
      // the same packed multiply-adds (each half rounds like the scalar operation)
      const int g = tid >> 5, rp = tid & 31;
      if (rp < ROWS / 2) {
        f32x2 win[8 + 2 * R], acc[8];
#pragma unroll
        for (int m = 0; m < 8 + 2 * R; ++m) win[m] = *reinterpret_cast<const f32x2*>(&in2[rp][8 * g + m]);
        const f32x2 f0 = pack2(fk[0], fk[0]);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = mul2(win[j], f0);
#pragma unroll
        for (int k = 1; k <= 2 * R; ++k) {
          const f32x2 f = pack2(fk[k], fk[k]);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = fma2(win[j + k], f, acc[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float lo, hi;
          unpack2(acc[j], lo, hi);
          rowp[2 * rp][8 * g + j] = lo;
          rowp[2 * rp + 1][8 * g + j] = hi;
        }
      }
    }
    __syncthreads();
    const int col = tid & 63, q = tid >> 6;
    float win[8 + 2 * R], acc[8];
#pragma unroll
    for (int m = 0; m < 8 + 2 * R; ++m) win[m] = rowp[8 * q + m][col];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = __fmul_rn(fk[R], win[j + R]);
#pragma unroll
    for (int k = 1; k <= R; ++k) {
      const float f = fk[R + k];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = __fmaf_rn(__fadd_rn(win[j + R - k], win[j + R + k]), f, acc[j]);
    }
    const int x = x0 + col;
    if (x < w) {
      float* dp = d + (size_t)(y0 + 8 * q) * w + x;
      const int left = h - (y0 + 8 * q);          // rows of this group inside the image
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < left) dp[(size_t)j * w] = acc[j];
    }
    return;
  }
  for (int i = tid; i < ROWS * TW; i += 256) {
    const int ry = i / TW, cx = i % TW;
    const int x = x0 + cx;
    if (x >= w) continue;
    float acc = __fmul_rn(SIFT_IN(ry, cx), fk[0]);
    if (x < row_tail) {
      for (int k = 1; k <= 2 * R; ++k) acc = __fmaf_rn(SIFT_IN(ry, cx + k), fk[k], acc);
    } else {
      for (int k = 1; k <= 2 * R; ++k) acc = __fadd_rn(__fmul_rn(SIFT_IN(ry, cx + k), fk[k]), acc);
    }
    rowp[ry][cx] = acc;
  }
  __syncthreads();
  for (int i = tid; i < TH * TW; i += 256) {
    const int ty = i / TW, cx = i % TW;
    const int x = x0 + cx, y = y0 + ty;
    if (x >= w || y >= h) continue;
    float acc = __fmul_rn(fk[R], rowp[ty + R][cx]);
    if (x < col_tail) {
      for (int k = 1; k <= R; ++k)
        acc = __fmaf_rn(__fadd_rn(rowp[ty + R - k][cx], rowp[ty + R + k][cx]), fk[R + k], acc);
    } else {
      for (int k = 1; k <= R; ++k)
        acc = __fadd_rn(__fmul_rn(__fadd_rn(rowp[ty + R - k][cx], rowp[ty + R + k][cx]), fk[R + k]), acc);
    }
    d[(size_t)y * w + x] = acc;
  }
}

#undef SIFT_IN

static int launch_blur(const float* src, float* dst, int h, int w, long long stride, const Kernels& K,
                       int ki, int n_frames, cudaStream_t st) {
  const dim3 grid((w + TW - 1) / TW, (h + TH - 1) / TH, n_frames);
  switch (K.r[ki]) {
    case 5: blur_kernel<5><<<grid, 256, 0, st>>>(src, dst, h, w, stride, K, ki); break;
    case 6: blur_kernel<6><<<grid, 256, 0, st>>>(src, dst, h, w, stride, K, ki); break;
    case 8: blur_kernel<8><<<grid, 256, 0, st>>>(src, dst, h, w, stride, K, ki); break;
    case 10: blur_kernel<10><<<grid, 256, 0, st>>>(src, dst, h, w, stride, K, ki); break;
    case 13: blur_kernel<13><<<grid, 256, 0, st>>>(src, dst, h, w, stride, K, ki); break;
    default: return -1;
  }
  return 0;
}

// ------------------------------------------------------------------ S3: next octave's first layer
// resize(INTER_NEAREST) to (w / 2, h / 2): sx = min(floor(x * (w / (w / 2))), w - 1) in double
__global__ void halve_kernel(const float* __restrict__ src, int h, int w, float* __restrict__ dst,
                             int dh, int dw, long long stride) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= dw) return;
  const double fx = (double)w / dw, fy = (double)h / dh;
  int sx = (int)floor(x * fx), sy = (int)floor(y * fy);
  sx = sx < w - 1 ? sx : w - 1;
  sy = sy < h - 1 ? sy : h - 1;
  dst[(size_t)blockIdx.z * stride + (size_t)y * dw + x] =
      src[(size_t)blockIdx.z * stride + (size_t)sy * w + sx];
}

// ------------------------------------------------------------------ S4: extrema + refinement
// DoG layer l (0 .. 4) of an octave = Gaussian layer l + 1 minus layer l, computed on the fly.
struct Octave {
  const float* g;       // Gaussian layer 0 of this octave for this frame
  int h, w;
  __device__ __forceinline__ float gauss(int layer, int r, int c) const {
    return g[(size_t)layer * h * w + (size_t)r * w + c];
  }
  __device__ __forceinline__ float dog(int layer, int r, int c) const {
    const size_t ls = (size_t)h * w, p = (size_t)r * w + c;
    return __fsub_rn(g[(layer + 1) * ls + p], g[layer * ls + p]);
  }
};

struct Cand {           // refined extremum, CAND_WORDS 32-bit words in the candidate list
  int o, layer, r, c;
  float x, y, size, response;
  int octave_word;
};

// Matx33f::solve(DECOMP_LU) = Cramer's rule in float; singular -> 0
__device__ __forceinline__ void solve3(const float a[3][3], const float b[3], float x[3]) {
  float d = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) -
            a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
            a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
  if (d == 0.f) { x[0] = x[1] = x[2] = 0.f; return; }
  d = 1.f / d;
  x[0] = d * (b[0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) -
              a[0][1] * (b[1] * a[2][2] - a[1][2] * b[2]) +
              a[0][2] * (b[1] * a[2][1] - a[1][1] * b[2]));
  x[1] = d * (a[0][0] * (b[1] * a[2][2] - a[1][2] * b[2]) -
              b[0] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
              a[0][2] * (a[1][0] * b[2] - b[1] * a[2][0]));
  x[2] = d * (a[0][0] * (a[1][1] * b[2] - b[1] * a[2][1]) -
              a[0][1] * (a[1][0] * b[2] - b[1] * a[2][0]) +
              b[0] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]));
}

// adjustLocalExtrema; `layer` is the DoG layer (1 .. 3)
__device__ bool adjust_local_extrema(const Octave& O, int octv, int layer, int r, int c, Cand* out) {
  const float img_scale = 1.f / 255.f;
  const float deriv_scale = img_scale * 0.5f;
  const float second_deriv_scale = img_scale;
  const float cross_deriv_scale = img_scale * 0.25f;
  float xi = 0.f, xr = 0.f, xc = 0.f;
  int i = 0;
  for (; i < 5; ++i) {
    const float dD[3] = {(O.dog(layer, r, c + 1) - O.dog(layer, r, c - 1)) * deriv_scale,
                         (O.dog(layer, r + 1, c) - O.dog(layer, r - 1, c)) * deriv_scale,
                         (O.dog(layer + 1, r, c) - O.dog(layer - 1, r, c)) * deriv_scale};
    const float v2 = O.dog(layer, r, c) * 2.f;
    const float dxx = (O.dog(layer, r, c + 1) + O.dog(layer, r, c - 1) - v2) * second_deriv_scale;
    const float dyy = (O.dog(layer, r + 1, c) + O.dog(layer, r - 1, c) - v2) * second_deriv_scale;
    const float dss = (O.dog(layer + 1, r, c) + O.dog(layer - 1, r, c) - v2) * second_deriv_scale;
    const float dxy = (O.dog(layer, r + 1, c + 1) - O.dog(layer, r + 1, c - 1) -
                       O.dog(layer, r - 1, c + 1) + O.dog(layer, r - 1, c - 1)) * cross_deriv_scale;
    const float dxs = (O.dog(layer + 1, r, c + 1) - O.dog(layer + 1, r, c - 1) -
                       O.dog(layer - 1, r, c + 1) + O.dog(layer - 1, r, c - 1)) * cross_deriv_scale;
    const float dys = (O.dog(layer + 1, r + 1, c) - O.dog(layer + 1, r - 1, c) -
                       O.dog(layer - 1, r + 1, c) + O.dog(layer - 1, r - 1, c)) * cross_deriv_scale;
    const float H[3][3] = {{dxx, dxy, dxs}, {dxy, dyy, dys}, {dxs, dys, dss}};
    float X[3];
    solve3(H, dD, X);
    xi = -X[2]; xr = -X[1]; xc = -X[0];
    if (fabsf(xi) < 0.5f && fabsf(xr) < 0.5f && fabsf(xc) < 0.5f) break;
    const float big = (float)(2147483647 / 3);
    if (fabsf(xi) > big || fabsf(xr) > big || fabsf(xc) > big) return false;
    c += __float2int_rn(xc);
    r += __float2int_rn(xr);
    layer += __float2int_rn(xi);
    if (layer < 1 || layer > NL || c < BORDER || c >= O.w - BORDER || r < BORDER || r >= O.h - BORDER)
      return false;
  }
  if (i >= 5) return false;
  const float dD[3] = {(O.dog(layer, r, c + 1) - O.dog(layer, r, c - 1)) * deriv_scale,
                       (O.dog(layer, r + 1, c) - O.dog(layer, r - 1, c)) * deriv_scale,
                       (O.dog(layer + 1, r, c) - O.dog(layer - 1, r, c)) * deriv_scale};
  const float t = dD[0] * xc + dD[1] * xr + dD[2] * xi;
  const float contr = O.dog(layer, r, c) * img_scale + t * 0.5f;
  if (fabsf(contr) * NL < 0.04f) return false;
  const float v2 = O.dog(layer, r, c) * 2.f;
  const float dxx = (O.dog(layer, r, c + 1) + O.dog(layer, r, c - 1) - v2) * second_deriv_scale;
  const float dyy = (O.dog(layer, r + 1, c) + O.dog(layer, r - 1, c) - v2) * second_deriv_scale;
  const float dxy = (O.dog(layer, r + 1, c + 1) - O.dog(layer, r + 1, c - 1) -
                     O.dog(layer, r - 1, c + 1) + O.dog(layer, r - 1, c - 1)) * cross_deriv_scale;
  const float tr = dxx + dyy, det = dxx * dyy - dxy * dxy;
  const float e = 10.f;
  if (det <= 0.f || tr * tr * e >= (e + 1.f) * (e + 1.f) * det) return false;
  out->o = octv; out->layer = layer; out->r = r; out->c = c;
  out->x = (c + xc) * (float)(1 << octv);
  out->y = (r + xr) * (float)(1 << octv);
  out->octave_word = octv + (layer << 8) + ((int)nearbyint(((double)xi + 0.5) * 255.0) << 16);
  out->size = 1.6f * powf(2.f, (layer + xi) / NL) * (float)(1 << octv) * 2.f;
  out->response = fabsf(contr);
  return true;
}

// one thread per pixel of one octave; all three DoG layers
__global__ void __launch_bounds__(256)
extrema_kernel(const float* __restrict__ pyr, Geom G, int octv, int cap, int32_t* __restrict__ n_cand,
               int32_t* __restrict__ cand, int32_t* __restrict__ status) {
  const int f = blockIdx.z;
  Octave O{pyr + (size_t)f * G.frame_stride + G.off[octv], G.h[octv], G.w[octv]};
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int r = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (c < BORDER || c >= O.w - BORDER || r < BORDER || r >= O.h - BORDER) return;
  // centre values of the five DoG layers first: most pixels fail the threshold on all three
  const size_t ls = (size_t)O.h * O.w, pc = (size_t)r * O.w + c;
  float gc[LAYERS];
#pragma unroll
  for (int l = 0; l < LAYERS; ++l) gc[l] = O.g[l * ls + pc];
  bool any = false;
#pragma unroll
  for (int l = 1; l <= NL; ++l) any |= fabsf(__fsub_rn(gc[l + 1], gc[l])) > 1.f;   // threshold = 1
  if (!any) return;
  float dc[LAYERS - 1];
#pragma unroll
  for (int l = 0; l < LAYERS - 1; ++l) dc[l] = __fsub_rn(gc[l + 1], gc[l]);
#pragma unroll
  for (int layer = 1; layer <= NL; ++layer) {
    const float val = dc[layer];
    if (!(fabsf(val) > 1.f)) continue;      // threshold = floor(0.5 * 0.04 / 3 * 255) = 1
    const bool pos = val > 0.f;
    // the 8 neighbours in the same DoG layer first (nine of ten candidates end here), then the
    // 9 + 9 of the layers below and above; the two Gaussian layers loaded here are reused there
    float lo[9], hi[9];
    bool ext = true;
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      const size_t p = pc + (q / 3 - 1) * O.w + (q % 3 - 1);
      lo[q] = O.g[layer * ls + p];
      hi[q] = O.g[(layer + 1) * ls + p];
      const float nb = __fsub_rn(hi[q], lo[q]);
      ext &= pos ? !(val < nb) : !(val > nb);
    }
    if (!ext) continue;
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      const size_t p = pc + (q / 3 - 1) * O.w + (q % 3 - 1);
      const float below = __fsub_rn(lo[q], O.g[(layer - 1) * ls + p]);
      const float above = __fsub_rn(O.g[(layer + 2) * ls + p], hi[q]);
      ext &= pos ? !(val < below) && !(val < above) : !(val > below) && !(val > above);
    }
    if (!ext) continue;
    Cand k;
    if (!adjust_local_extrema(O, octv, layer, r, c, &k)) continue;
    const int slot = atomicAdd(&n_cand[f], 1);
    if (slot >= cap) { status[f] = 1; continue; }
    int32_t* w = cand + ((size_t)f * cap + slot) * CAND_WORDS;
    w[0] = k.o; w[1] = k.layer; w[2] = k.r; w[3] = k.c;
    w[4] = __float_as_int(k.x); w[5] = __float_as_int(k.y); w[6] = __float_as_int(k.size);
    w[7] = __float_as_int(k.response); w[8] = k.octave_word;
  }
}

// ------------------------------------------------------------------ hal::fastAtan2 (degrees)
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float k = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k;
  const float p5 = 0.1555786518463281f * k, p7 = -0.04432655554792128f * k;
  const float ax = fabsf(x), ay = fabsf(y);
  const float eps = (float)2.2204460492503131e-16;
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + eps);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + eps);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0.f) a = 180.f - a;
  if (y < 0.f) a = 360.f - a;
  return a;
}

// ------------------------------------------------------------------ S5: orientations
// One block of 64 threads per refined extremum: the window's samples (bin, weight * magnitude) go
// to shared memory in row-major order, then thread b sums bin b IN THAT ORDER (cv2's scalar
// accumulation), so the histogram does not depend on scheduling.
constexpr int ORI_MAX_RADIUS = 24;
constexpr int ORI_MAX_SAMPLES = (2 * ORI_MAX_RADIUS + 1) * (2 * ORI_MAX_RADIUS + 1);
constexpr int KP_WORDS = 6;
__global__ void __launch_bounds__(64)
orient_kernel(const float* __restrict__ pyr, Geom G, int cap, const int32_t* __restrict__ n_cand,
              const int32_t* __restrict__ cand, int32_t* __restrict__ n_raw, float* __restrict__ raw,
              int32_t* __restrict__ status) {
  const int f = blockIdx.y;
  const int n_c = n_cand[f] < cap ? n_cand[f] : cap;
  __shared__ float contrib[ORI_MAX_SAMPLES];
  __shared__ uint8_t bins[ORI_MAX_SAMPLES];
  __shared__ float temp[ORI_BINS + 4], hist[ORI_BINS];
  for (int ci = blockIdx.x; ci < n_c; ci += gridDim.x) {
    const int32_t* w = cand + ((size_t)f * cap + ci) * CAND_WORDS;
    const int octv = w[0], layer = w[1], pr = w[2], pc = w[3];
    const float size = __int_as_float(w[6]);
    const float* img = pyr + (size_t)f * G.frame_stride + G.off[octv] +
                       (size_t)layer * G.h[octv] * G.w[octv];
    const int h = G.h[octv], wd = G.w[octv];
    const float scl_octv = size * 0.5f / (float)(1 << octv);
    int radius = __float2int_rn(4.5f * scl_octv);       // SIFT_ORI_RADIUS = 3 * 1.5
    if (radius > ORI_MAX_RADIUS) radius = ORI_MAX_RADIUS;   // unreachable: scl_octv < 4.1
    const float sigma = 1.5f * scl_octv;
    const float expf_scale = -1.f / (2.f * sigma * sigma);
    const int side = 2 * radius + 1, len = side * side;
    for (int k = threadIdx.x; k < len; k += 64) {
      const int i = k / side - radius, j = k % side - radius;
      const int y = pr + i, x = pc + j;
      uint8_t b = 255;
      float v = 0.f;
      if (y > 0 && y < h - 1 && x > 0 && x < wd - 1) {
        const float dx = img[(size_t)y * wd + x + 1] - img[(size_t)y * wd + x - 1];
        const float dy = img[(size_t)(y - 1) * wd + x] - img[(size_t)(y + 1) * wd + x];
        const float wgt = expf((float)(i * i + j * j) * expf_scale);
        const float ori = fast_atan2_deg(dy, dx);
        const float mag = sqrtf(dx * dx + dy * dy);
        int bin = __float2int_rn((ORI_BINS / 360.f) * ori);
        if (bin >= ORI_BINS) bin -= ORI_BINS;
        if (bin < 0) bin += ORI_BINS;
        b = (uint8_t)bin;
        v = wgt * mag;
      }
      bins[k] = b;
      contrib[k] = v;
    }
    __syncthreads();
    if (threadIdx.x < ORI_BINS) {
      float s = 0.f;
      const uint8_t me = (uint8_t)threadIdx.x;
      for (int k = 0; k < len; ++k)
        if (bins[k] == me) s += contrib[k];
      temp[threadIdx.x + 2] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      temp[0] = temp[ORI_BINS];
      temp[1] = temp[ORI_BINS + 1];
      temp[ORI_BINS + 2] = temp[2];
      temp[ORI_BINS + 3] = temp[3];
    }
    __syncthreads();
    if (threadIdx.x < ORI_BINS) {
      const float* t = temp + threadIdx.x + 2;
      hist[threadIdx.x] = (t[-2] + t[2]) * (1.f / 16.f) + (t[-1] + t[1]) * (4.f / 16.f) + t[0] * (6.f / 16.f);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float omax = hist[0];
      for (int b = 1; b < ORI_BINS; ++b) omax = fmaxf(omax, hist[b]);
      const float mag_thr = omax * 0.8f;
      for (int j = 0; j < ORI_BINS; ++j) {
        const int l = j > 0 ? j - 1 : ORI_BINS - 1, r2 = j < ORI_BINS - 1 ? j + 1 : 0;
        if (hist[j] > hist[l] && hist[j] > hist[r2] && hist[j] >= mag_thr) {
          float bin = j + 0.5f * (hist[l] - hist[r2]) / (hist[l] - 2.f * hist[j] + hist[r2]);
          bin = bin < 0.f ? ORI_BINS + bin : (bin >= ORI_BINS ? bin - ORI_BINS : bin);
          float angle = 360.f - (360.f / ORI_BINS) * bin;
          if (fabsf(angle - 360.f) < 1.1920929e-07f) angle = 0.f;
          const int slot = atomicAdd(&n_raw[f], 1);
          if (slot >= cap) { status[f] = 1; continue; }
          float* o = raw + ((size_t)f * cap + slot) * KP_WORDS;
          o[0] = __int_as_float(w[4]); o[1] = __int_as_float(w[5]); o[2] = size; o[3] = angle;
          o[4] = __int_as_float(w[7]); o[5] = __int_as_float(w[8]);     // octave word as raw bits
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ S6: cv2's output order
// KeyPoint12_LessThan: x, y ascending; size descending; angle ascending; response, octave
// descending.  rank = number of key points that sort before this one (index breaks full ties).
__device__ __forceinline__ bool kp_less(const float* a, int ia, const float* b, int ib) {
  if (a[0] != b[0]) return a[0] < b[0];
  if (a[1] != b[1]) return a[1] < b[1];
  if (a[2] != b[2]) return a[2] > b[2];
  if (a[3] != b[3]) return a[3] < b[3];
  if (a[4] != b[4]) return a[4] > b[4];
  const int oa = __float_as_int(a[5]), ob = __float_as_int(b[5]);
  if (oa != ob) return oa > ob;
  return ia < ib;
}

__global__ void __launch_bounds__(256)
rank_kernel(int cap, const int32_t* __restrict__ n_raw, const float* __restrict__ raw,
            float* __restrict__ sorted) {
  const int f = blockIdx.y;
  const int n = n_raw[f] < cap ? n_raw[f] : cap;
  __shared__ float tile[256][KP_WORDS];
  const float* base = raw + (size_t)f * cap * KP_WORDS;
  for (int i0 = blockIdx.x * 256; i0 < n; i0 += gridDim.x * 256) {
    const int i = i0 + threadIdx.x;
    float me[KP_WORDS];
    if (i < n)
      for (int q = 0; q < KP_WORDS; ++q) me[q] = base[(size_t)i * KP_WORDS + q];
    int rank = 0;
    for (int j0 = 0; j0 < n; j0 += 256) {
      __syncthreads();
      if (j0 + threadIdx.x < n)
        for (int q = 0; q < KP_WORDS; ++q) tile[threadIdx.x][q] = base[(size_t)(j0 + threadIdx.x) * KP_WORDS + q];
      __syncthreads();
      if (i < n) {
        const int m = n - j0 < 256 ? n - j0 : 256;
        for (int j = 0; j < m; ++j) rank += kp_less(tile[j], j0 + j, me, i) ? 1 : 0;
      }
    }
    if (i < n) {
      float* o = sorted + ((size_t)f * cap + rank) * KP_WORDS;
      for (int q = 0; q < KP_WORDS; ++q) o[q] = me[q];
    }
    __syncthreads();
  }
}

// removeDuplicatedSorted (equal x, y, size, angle as the predecessor), KeyPointsFilter::retainBest
// for nfeatures > 0 (everything whose response is >= the nfeatures-th largest one, ties kept: a
// radix select over the response bits), then the firstOctave = -1 correction: octave byte - 1,
// point and size halved.  One block per frame.
__device__ __forceinline__ bool sift_not_duplicate(const float* __restrict__ s, int i) {
  if (i == 0) return true;
  const float* a = s + (size_t)(i - 1) * KP_WORDS;
  const float* b = s + (size_t)i * KP_WORDS;
  return !(a[0] == b[0] && a[1] == b[1] && a[2] == b[2] && a[3] == b[3]);
}

__global__ void __launch_bounds__(256)
finalize_kernel(int cap, int nfeatures, const int32_t* __restrict__ n_raw,
                const float* __restrict__ sorted, float* __restrict__ kp, int32_t* __restrict__ n_out) {
  const int f = blockIdx.x;
  const int n = n_raw[f] < cap ? n_raw[f] : cap;
  const float* s = sorted + (size_t)f * cap * KP_WORDS;
  __shared__ int warp_sums[8];
  __shared__ int carry;
  __shared__ int hist[256];
  __shared__ unsigned sel_prefix, sel_mask;
  __shared__ int sel_want;
  unsigned thr_bits = 0u;                      // responses are >= 0: their bit patterns order them
  if (nfeatures > 0) {
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    int mine = 0;
    for (int i = threadIdx.x; i < n; i += 256) mine += sift_not_duplicate(s, i) ? 1 : 0;
    atomicAdd(&carry, mine);
    __syncthreads();
    const int n_unique = carry;
    __syncthreads();
    if (n_unique > nfeatures) {
      if (threadIdx.x == 0) { sel_prefix = 0u; sel_mask = 0u; sel_want = nfeatures; }
      for (int shift = 24; shift >= 0; shift -= 8) {
        hist[threadIdx.x] = 0;
        __syncthreads();
        const unsigned prefix = sel_prefix, mask = sel_mask;
        for (int i = threadIdx.x; i < n; i += 256) {
          const unsigned bits = (unsigned)__float_as_int(s[(size_t)i * KP_WORDS + 4]);
          if ((bits & mask) == prefix && sift_not_duplicate(s, i)) atomicAdd(&hist[(bits >> shift) & 255u], 1);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
          int want = sel_want, b = 255;
          for (; b > 0; --b) {
            if (hist[b] >= want) break;
            want -= hist[b];
          }
          sel_want = want;
          sel_prefix = prefix | ((unsigned)b << shift);
          sel_mask = mask | (255u << shift);
        }
        __syncthreads();
      }
      thr_bits = sel_prefix;
    }
  }
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 256) {
    const int i = i0 + threadIdx.x;
    int keep = 0;
    if (i < n)
      keep = sift_not_duplicate(s, i) && (unsigned)__float_as_int(s[(size_t)i * KP_WORDS + 4]) >= thr_bits;
    const unsigned ball = __ballot_sync(0xffffffffu, keep);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) warp_sums[wid] = __popc(ball);
    __syncthreads();
    int before = carry;
    for (int q = 0; q < wid; ++q) before += warp_sums[q];
    const int pos = before + __popc(ball & ((1u << lane) - 1u));
    if (keep) {
      const float* b = s + (size_t)i * KP_WORDS;
      float* o = kp + ((size_t)f * cap + pos) * KP_WORDS;
      int octw = __float_as_int(b[5]);
      octw = (octw & ~255) | ((octw - 1) & 255);
      o[0] = b[0] * 0.5f; o[1] = b[1] * 0.5f; o[2] = b[2] * 0.5f; o[3] = b[3]; o[4] = b[4];
      o[5] = (float)octw;                      // < 2^24: exact
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int q = 0; q < 8; ++q) t += warp_sums[q];
      carry += t;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) n_out[f] = carry;
}

// ------------------------------------------------------------------ S7: descriptors
// One block of 128 threads per key point; warp i owns row i of the 4 x 4 x 8 histogram and lane
// (j, k) its bin.  The window is walked in row-major order in pieces of 4 x DESC_PART samples, a
// quarter per warp: the warp computes the samples that fall inside the descriptor (bin coordinates,
// trilinear shares in cv2's operation order) and routes each to the lists of the one or two
// histogram rows it touches — ballot compaction keeps every list in sample order, one segment per
// producing warp.  Then warp i walks row i's four segments in order, eight samples at a time, one
// lane per (sample, column, orientation) share; shares of different samples for the same bin are
// added in sample order (match.any rank), the rest in parallel: the float sums are those of cv2's
// scalar loop, independent of scheduling, no atomics.
constexpr int DESC_PART = 128;                   // samples per warp and piece
constexpr int DESC_STRIDE = DESC_PART + 4;       // segment padded with empty words to a multiple of 4
__global__ void __launch_bounds__(128)
describe_kernel(const float* __restrict__ pyr, Geom G, int cap, const int32_t* __restrict__ n_kp,
                const float* __restrict__ kp, uint8_t* __restrict__ desc) {
  const int f = blockIdx.y;
  const int n = n_kp[f];
  // segment (row R, producing warp p) = index R * 4 + p
  __shared__ float share[16][4][DESC_STRIDE];                 // [segment][b * 2 + e][sample]
  __shared__ __align__(4) uint8_t meta[16][DESC_STRIDE];      // (c0 + 1) | o0 << 3; 0xff = none
  __shared__ int seg_count[16];
  __shared__ float hrow[4][36];
  __shared__ float rawd[128];
  __shared__ float s_scale;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned tj = ((t >> 3) & 3) + 1, tk = t & 7;
  constexpr int d = 4, nb = 8;
  for (int idx = blockIdx.x; idx < n; idx += gridDim.x) {
    const float* k = kp + ((size_t)f * cap + idx) * KP_WORDS;
    // calcDescriptors: unpackOctave and back to the octave's own pixel grid
    const int octw = (int)k[5];
    int octave = octw & 255;
    const int layer = (octw >> 8) & 255;
    octave = octave < 128 ? octave : (-128 | octave);
    const float scale = octave >= 0 ? 1.f / (float)(1 << octave) : (float)(1 << -octave);
    const int o = octave + 1;                                 // firstOctave = -1
    const float* img = pyr + (size_t)f * G.frame_stride + G.off[o] + (size_t)layer * G.h[o] * G.w[o];
    const int rows = G.h[o], cols = G.w[o];
    const float ptx = k[0] * scale, pty = k[1] * scale;
    float ori = 360.f - k[3];
    if (fabsf(ori - 360.f) < 1.1920929e-07f) ori = 0.f;
    const float scl = k[2] * scale * 0.5f;
    const int px = __float2int_rn(ptx), py = __float2int_rn(pty);
    float cos_t = cosf(ori * (float)(3.14159265358979323846 / 180.0));
    float sin_t = sinf(ori * (float)(3.14159265358979323846 / 180.0));
    const float bins_per_rad = nb / 360.f;
    const float exp_scale = -1.f / (d * d * 0.5f);
    const float hist_width = 3.f * scl;
    int radius = __float2int_rn(hist_width * 1.4142135623730951f * (d + 1) * 0.5f);
    const int diag = (int)sqrt((double)cols * cols + (double)rows * rows);
    radius = radius < diag ? radius : diag;
    cos_t /= hist_width;
    sin_t /= hist_width;
    const int side = 2 * radius + 1;
    const int len = side * side;              // radius <= image diagonal: fits an int
    for (int q = lane; q < 36; q += 32) hrow[warp][q] = 0.f;   // bins (j, k = 0 .. 8) of this warp's row
    for (int base = 0; base < len; base += 4 * DESC_PART) {
      __syncthreads();
      int fill[4] = {0, 0, 0, 0};             // samples this warp has routed to rows 0 .. 3
      for (int u = 0; u < DESC_PART / 32; ++u) {
        const int s = base + warp * DESC_PART + u * 32 + lane;
        bool ok = false;
        float rbin = 0.f, cbin = 0.f, c_rot = 0.f, r_rot = 0.f;
        int r = 0, c = 0;
        if (s < len) {
          const int si = s / side;
          const int i = si - radius, j = s - si * side - radius;
          c_rot = j * cos_t - i * sin_t;
          r_rot = j * sin_t + i * cos_t;
          rbin = r_rot + d / 2 - 0.5f;
          cbin = c_rot + d / 2 - 0.5f;
          r = py + i;
          c = px + j;
          ok = rbin > -1.f && rbin < d && cbin > -1.f && cbin < d && r > 0 && r < rows - 1 && c > 0 &&
               c < cols - 1;
        }
        if (!__any_sync(0xffffffffu, ok)) continue;
        int r0 = -9, c0 = 0, o0 = 0;
        float v_r0 = 0.f, v_r1 = 0.f, obin = 0.f;
        if (ok) {
          const float dx = img[(size_t)r * cols + c + 1] - img[(size_t)r * cols + c - 1];
          const float dy = img[(size_t)(r - 1) * cols + c] - img[(size_t)(r + 1) * cols + c];
          const float wgt = expf((c_rot * c_rot + r_rot * r_rot) * exp_scale);
          const float og = fast_atan2_deg(dy, dx);
          const float mag = sqrtf(dx * dx + dy * dy) * wgt;
          obin = (og - ori) * bins_per_rad;
          r0 = (int)floorf(rbin);
          c0 = (int)floorf(cbin);
          o0 = (int)floorf(obin);
          rbin -= r0; cbin -= c0; obin -= o0;
          if (o0 < 0) o0 += nb;
          if (o0 >= nb) o0 -= nb;
          v_r1 = mag * rbin;
          v_r0 = mag - v_r1;
        }
#pragma unroll
        for (int R = 0; R < 4; ++R) {
          const bool lower = r0 == R, upper = r0 + 1 == R;      // share of row r0 / of row r0 + 1
          const bool keep = lower || upper;
          const unsigned ball = __ballot_sync(0xffffffffu, keep);
          if (keep) {
            const int q = fill[R] + __popc(ball & ((1u << lane) - 1u));
            const int seg = R * 4 + warp;
            const float v_r = upper ? v_r1 : v_r0;
            const float v_c1 = v_r * cbin, v_c0 = v_r - v_c1;
            const float v_11 = v_c1 * obin, v_10 = v_c1 - v_11;
            const float v_01 = v_c0 * obin, v_00 = v_c0 - v_01;
            share[seg][0][q] = v_00; share[seg][1][q] = v_01;
            share[seg][2][q] = v_10; share[seg][3][q] = v_11;
            meta[seg][q] = (uint8_t)((c0 + 1) | (o0 << 3));
          }
          fill[R] += __popc(ball);
        }
      }
#pragma unroll
      for (int R = 0; R < 4; ++R) {
        if (lane < 4) meta[R * 4 + warp][fill[R] + lane] = 0xff;
        if (lane == 0) seg_count[R * 4 + warp] = fill[R];
      }
      __syncthreads();
      // row `warp`: the segments of producers 0 .. 3 in order, eight samples at a time: lane
      // (s, b, e) carries the share of sample s for column c0 + b and orientation o0 + e; shares
      // of different samples that land in the same bin are added in sample order (rank among
      // the lanes with the same bin), everything else in parallel
      float* hist = hrow[warp];
      for (int p = 0; p < 4; ++p) {
        const int seg = warp * 4 + p;
        const int m = seg_count[seg];
        for (int q8 = 0; q8 < m; q8 += 8) {
          const int q = q8 + (lane >> 2);
          const unsigned b = (lane >> 1) & 1u, e = lane & 1u;
          const unsigned mv = q < m ? meta[seg][q] : 0xffu;
          const int j = (int)(mv & 7u) - 1 + (int)b;
          const bool ok = mv != 0xffu && j >= 0 && j < d;
          const int bin = j * 9 + (int)((mv >> 3) & 7u) + (int)e;
          const float v = ok ? share[seg][b * 2 + e][q] : 0.f;
          const unsigned peers = __match_any_sync(0xffffffffu, ok ? bin : 64 + lane);
          const int rank = __popc(peers & ((1u << lane) - 1u));
          const int steps = __reduce_max_sync(0xffffffffu, ok ? rank + 1 : 0);
          for (int st = 0; st < steps; ++st) {
            if (ok && rank == st) hist[bin] += v;
            __syncwarp();
          }
        }
      }
    }
    __syncthreads();
    const float acc = hrow[warp][(tj - 1) * 9 + tk];
    const float acc_wrap = tk == 0 ? hrow[warp][(tj - 1) * 9 + 8] : 0.f;
    // circular orientation bins: hist[0] += hist[n] (hist[n + 1] never receives anything)
    rawd[t] = tk == 0 ? acc + acc_wrap : acc;
    __syncthreads();
    if (t == 0) {
      float nrm2 = 0.f;
      for (int q = 0; q < 128; ++q) nrm2 += rawd[q] * rawd[q];
      const float thr = sqrtf(nrm2) * 0.2f;
      nrm2 = 0.f;
      for (int q = 0; q < 128; ++q) {
        const float v = fminf(rawd[q], thr);
        rawd[q] = v;
        nrm2 += v * v;
      }
      s_scale = 512.f / fmaxf(sqrtf(nrm2), 1.1920929e-07f);
    }
    __syncthreads();
    const int v = __float2int_rn(rawd[t] * s_scale);
    desc[((size_t)f * cap + idx) * 128 + t] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
    __syncthreads();
  }
}

// ------------------------------------------------------------------ workspace
struct Ws {
  float* pyr;        // [F] Gaussian pyramids
  int32_t* n_cand;   // [F]
  int32_t* n_raw;    // [F]
  int32_t* cand;     // [F, cap, CAND_WORDS]
  float* raw;        // [F, cap, KP_WORDS]
  float* sorted;     // [F, cap, KP_WORDS]
  size_t bytes;
};

static Ws carve(void* base, const Geom& G, int n_frames, int cap) {
  Ws w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += (bytes + 255) / 256 * 256;
    return p;
  };
  w.pyr = (float*)take((size_t)n_frames * G.frame_stride * 4);
  w.n_cand = (int32_t*)take((size_t)n_frames * 4);
  w.n_raw = (int32_t*)take((size_t)n_frames * 4);
  w.cand = (int32_t*)take((size_t)n_frames * cap * CAND_WORDS * 4);
  w.raw = (float*)take((size_t)n_frames * cap * KP_WORDS * 4);
  w.sorted = (float*)take((size_t)n_frames * cap * KP_WORDS * 4);
  w.bytes = off;
  return w;
}

}  // namespace sift
}  // namespace vo

using namespace vo::sift;

extern "C" int vo_sift_octaves(int width, int height, int32_t* w_h, int32_t* h_h) {
  Geom G;
  if (make_geom(width, height, &G)) return 0;
  for (int o = 0; o < G.n_oct; ++o) {
    if (w_h) w_h[o] = G.w[o];
    if (h_h) h_h[o] = G.h[o];
  }
  return G.n_oct;
}

extern "C" size_t vo_sift_workspace_bytes(int n_frames, int width, int height, int cap) {
  Geom G;
  if (n_frames < 1 || cap < 1 || width < 2 || height < 2 || make_geom(width, height, &G)) return 0;
  return carve(nullptr, G, n_frames, cap).bytes;
}

extern "C" int vo_sift_detect_batch(const uint8_t* images, int n_frames, int width, int height,
                                    int cap, int nfeatures, float* kp, uint8_t* desc, int32_t* n_out,
                                    int32_t* status, float* pyramid_out, void* ws, size_t ws_bytes,
                                    void* stream) {
  VO_REQUIRE(images && kp && desc && n_out && status && ws, VO_ERR_INVALID_ARG,
             "vo_sift_detect_batch: null pointer");
  VO_REQUIRE(n_frames >= 1 && cap >= 1 && width >= 2 && height >= 2, VO_ERR_INVALID_ARG,
             "vo_sift_detect_batch: bad sizes (%d frames, %d x %d, cap %d)", n_frames, width, height, cap);
  VO_REQUIRE(n_frames <= 65535, VO_ERR_INVALID_ARG, "vo_sift_detect_batch: at most 65535 frames per call");
  VO_REQUIRE(nfeatures >= 0, VO_ERR_INVALID_ARG, "vo_sift_detect_batch: nfeatures %d < 0", nfeatures);
  Geom G;
  VO_REQUIRE(make_geom(width, height, &G) == 0, VO_ERR_INVALID_ARG,
             "vo_sift_detect_batch: image size %d x %d gives no valid octave count", width, height);
  Kernels K;
  VO_REQUIRE(make_kernels(&K) == 0, VO_ERR_INVALID_ARG, "vo_sift_detect_batch: Gaussian kernel too wide");
  const Ws w = carve(ws, G, n_frames, cap);
  VO_REQUIRE(ws_bytes >= w.bytes, VO_ERR_WORKSPACE, "vo_sift_detect_batch: workspace %zu < %zu bytes",
             ws_bytes, w.bytes);
  cudaStream_t st = (cudaStream_t)stream;
  const int F = n_frames;
  cudaMemsetAsync(w.n_cand, 0, (size_t)F * 4, st);
  cudaMemsetAsync(w.n_raw, 0, (size_t)F * 4, st);
  cudaMemsetAsync(status, 0, (size_t)F * 4, st);
  cudaMemsetAsync(n_out, 0, (size_t)F * 4, st);
  {
    vo::ProfScope prof(10, st);
    // base image: 2x upsampling into layer 1's slot, blurred into layer 0
    float* l0 = w.pyr + G.off[0];
    float* tmp = l0 + (size_t)G.h[0] * G.w[0];
    upsample_kernel<<<dim3((2 * width + 255) / 256, 2 * height, F), 256, 0, st>>>(images, height, width,
                                                                                tmp, G.frame_stride);
    VO_REQUIRE(launch_blur(tmp, l0, G.h[0], G.w[0], G.frame_stride, K, 0, F, st) == 0, VO_ERR_INVALID_ARG,
               "vo_sift_detect_batch: no blur kernel for half width %d", K.r[0]);
    for (int o = 0; o < G.n_oct; ++o) {
      const size_t ls = (size_t)G.h[o] * G.w[o];
      float* base = w.pyr + G.off[o];
      if (o > 0) {
        const float* src = w.pyr + G.off[o - 1] + (size_t)NL * G.h[o - 1] * G.w[o - 1];
        halve_kernel<<<dim3((G.w[o] + 255) / 256, G.h[o], F), 256, 0, st>>>(
            src, G.h[o - 1], G.w[o - 1], base, G.h[o], G.w[o], G.frame_stride);
      }
      for (int i = 1; i < LAYERS; ++i)
        VO_REQUIRE(launch_blur(base + (i - 1) * ls, base + i * ls, G.h[o], G.w[o], G.frame_stride, K, i, F,
                               st) == 0,
                   VO_ERR_INVALID_ARG, "vo_sift_detect_batch: no blur kernel for half width %d", K.r[i]);
    }
  }
  {
    vo::ProfScope prof(11, st);
    for (int o = 0; o < G.n_oct; ++o) {
      if (G.h[o] <= 2 * BORDER || G.w[o] <= 2 * BORDER) continue;
      extrema_kernel<<<dim3((G.w[o] + 31) / 32, (G.h[o] + 7) / 8, F), 256, 0, st>>>(
          w.pyr, G, o, cap, w.n_cand, w.cand, status);
    }
    const int bx = cap < 2048 ? cap : 2048;
    orient_kernel<<<dim3(bx, F), 64, 0, st>>>(w.pyr, G, cap, w.n_cand, w.cand, w.n_raw, w.raw, status);
    rank_kernel<<<dim3((cap + 255) / 256 < 16 ? (cap + 255) / 256 : 16, F), 256, 0, st>>>(cap, w.n_raw, w.raw,
                                                                                         w.sorted);
    finalize_kernel<<<F, 256, 0, st>>>(cap, nfeatures, w.n_raw, w.sorted, kp, n_out);
  }
  {
    vo::ProfScope prof(12, st);
    const int bx = cap < 2048 ? cap : 2048;
    describe_kernel<<<dim3(bx, F), 128, 0, st>>>(w.pyr, G, cap, n_out, kp, desc);
  }
  if (pyramid_out)
    cudaMemcpyAsync(pyramid_out, w.pyr, (size_t)F * G.frame_stride * 4, cudaMemcpyDeviceToDevice, st);
  return vo::check_launch("vo_sift_detect_batch");
}
