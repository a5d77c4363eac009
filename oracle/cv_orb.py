"""TEST INFRASTRUCTURE ONLY - CPU restatement (numpy) of OpenCV's ORB front end, the detector the
reference hands to FrameGenerator (reference src/frame_generator.py:25 `detectAndCompute`,
src/visual_slam.py:16 `cv2.ORB_create()`, src/image_and_keypoints.py:8).

The arithmetic lives in the third-party dependency opencv-python (Pipfile.lock pins 4.7.0.72; this
container has 4.13.0), whose source is not on this box.  The restatement follows the published
algorithm (upstream modules/features2d/src/orb.cpp, fast.cpp, fast_score.cpp, keypoint.cpp,
imgproc resize / smooth bit-exact paths) and is PINNED by running cv2 itself on the same images:
tests/test_oracle_orb.py compares every stage with cv2.FastFeatureDetector, cv2.GaussianBlur,
cv2.resize, cv2.fastAtan2 and cv2.ORB_create().detectAndCompute, and tests/golden/orb_*.npz holds
cv2's outputs for the GPU box (generator: tests/tools/make_golden_orb.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.

Order of the keypoints: cv2 truncates every pyramid level with std::nth_element
(KeyPointsFilter::retainBest), which leaves the survivors in an implementation-defined order.  The
SET of survivors is defined (everything at least as good as the n-th best response, ties kept);
this restatement and the CUDA path return them in (level, y, x) order and the parity tests compare
after sorting cv2's output the same way.
"""
import numpy as np

f32 = np.float32

# Bresenham circle of radius 3 (x, y), clockwise from the bottom
CIRCLE16 = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3),
            (0, -3), (-1, -3), (-2, -2), (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]
HARRIS_K = f32(0.04)
HARRIS_BLOCK = 7
PATCH = 31
HALF_PATCH = 15
BLUR_K = np.array([18, 34, 48, 56, 48, 34, 18], np.int64)   # 7 taps, sigma 2, 8 fractional bits


# ----------------------------------------------------------------------------- FAST-9/16
def fast9_score_map(img, threshold):
    """score(x, y) = max(t, max_arc min(v - p), max_arc min(p - v)) - 1 over the 16 arcs of 9
    contiguous circle pixels when the pixel is a corner (some arc entirely > t brighter or darker),
    else 0; only 3 <= x < W-3, 3 <= y < H-3 are tested (fast.cpp FAST_t<16>, fast_score.cpp
    cornerScore<16>)."""
    img = np.asarray(img, np.uint8)
    H, W = img.shape
    out = np.zeros((H, W), np.int32)
    if H < 7 or W < 7:
        return out
    v = img[3:H - 3, 3:W - 3].astype(np.int32)
    d = np.stack([v - img[3 + dy:H - 3 + dy, 3 + dx:W - 3 + dx].astype(np.int32) for dx, dy in CIRCLE16])
    d2 = np.concatenate([d, d[:8]])
    amin = np.full(v.shape, -1000, np.int32)
    bmin = np.full(v.shape, -1000, np.int32)
    for s in range(16):
        arc = d2[s:s + 9]
        amin = np.maximum(amin, arc.min(0))
        bmin = np.maximum(bmin, (-arc).min(0))
    raw = np.maximum(amin, bmin)
    out[3:H - 3, 3:W - 3] = np.where(raw > threshold, raw - 1, 0)
    return out


def fast9_detect(img, threshold=20, nms=True):
    """-> (xy int32 [n,2], score int32 [n]) in row-major order (cv2.FastFeatureDetector.detect).
    Non-maximum suppression keeps a corner whose score is strictly greater than all 8 neighbours."""
    s = fast9_score_map(img, threshold)
    H, W = s.shape
    keep = s > 0
    if nms:
        p = np.pad(s, 1)
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if dx or dy:
                    keep &= s > p[1 + dy:1 + dy + H, 1 + dx:1 + dx + W]
    ys, xs = np.nonzero(keep)
    return np.stack([xs, ys], 1).astype(np.int32), s[ys, xs]


# ----------------------------------------------------------------------------- filters
def run_by_image_border(xy, shape, border):
    """KeyPointsFilter::runByImageBorder: keep border <= x < W - border, border <= y < H - border."""
    H, W = shape
    if border > 0 and (H <= 2 * border or W <= 2 * border):
        return np.zeros(len(xy), bool)
    return (xy[:, 0] >= border) & (xy[:, 0] < W - border) & (xy[:, 1] >= border) & (xy[:, 1] < H - border)


def retain_best(response, n):
    """KeyPointsFilter::retainBest as a mask: everything >= the n-th largest response (ties kept)."""
    m = len(response)
    if n < 0 or m <= n:
        return np.ones(m, bool)
    if n == 0:
        return np.zeros(m, bool)
    thr = np.sort(response)[::-1][n - 1]
    return response >= thr


# ----------------------------------------------------------------------------- Harris / angle
def harris_response(img, xy):
    """orb.cpp HarrisResponses: 7x7 block of Sobel-like integer gradients on the raw level image."""
    img = np.asarray(img, np.int64)
    r = HARRIS_BLOCK // 2
    scale = f32(1.0) / (f32(4 * HARRIS_BLOCK) * f32(255.0))
    scale4 = scale * scale * scale * scale
    out = np.zeros(len(xy), f32)
    for k, (x0, y0) in enumerate(xy):
        p = img[y0 - r - 1:y0 + r + 2, x0 - r - 1:x0 + r + 2]
        ix = (p[1:-1, 2:] - p[1:-1, :-2]) * 2 + (p[:-2, 2:] - p[:-2, :-2]) + (p[2:, 2:] - p[2:, :-2])
        iy = (p[2:, 1:-1] - p[:-2, 1:-1]) * 2 + (p[2:, :-2] - p[:-2, :-2]) + (p[2:, 2:] - p[:-2, 2:])
        a, b, c = int((ix * ix).sum()), int((iy * iy).sum()), int((ix * iy).sum())
        # int accumulators wrap at 32 bits in C; a 7x7 block of |I| <= 1020 stays below 2^31
        fa, fb, fc = f32(a), f32(b), f32(c)
        out[k] = (fa * fb - fc * fc - HARRIS_K * (fa + fb) * (fa + fb)) * scale4
    return out


def umax_table():
    """end of every row of the circular patch (orb.cpp computeKeyPoints)."""
    umax = np.zeros(HALF_PATCH + 2, np.int64)
    vmax = int(np.floor(HALF_PATCH * np.sqrt(f32(2.0)) / 2 + 1))
    vmin = int(np.ceil(HALF_PATCH * np.sqrt(f32(2.0)) / 2))
    for v in range(vmax + 1):
        umax[v] = int(np.rint(np.sqrt(float(HALF_PATCH * HALF_PATCH - v * v))))
    v0 = 0
    for v in range(HALF_PATCH, vmin - 1, -1):
        while umax[v0] == umax[v0 + 1]:
            v0 += 1
        umax[v] = v0
        v0 += 1
    return umax[:HALF_PATCH + 1]


_K = f32(180.0 / np.pi)
_P1, _P3, _P5, _P7 = (f32(0.9997878412794807) * _K, f32(-0.3258083974640975) * _K,
                      f32(0.1555786518463281) * _K, f32(-0.04432655554792128) * _K)
_EPS = f32(2.220446049250313e-16)


def fast_atan2(y, x):
    """cv::fastAtan2 (degrees, 7th-order odd polynomial, plain float multiply-adds)."""
    y, x = f32(y), f32(x)
    ax, ay = np.abs(x), np.abs(y)
    if ax >= ay:
        c = ay / (ax + _EPS)
        c2 = c * c
        a = (((_P7 * c2 + _P5) * c2 + _P3) * c2 + _P1) * c
    else:
        c = ax / (ay + _EPS)
        c2 = c * c
        a = f32(90.0) - (((_P7 * c2 + _P5) * c2 + _P3) * c2 + _P1) * c
    if x < 0:
        a = f32(180.0) - a
    if y < 0:
        a = f32(360.0) - a
    return f32(a)


def ic_angle(img, xy):
    """orb.cpp ICAngles: intensity-centroid orientation over the circular patch of radius 15."""
    img = np.asarray(img, np.int64)
    um = umax_table()
    out = np.zeros(len(xy), f32)
    for k, (x0, y0) in enumerate(xy):
        m01 = m10 = 0
        for v in range(-HALF_PATCH, HALF_PATCH + 1):
            d = int(um[abs(v)])
            row = img[y0 + v, x0 - d:x0 + d + 1]
            m10 += int((np.arange(-d, d + 1) * row).sum())
            m01 += v * int(row.sum())
        with np.errstate(all="ignore"):
            out[k] = fast_atan2(m01, m10)
    return out


# ----------------------------------------------------------------------------- pyramid, blur
def _linear_coefs(src, dst):
    scale = np.float64(src) / np.float64(dst)
    x = (np.arange(dst) + 0.5) * scale - 0.5
    fx = np.floor(x).astype(np.int64)
    fr = x - fx
    lo = fx < 0
    fx[lo], fr[lo] = 0, 0.0
    hi = fx >= src - 1
    fx[hi], fr[hi] = src - 1, 0.0
    a1 = np.rint(fr * 256).astype(np.int64)
    return fx, np.minimum(fx + 1, src - 1), 256 - a1, a1


def resize_linear_exact(img, w2, h2):
    """cv2.resize(..., INTER_LINEAR_EXACT) for uint8: 8-bit fixed-point taps, horizontal pass kept
    in 8.8, vertical pass in 16.16, round half up."""
    img = np.asarray(img, np.int64)
    H, W = img.shape
    x0, x1, ax0, ax1 = _linear_coefs(W, w2)
    y0, y1, ay0, ay1 = _linear_coefs(H, h2)
    h = img[:, x0] * ax0 + img[:, x1] * ax1
    v = h[y0, :] * ay0[:, None] + h[y1, :] * ay1[:, None]
    return ((v + (1 << 15)) >> 16).astype(np.uint8)


def gaussian_blur_fixed7(img):
    """cv2.GaussianBlur(img, (7,7), 2, 2, BORDER_REFLECT_101) for a CONTINUOUS uint8 image
    (bit-exact fixed-point path: taps [18,34,48,56,48,34,18]/256, round half up after both passes).
    NOT what ORB uses (see gaussian_blur7); kept because the probe of the cv2 binary showed the two
    differ and the test pins both."""
    img = np.asarray(img, np.uint8)
    H, W = img.shape
    p = np.pad(img.astype(np.int64), 3, mode="reflect")
    h = sum(BLUR_K[i] * p[:, i:i + W] for i in range(7))
    v = sum(BLUR_K[i] * h[i:i + H, :] for i in range(7))
    return ((v + (1 << 15)) >> 16).astype(np.uint8)


def gaussian_taps7():
    """float32 taps of getGaussianKernel(7, 2, CV_32F): exp(-x^2 / 8) normalised in double."""
    x = np.arange(-3, 4, dtype=np.float64)
    k = np.exp(-(x * x) / 8.0)
    return (k / k.sum()).astype(f32)


def _fma(a, b, c):
    return (a.astype(np.float64) * np.float64(b) + c.astype(np.float64)).astype(f32)


def gaussian_blur7(img):
    """The blur ORB applies before sampling (orb.cpp: GaussianBlur(level ROI, 7x7, sigma 2,
    BORDER_REFLECT_101)).  The level is a SUBMATRIX of the pyramid buffer, so OpenCV takes the
    generic separable float filter, not the bit-exact fixed-point one: float32 taps; row pass
    k0*p0 then fused multiply-adds in tap order; column pass in the symmetric form
    k3*c + fma(k4, (r+1)+(r-1)) + ...; round half to even, saturate.  Summation order and fusing
    were determined against cv2.sepFilter2D on 12 Mpixel (0 differing pixels; every other order
    differs in 24-69)."""
    img = np.asarray(img, np.uint8)
    H, W = img.shape
    kf = gaussian_taps7()
    p = np.pad(img.astype(f32), 3, mode="reflect")
    h = kf[0] * p[:, 0:W]
    for i in range(1, 7):
        h = _fma(p[:, i:i + W], kf[i], h)
    v = kf[3] * h[3:3 + H]
    for i in (1, 2, 3):
        v = _fma(h[3 + i:3 + i + H] + h[3 - i:3 - i + H], kf[3 + i], v)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


def _sf(scale_factor):
    # ORB::create takes scaleFactor as a float and keeps it in a double member: 1.2f, not 1.2
    return np.float64(f32(scale_factor))


def level_scales(nlevels, scale_factor=1.2):
    return np.array([f32(np.power(_sf(scale_factor), np.float64(l))) for l in range(nlevels)], f32)


def level_sizes(shape, nlevels, scale_factor=1.2):
    H, W = shape
    sc = level_scales(nlevels, scale_factor)
    return [(int(np.rint(f32(W) / s)), int(np.rint(f32(H) / s))) for s in sc]   # (w, h)


def features_per_level(nfeatures, nlevels, scale_factor=1.2):
    factor = f32(1.0 / _sf(scale_factor))
    nd = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(np.power(np.float64(factor), np.float64(nlevels))))
    out, total = [], 0
    for _ in range(nlevels - 1):
        out.append(int(np.rint(nd)))
        total += out[-1]
        nd = f32(nd * factor)
    out.append(max(nfeatures - total, 0))
    return out


def build_pyramid(img, nlevels, scale_factor=1.2):
    levels = [np.asarray(img, np.uint8)]
    sizes = level_sizes(img.shape, nlevels, scale_factor)
    for l in range(1, nlevels):
        levels.append(resize_linear_exact(levels[-1], *sizes[l]))
    return levels


# ----------------------------------------------------------------------------- rBRIEF
def describe(blurred, pts, angles_deg, pattern):
    """orb.cpp computeOrbDescriptors, WTA_K = 2: bit i = blur(a_i) < blur(b_i), test points rotated
    by the keypoint angle in float arithmetic and rounded half-to-even.
    pts = integer level coordinates [n,2]; pattern = int8 [256,4] (ax, ay, bx, by)."""
    blurred = np.asarray(blurred, np.uint8)
    pat = np.asarray(pattern, np.float32)
    n = len(pts)
    out = np.zeros((n, 32), np.uint8)
    for k in range(n):
        ang = f32(angles_deg[k]) * f32(np.pi / 180.0)
        a, b = f32(np.cos(np.float64(ang))), f32(np.sin(np.float64(ang)))
        cx, cy = int(pts[k][0]), int(pts[k][1])
        xa = np.rint(pat[:, 0] * a - pat[:, 1] * b).astype(np.int64)
        ya = np.rint(pat[:, 0] * b + pat[:, 1] * a).astype(np.int64)
        xb = np.rint(pat[:, 2] * a - pat[:, 3] * b).astype(np.int64)
        yb = np.rint(pat[:, 2] * b + pat[:, 3] * a).astype(np.int64)
        bits = blurred[cy + ya, cx + xa] < blurred[cy + yb, cx + xb]
        out[k] = np.packbits(bits, bitorder="little")
    return out


# ----------------------------------------------------------------------------- whole detector
def orb_detect_and_compute(img, pattern, nfeatures=500, scale_factor=1.2, nlevels=8, edge_threshold=31,
                           fast_threshold=20):
    """cv2.ORB_create(nfeatures, scale_factor, nlevels, edge_threshold, 0, 2, HARRIS_SCORE, 31,
    fast_threshold).detectAndCompute(img, None), survivors of every level in (y, x) order.
    -> dict(pt f32 [n,2], size f32, angle f32, response f32, octave i32, desc u8 [n,32])"""
    levels = build_pyramid(img, nlevels, scale_factor)
    scales = level_scales(nlevels, scale_factor)
    per_level = features_per_level(nfeatures, nlevels, scale_factor)
    res = dict(pt=[], size=[], angle=[], response=[], octave=[], desc=[], level_xy=[])
    for l, lev in enumerate(levels):
        xy, score = fast9_detect(lev, fast_threshold, True)
        m = run_by_image_border(xy, lev.shape, edge_threshold)
        xy, score = xy[m], score[m]
        m = retain_best(score, 2 * per_level[l])
        xy = xy[m]
        hr = harris_response(lev, xy)
        m = retain_best(hr, per_level[l])
        xy, hr = xy[m], hr[m]
        ang = ic_angle(lev, xy)
        desc = describe(gaussian_blur7(lev), xy, ang, pattern)
        res["pt"].append(xy.astype(f32) * scales[l])
        res["size"].append(np.full(len(xy), f32(PATCH) * scales[l], f32))
        res["angle"].append(ang)
        res["response"].append(hr)
        res["octave"].append(np.full(len(xy), l, np.int32))
        res["desc"].append(desc)
        res["level_xy"].append(xy)
    return {k: np.concatenate(v) for k, v in res.items()}
