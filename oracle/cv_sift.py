"""TEST INFRASTRUCTURE (oracle): CPU restatement of cv2.SIFT_create().detectAndCompute, the detector
the reference builds at /root/reference/src/visual_slam.py:17 and
/root/reference/src/image_and_keypoints.py:12 and calls at /root/reference/src/frame_generator.py:25
/ src/image_and_keypoints.py:46.  The algorithm lives in OpenCV (opencv-python pinned 4.7.0.72 in
the reference's Pipfile.lock, 4.13.0 in this image; modules/features2d/src/sift.dispatch.cpp and
sift.simd.hpp), not in /root/reference, so this file restates OpenCV's published algorithm
(Lowe 2004 as implemented by OpenCV: base image doubled, sigma 1.6, 3 layers per octave,
contrast 0.04, edge 10) in numpy float32 and is PINNED against the cv2 binary itself
(tests/test_oracle_sift.py: key-point sets, angles, sizes, responses, octave words, descriptors)
and against the committed golden vectors tests/golden/siftdet_*.npz.

Float work: cv2 is built with SIMD/FMA dispatch, so single pixels of the pyramid differ from this
restatement in the last float bits; discrete decisions next to a threshold can therefore flip for
a few key points in a thousand.  The tolerance is stated in the tests, not here.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np

F = np.float32

N_OCTAVE_LAYERS = 3
CONTRAST_THRESHOLD = 0.04
EDGE_THRESHOLD = 10.0
SIGMA = 1.6
INIT_SIGMA = 0.5
IMG_BORDER = 5
MAX_INTERP_STEPS = 5
ORI_HIST_BINS = 36
ORI_SIG_FCTR = 1.5
ORI_RADIUS = 3 * ORI_SIG_FCTR
ORI_PEAK_RATIO = 0.8
DESCR_WIDTH = 4
DESCR_HIST_BINS = 8
DESCR_SCL_FCTR = 3.0
DESCR_MAG_THR = 0.2
INT_DESCR_FCTR = 512.0
FLT_EPSILON = np.finfo(np.float32).eps


def cv_round(x):
    """cvRound: round half to even (lrint)."""
    return int(np.rint(x))


def bgr_to_gray(img):
    """cvtColor(BGR2GRAY) for uint8 as this cv2 computes it (pinned in tests/test_oracle_orb.py)."""
    x = img.astype(np.int32)
    return ((x[..., 0] * 3735 + x[..., 1] * 19235 + x[..., 2] * 9798 + (1 << 14)) >> 15).astype(np.uint8)


def gaussian_kernel(sigma):
    """GaussianBlur(src, dst, Size(), sigma) on a float image: ksize = cvRound(sigma*4*2+1)|1,
    getGaussianKernel(ksize, sigma, CV_32F): exp(-x^2 / 2 sigma^2) normalised in double, stored
    as float.  Returns the half kernel f[0..R] (symmetric)."""
    ksize = cv_round(float(sigma) * 8 + 1) | 1
    r = ksize // 2
    x = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-(x * x) / (2.0 * float(sigma) * float(sigma)))
    k /= k.sum()
    return k.astype(F)[r:]


def _fma(a, b, c):
    """fmaf emulated through float64 (the product is exact there)."""
    return (a.astype(np.float64) * np.float64(b) + c.astype(np.float64)).astype(F)


def blur(img, sigma):
    """Separable filter, BORDER_REFLECT_101, in the order the cv2 binary computes it on an
    AVX2/FMA host (found by matching cv2.GaussianBlur bit for bit, tests/test_oracle_sift.py):
    rows: acc = 0; acc = fma(x[c - r + k], f[k], acc) for k = 0 .. 2r (RowVec_32f);
    columns: acc = f0*x[c]; acc = fma(x[c-k] + x[c+k], f[k], acc) for k = 1 .. r
    (SymmColumnVec_32f).  The scalar tails of the vector loops (the last w % 4 columns of the row
    pass, the last w % 8 columns of the column pass) multiply and add with separate roundings."""
    f = gaussian_kernel(sigma)
    r = len(f) - 1
    full = np.concatenate([f[:0:-1], f])
    h, w = img.shape

    def reflect(n, r):
        # REFLECT_101 index map for positions -r .. n-1+r (OpenCV's borderInterpolate, also valid
        # when r >= n)
        idx = np.arange(-r, n + r)
        if n == 1:
            return np.zeros_like(idx)
        period = 2 * (n - 1)
        idx = np.mod(idx, period)
        return np.where(idx >= n, period - idx, idx)

    p = img[:, reflect(w, r)]
    acc = np.zeros((h, w), F)
    t = (w // 4) * 4                  # row pass: the last w % 4 columns are the scalar tail
    tail = (p[:, t:t + w - t] * full[0]).astype(F)
    for k in range(2 * r + 1):
        acc = _fma(p[:, k:k + w], full[k], acc)
        if k:
            tail = (p[:, t + k:k + w] * full[k]).astype(F) + tail
    acc[:, t:] = tail
    p = acc[reflect(h, r), :]
    out = (f[0] * p[r:r + h, :]).astype(F)
    t = (w // 8) * 8                  # column pass: the last w % 8 columns
    tail = out[:, t:].copy()
    for k in range(1, r + 1):
        pair = (p[r - k:r - k + h, :] + p[r + k:r + k + h, :]).astype(F)
        out = _fma(pair, f[k], out)
        tail = (pair[:, t:] * f[k]).astype(F) + tail
    out[:, t:] = tail
    return out


def upsample2_linear(img):
    """resize(src, Size(2w, 2h), INTER_LINEAR) on float: source coordinate (d + 0.5)/2 - 0.5,
    clamped at the borders; weights are exactly 0.25 / 0.75."""
    h, w = img.shape

    def taps(n):
        d = np.arange(2 * n)
        fx = (d + 0.5) * 0.5 - 0.5
        i0 = np.floor(fx).astype(np.int64)
        a = (fx - i0).astype(F)
        lo = i0 < 0
        i0 = np.where(lo, 0, i0)
        a = np.where(lo, F(0), a)
        hi = i0 >= n - 1
        i1 = np.where(hi, n - 1, i0 + 1)
        i0 = np.minimum(i0, n - 1)
        a = np.where(hi, F(0), a)
        return i0, i1, a

    x0, x1, ax = taps(w)
    y0, y1, ay = taps(h)
    # OpenCV resizes rows horizontally first, then blends two rows vertically
    hr = img[:, x0] * (F(1) - ax)[None, :] + img[:, x1] * ax[None, :]
    out = hr[y0, :] * (F(1) - ay)[:, None] + hr[y1, :] * ay[:, None]
    return out.astype(F)


def halve_nearest(img):
    """resize(src, Size(w/2, h/2), INTER_NEAREST): sx = min(floor(x * (w / (w/2))), w - 1)."""
    h, w = img.shape
    dh, dw = h // 2, w // 2
    sy = np.minimum(np.floor(np.arange(dh) * (h / dh)).astype(np.int64), h - 1)
    sx = np.minimum(np.floor(np.arange(dw) * (w / dw)).astype(np.int64), w - 1)
    return np.ascontiguousarray(img[sy][:, sx])
