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


# ---------------------------------------------------------------------------- pyramids
def n_octaves(height, width, first_octave=-1):
    """cvRound(log(min(cols, rows)) / log(2) - 2) - firstOctave, on the ORIGINAL image size."""
    return cv_round(np.log(float(min(width, height))) / np.log(2.0) - 2) - first_octave


def layer_sigmas():
    """incremental blur of layers 1 .. nOctaveLayers + 2 (buildGaussianPyramid), in double"""
    k = 2.0 ** (1.0 / N_OCTAVE_LAYERS)
    sig = [SIGMA]
    for i in range(1, N_OCTAVE_LAYERS + 3):
        prev = (k ** (i - 1)) * SIGMA
        total = prev * k
        sig.append(float(np.sqrt(total * total - prev * prev)))
    return sig


def initial_image(gray_u8):
    """createInitialImage(img, doubleImageSize = true, sigma): float copy, 2x linear upsampling,
    blur by sqrt(max(sigma^2 - (2 * 0.5)^2, 0.01)) (float arithmetic)."""
    g = gray_u8.astype(F)
    sig_diff = np.sqrt(max(F(SIGMA) * F(SIGMA) - F(INIT_SIGMA) * F(INIT_SIGMA) * F(4), F(0.01)), dtype=F)
    return blur(upsample2_linear(g), float(sig_diff))


def gaussian_pyramid(base, n_oct):
    sig = layer_sigmas()
    L = N_OCTAVE_LAYERS + 3
    pyr = []
    for o in range(n_oct):
        for i in range(L):
            if o == 0 and i == 0:
                pyr.append(base)
            elif i == 0:
                pyr.append(halve_nearest(pyr[(o - 1) * L + N_OCTAVE_LAYERS]))
            else:
                pyr.append(blur(pyr[o * L + i - 1], sig[i]))
    return pyr


def dog_pyramid(gpyr, n_oct):
    L = N_OCTAVE_LAYERS + 3
    return [gpyr[o * L + i + 1] - gpyr[o * L + i] for o in range(n_oct) for i in range(L - 1)]


# ---------------------------------------------------------------------------- hal functions
_ATAN2_P1 = F(0.9997878412794807) * F(180 / np.pi)
_ATAN2_P3 = F(-0.3258083974640975) * F(180 / np.pi)
_ATAN2_P5 = F(0.1555786518463281) * F(180 / np.pi)
_ATAN2_P7 = F(-0.04432655554792128) * F(180 / np.pi)


def fast_atan2(y, x):
    """hal::fastAtan2 in degrees (float polynomial, vectorised)."""
    y = np.asarray(y, F)
    x = np.asarray(x, F)
    ax, ay = np.abs(x), np.abs(y)
    eps = F(2.220446049250313e-16)
    swap = ax < ay
    num = np.where(swap, ax, ay)
    den = np.where(swap, ay, ax) + eps
    c = (num / den).astype(F)
    c2 = (c * c).astype(F)
    a = ((((_ATAN2_P7 * c2 + _ATAN2_P5) * c2 + _ATAN2_P3) * c2 + _ATAN2_P1) * c).astype(F)
    a = np.where(swap, F(90) - a, a)
    a = np.where(x < 0, F(180) - a, a)
    a = np.where(y < 0, F(360) - a, a)
    return a.astype(F)


# ---------------------------------------------------------------------------- extrema
def _solve3(H, b):
    """Matx33f::solve(b, DECOMP_LU) = Cramer's rule in float (matx.hpp Matx_FastSolveOp<3,3,1>);
    a singular matrix gives x = 0."""
    a = H
    d = (a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1])
         - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0])
         + a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]))
    if d == 0:
        return [F(0), F(0), F(0)]
    d = F(1) / d
    x0 = d * (b[0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1])
              - a[0][1] * (b[1] * a[2][2] - a[1][2] * b[2])
              + a[0][2] * (b[1] * a[2][1] - a[1][1] * b[2]))
    x1 = d * (a[0][0] * (b[1] * a[2][2] - a[1][2] * b[2])
              - b[0] * (a[1][0] * a[2][2] - a[1][2] * a[2][0])
              + a[0][2] * (a[1][0] * b[2] - b[1] * a[2][0]))
    x2 = d * (a[0][0] * (a[1][1] * b[2] - b[1] * a[2][1])
              - a[0][1] * (a[1][0] * b[2] - b[1] * a[2][0])
              + b[0] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]))
    return [x0, x1, x2]


def adjust_local_extrema(dog, octv, layer, r, c):
    """adjustLocalExtrema (sift.simd.hpp): quadratic fit in (x, y, scale), up to 5 moves, contrast
    and edge tests.  Returns None or dict(x, y, octave word, size, response, r, c, layer)."""
    img_scale = F(1.0 / 255)
    deriv_scale = img_scale * F(0.5)
    second_deriv_scale = img_scale
    cross_deriv_scale = img_scale * F(0.25)
    nl = N_OCTAVE_LAYERS
    xi = xr = xc = F(0)
    i = 0
    while i < MAX_INTERP_STEPS:
        idx = octv * (nl + 2) + layer
        img, prev, nxt = dog[idx], dog[idx - 1], dog[idx + 1]
        dD = [(img[r, c + 1] - img[r, c - 1]) * deriv_scale,
              (img[r + 1, c] - img[r - 1, c]) * deriv_scale,
              (nxt[r, c] - prev[r, c]) * deriv_scale]
        v2 = img[r, c] * F(2)
        dxx = (img[r, c + 1] + img[r, c - 1] - v2) * second_deriv_scale
        dyy = (img[r + 1, c] + img[r - 1, c] - v2) * second_deriv_scale
        dss = (nxt[r, c] + prev[r, c] - v2) * second_deriv_scale
        dxy = (img[r + 1, c + 1] - img[r + 1, c - 1] - img[r - 1, c + 1] + img[r - 1, c - 1]) * cross_deriv_scale
        dxs = (nxt[r, c + 1] - nxt[r, c - 1] - prev[r, c + 1] + prev[r, c - 1]) * cross_deriv_scale
        dys = (nxt[r + 1, c] - nxt[r - 1, c] - prev[r + 1, c] + prev[r - 1, c]) * cross_deriv_scale
        X = _solve3([[dxx, dxy, dxs], [dxy, dyy, dys], [dxs, dys, dss]], dD)
        xi, xr, xc = -X[2], -X[1], -X[0]
        if abs(xi) < 0.5 and abs(xr) < 0.5 and abs(xc) < 0.5:
            break
        big = float(np.iinfo(np.int32).max // 3)
        if abs(xi) > big or abs(xr) > big or abs(xc) > big:
            return None
        c += cv_round(xc)
        r += cv_round(xr)
        layer += cv_round(xi)
        h, w = img.shape
        if (layer < 1 or layer > nl or c < IMG_BORDER or c >= w - IMG_BORDER
                or r < IMG_BORDER or r >= h - IMG_BORDER):
            return None
        i += 1
    if i >= MAX_INTERP_STEPS:
        return None
    idx = octv * (nl + 2) + layer
    img, prev, nxt = dog[idx], dog[idx - 1], dog[idx + 1]
    dD = [(img[r, c + 1] - img[r, c - 1]) * deriv_scale,
          (img[r + 1, c] - img[r - 1, c]) * deriv_scale,
          (nxt[r, c] - prev[r, c]) * deriv_scale]
    t = dD[0] * xc + dD[1] * xr + dD[2] * xi
    contr = img[r, c] * img_scale + t * F(0.5)
    if abs(contr) * nl < F(CONTRAST_THRESHOLD):
        return None
    v2 = img[r, c] * F(2)
    dxx = (img[r, c + 1] + img[r, c - 1] - v2) * second_deriv_scale
    dyy = (img[r + 1, c] + img[r - 1, c] - v2) * second_deriv_scale
    dxy = (img[r + 1, c + 1] - img[r + 1, c - 1] - img[r - 1, c + 1] + img[r - 1, c - 1]) * cross_deriv_scale
    tr = dxx + dyy
    det = dxx * dyy - dxy * dxy
    e = F(EDGE_THRESHOLD)
    if det <= 0 or tr * tr * e >= (e + 1) * (e + 1) * det:
        return None
    size = F(SIGMA) * np.power(F(2), (F(layer) + xi) / F(nl), dtype=F) * F(1 << octv) * F(2)
    return dict(x=(F(c) + xc) * F(1 << octv), y=(F(r) + xr) * F(1 << octv),
                octave=octv + (layer << 8) + (cv_round((float(xi) + 0.5) * 255) << 16),
                size=F(size), response=F(abs(contr)), r=r, c=c, layer=layer)


def extrema_candidates(dog, octv, layer):
    """findScaleSpaceExtremaT: pixels inside the 5-pixel border with |val| > threshold
    (= floor(0.5 * 0.04 / 3 * 255) = 1) that are >= (val > 0) or <= (val < 0) all 26 neighbours.
    Returns (r, c) arrays in row-major order."""
    thr = int(np.floor(0.5 * CONTRAST_THRESHOLD / N_OCTAVE_LAYERS * 255))
    idx = octv * (N_OCTAVE_LAYERS + 2) + layer
    cur, prev, nxt = dog[idx], dog[idx - 1], dog[idx + 1]
    h, w = cur.shape
    b = IMG_BORDER
    if h <= 2 * b or w <= 2 * b:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    val = cur[b:h - b, b:w - b]
    nmax = np.full(val.shape, -np.inf, F)
    nmin = np.full(val.shape, np.inf, F)
    for im in (cur, prev, nxt):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if im is cur and dy == 0 and dx == 0:
                    continue
                nb = im[b + dy:h - b + dy, b + dx:w - b + dx]
                nmax = np.maximum(nmax, nb)
                nmin = np.minimum(nmin, nb)
    ok = (np.abs(val) > thr) & (((val > 0) & (val >= nmax)) | ((val < 0) & (val <= nmin)))
    r, c = np.nonzero(ok)
    return r + b, c + b


def orientation_hist(img, px, py, radius, sigma):
    """calcOrientationHist: 36-bin histogram of gradient directions weighted by magnitude and a
    Gaussian of `sigma`, smoothed with [1 4 6 4 1]/16.  Returns (hist, max)."""
    n = ORI_HIST_BINS
    h, w = img.shape
    expf_scale = F(-1.0) / (F(2.0) * F(sigma) * F(sigma))
    ii, jj = np.meshgrid(np.arange(-radius, radius + 1), np.arange(-radius, radius + 1), indexing="ij")
    y = py + ii
    x = px + jj
    ok = (y > 0) & (y < h - 1) & (x > 0) & (x < w - 1)
    y, x, ii, jj = y[ok], x[ok], ii[ok], jj[ok]
    dx = img[y, x + 1] - img[y, x - 1]
    dy = img[y - 1, x] - img[y + 1, x]
    wgt = np.exp(((ii * ii + jj * jj).astype(F) * expf_scale).astype(F)).astype(F)
    ori = fast_atan2(dy, dx)
    mag = np.sqrt((dx * dx + dy * dy).astype(F)).astype(F)
    bins = np.rint((F(n) / F(360)) * ori).astype(np.int64)
    bins = np.where(bins >= n, bins - n, bins)
    bins = np.where(bins < 0, bins + n, bins)
    temp = np.zeros(n, F)
    contrib = (wgt * mag).astype(F)
    for k in range(len(bins)):          # sequential float accumulation in sample order
        temp[bins[k]] += contrib[k]
    t = np.concatenate([temp[-2:], temp, temp[:2]])
    hist = ((t[0:n] + t[4:n + 4]) * F(1.0 / 16) + (t[1:n + 1] + t[3:n + 3]) * F(4.0 / 16)
            + t[2:n + 2] * F(6.0 / 16)).astype(F)
    return hist, hist.max()


def orientations(hist, omax):
    """peaks >= 0.8 * max, parabolic interpolation -> angles (degrees, 360 - bin * 10)"""
    n = ORI_HIST_BINS
    mag_thr = F(omax * F(ORI_PEAK_RATIO))
    out = []
    for j in range(n):
        l = j - 1 if j > 0 else n - 1
        r2 = j + 1 if j < n - 1 else 0
        if hist[j] > hist[l] and hist[j] > hist[r2] and hist[j] >= mag_thr:
            b = F(j) + F(0.5) * (hist[l] - hist[r2]) / (hist[l] - F(2) * hist[j] + hist[r2])
            b = F(n) + b if b < 0 else (b - F(n) if b >= n else b)
            ang = F(360) - F(F(360.0 / n) * b)
            if abs(ang - F(360)) < FLT_EPSILON:
                ang = F(0)
            out.append(F(ang))
    return out


# ---------------------------------------------------------------------------- descriptors
def descriptor(img, ptx, pty, ori, scl):
    """calcSIFTDescriptor with d = 4, n = 8: 4 x 4 x 8 trilinear histogram over a rotated window,
    clipped at 0.2 of its norm, scaled by 512 / norm and saturated to uint8 values (stored as
    float)."""
    d, n = DESCR_WIDTH, DESCR_HIST_BINS
    h, w = img.shape
    px, py = cv_round(ptx), cv_round(pty)
    cos_t = np.cos(F(ori) * F(np.pi / 180), dtype=F)
    sin_t = np.sin(F(ori) * F(np.pi / 180), dtype=F)
    bins_per_rad = F(n / 360.0)
    exp_scale = F(-1.0) / F(d * d * 0.5)
    hist_width = F(DESCR_SCL_FCTR) * F(scl)
    radius = cv_round(hist_width * F(1.4142135623730951) * F(d + 1) * F(0.5))
    radius = min(radius, int(np.sqrt(float(w) * w + float(h) * h)))
    cos_t = cos_t / hist_width
    sin_t = sin_t / hist_width
    ii, jj = np.meshgrid(np.arange(-radius, radius + 1), np.arange(-radius, radius + 1), indexing="ij")
    ii = ii.ravel()
    jj = jj.ravel()
    fi, fj = ii.astype(F), jj.astype(F)
    c_rot = (fj * cos_t - fi * sin_t).astype(F)
    r_rot = (fj * sin_t + fi * cos_t).astype(F)
    rbin = (r_rot + F(d // 2) - F(0.5)).astype(F)
    cbin = (c_rot + F(d // 2) - F(0.5)).astype(F)
    r = py + ii
    c = px + jj
    ok = (rbin > -1) & (rbin < d) & (cbin > -1) & (cbin < d) & (r > 0) & (r < h - 1) & (c > 0) & (c < w - 1)
    r, c, rbin, cbin, c_rot, r_rot = r[ok], c[ok], rbin[ok], cbin[ok], c_rot[ok], r_rot[ok]
    dx = img[r, c + 1] - img[r, c - 1]
    dy = img[r - 1, c] - img[r + 1, c]
    wgt = np.exp(((c_rot * c_rot + r_rot * r_rot) * exp_scale).astype(F)).astype(F)
    o = fast_atan2(dy, dx)
    mag = (np.sqrt((dx * dx + dy * dy).astype(F)).astype(F) * wgt).astype(F)
    obin = ((o - F(ori)) * bins_per_rad).astype(F)
    r0 = np.floor(rbin).astype(np.int64)
    c0 = np.floor(cbin).astype(np.int64)
    o0 = np.floor(obin).astype(np.int64)
    rbin = (rbin - r0).astype(F)
    cbin = (cbin - c0).astype(F)
    obin = (obin - o0).astype(F)
    o0 = np.where(o0 < 0, o0 + n, o0)
    o0 = np.where(o0 >= n, o0 - n, o0)
    v_r1 = mag * rbin
    v_r0 = mag - v_r1
    v_rc11 = v_r1 * cbin
    v_rc10 = v_r1 - v_rc11
    v_rc01 = v_r0 * cbin
    v_rc00 = v_r0 - v_rc01
    v111 = v_rc11 * obin
    v110 = v_rc11 - v111
    v101 = v_rc10 * obin
    v100 = v_rc10 - v101
    v011 = v_rc01 * obin
    v010 = v_rc01 - v011
    v001 = v_rc00 * obin
    v000 = v_rc00 - v001
    hist = np.zeros((d + 2) * (d + 2) * (n + 2), F)
    idx = ((r0 + 1) * (d + 2) + c0 + 1) * (n + 2) + o0
    s1, s2 = n + 2, (d + 2) * (n + 2)
    for k in range(len(idx)):           # sequential float accumulation in sample order
        i = idx[k]
        hist[i] += v000[k]
        hist[i + 1] += v001[k]
        hist[i + s1] += v010[k]
        hist[i + s1 + 1] += v011[k]
        hist[i + s2] += v100[k]
        hist[i + s2 + 1] += v101[k]
        hist[i + s2 + s1] += v110[k]
        hist[i + s2 + s1 + 1] += v111[k]
    raw = np.zeros(d * d * n, F)
    for i in range(d):
        for j in range(d):
            b = ((i + 1) * (d + 2) + (j + 1)) * (n + 2)
            hist[b] += hist[b + n]
            hist[b + 1] += hist[b + n + 1]
            raw[(i * d + j) * n:(i * d + j + 1) * n] = hist[b:b + n]
    nrm2 = F(0)
    for v in raw:
        nrm2 += v * v
    thr = np.sqrt(nrm2, dtype=F) * F(DESCR_MAG_THR)
    raw = np.minimum(raw, thr)
    nrm2 = F(0)
    for v in raw:
        nrm2 += v * v
    scale = F(INT_DESCR_FCTR) / max(np.sqrt(nrm2, dtype=F), F(FLT_EPSILON))
    return np.clip(np.rint(raw * scale), 0, 255).astype(F)


# ---------------------------------------------------------------------------- the detector
def detect_and_compute(image, with_pyramid=False):
    """cv2.SIFT_create().detectAndCompute(image, None) -> (kp float32 [n, 6] = x, y, size, angle,
    response, octave word; desc float32 [n, 128]), in cv2's output order (sorted by
    KeyPoint12_LessThan, duplicates removed)."""
    img = np.asarray(image)
    if img.ndim == 3:
        img = bgr_to_gray(img)
    first_octave = -1
    n_oct = n_octaves(img.shape[0], img.shape[1], first_octave)
    gpyr = gaussian_pyramid(initial_image(img), n_oct)
    dog = dog_pyramid(gpyr, n_oct)
    L = N_OCTAVE_LAYERS + 3
    kps = []
    for o in range(n_oct):
        for layer in range(1, N_OCTAVE_LAYERS + 1):
            rr, cc = extrema_candidates(dog, o, layer)
            for r, c in zip(rr.tolist(), cc.tolist()):
                k = adjust_local_extrema(dog, o, layer, r, c)
                if k is None:
                    continue
                scl_octv = k["size"] * F(0.5) / F(1 << o)
                hist, omax = orientation_hist(gpyr[o * L + k["layer"]], k["c"], k["r"],
                                              cv_round(F(ORI_RADIUS) * scl_octv),
                                              F(ORI_SIG_FCTR) * scl_octv)
                for ang in orientations(hist, omax):
                    kps.append((k["x"], k["y"], k["size"], ang, k["response"], k["octave"]))
    # KeyPointsFilter::removeDuplicatedSorted
    kps.sort(key=lambda k: (k[0], k[1], -k[2], k[3], -k[4], -k[5]))
    uniq = []
    for k in kps:
        if uniq and uniq[-1][0] == k[0] and uniq[-1][1] == k[1] and uniq[-1][2] == k[2] and uniq[-1][3] == k[3]:
            continue
        uniq.append(k)
    # firstOctave = -1: octave byte + firstOctave, pt and size halved
    out = np.zeros((len(uniq), 6), np.float64)
    desc = np.zeros((len(uniq), 128), F)
    for i, (x, y, size, ang, resp, octw) in enumerate(uniq):
        octw = (octw & ~255) | ((octw + first_octave) & 255)
        x, y, size = F(x) * F(0.5), F(y) * F(0.5), F(size) * F(0.5)
        out[i] = (x, y, size, ang, resp, octw)
        # calcDescriptors: unpackOctave, back to the octave's own pixel grid
        octave = octw & 255
        layer = (octw >> 8) & 255
        octave = octave if octave < 128 else (-128 | octave)
        scale = F(1.0 / (1 << octave)) if octave >= 0 else F(1 << -octave)
        angle = F(360) - F(ang)
        if abs(angle - F(360)) < FLT_EPSILON:
            angle = F(0)
        g = gpyr[(octave - first_octave) * L + layer]
        desc[i] = descriptor(g, x * scale, y * scale, angle, size * scale * F(0.5))
    if with_pyramid:
        return out, desc, gpyr, dog
    return out, desc
