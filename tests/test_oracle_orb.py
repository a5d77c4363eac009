"""CPU: the ORB front-end restatement (oracle/cv_orb.py) against the cv2 binary and
against the committed golden vectors (tests/golden/orbdet_*.npz, made by cv2 in the build
container).  Every stage is pinned separately, then the whole detector."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import cv_orb
from visual_odometry_b200 import synth

cv2 = pytest.importorskip("cv2")
PATTERN = np.load(os.path.join(GOLDEN, "orb_pattern_31.npy"))
ORBDET = ["orbdet_480x640_500", "orbdet_1080x1920_2000", "orbdet_240x320_all_1level",
          "orbdet_150x200_300_thr10"]


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    h, w, seed, nf, nl, edge, thr = (int(v) for v in z["params"])
    img = synth.textured_image(h, w, seed)
    assert int(img.astype(np.uint64).sum()) == int(z["image_crc"][0]), "scene generator changed"
    return img, dict(nfeatures=nf, nlevels=nl, edge_threshold=edge, fast_threshold=thr), z


def test_pattern_is_the_published_table():
    # first and last rows of OpenCV's bit_pattern_31_ as printed in the ORB literature
    assert PATTERN.shape == (256, 4) and PATTERN.dtype == np.int8
    assert PATTERN[0].tolist() == [8, -3, 9, 5] and PATTERN[1].tolist() == [4, 2, 7, -12]
    assert np.abs(PATTERN).max() == 13


@pytest.mark.parametrize("thr,nms", [(20, True), (10, True), (35, False)])
def test_fast_equals_cv2(thr, nms):
    img = synth.textured_image(300, 400, 3)
    kps = cv2.FastFeatureDetector_create(thr, nms).detect(img)
    xy, score = cv_orb.fast9_detect(img, thr, nms)
    assert len(kps) > 50
    assert np.array_equal(np.array([k.pt for k in kps], np.int32), xy)
    if nms:      # without suppression cv2 does not compute the score (response 0)
        assert np.array_equal(np.array([k.response for k in kps], np.int32), score)


@pytest.mark.parametrize("name", ORBDET)
def test_fast_equals_golden(name):
    img, p, z = load(name)
    xy, score = cv_orb.fast9_detect(img, p["fast_threshold"], True)
    assert np.array_equal(z["fast"][:, :2].astype(np.int32), xy)
    assert np.array_equal(z["fast"][:, 2].astype(np.int32), score)


def test_fast_atan2_equals_cv2():
    rng = np.random.default_rng(5)
    for _ in range(3000):
        y, x = float(rng.integers(-300000, 300000)), float(rng.integers(-300000, 300000))
        assert np.float32(cv2.fastAtan2(y, x)) == cv_orb.fast_atan2(y, x)
    assert cv_orb.fast_atan2(0, 0) == 0


def test_resize_and_blurs_equal_cv2():
    img = synth.textured_image(270, 480, 4)
    for (w2, h2) in cv_orb.level_sizes(img.shape, 5)[1:]:
        assert np.array_equal(cv2.resize(img, (w2, h2), interpolation=cv2.INTER_LINEAR_EXACT),
                              cv_orb.resize_linear_exact(img, w2, h2))
    kf = cv2.getGaussianKernel(7, 2, cv2.CV_32F).ravel()
    assert np.array_equal(kf, cv_orb.gaussian_taps7())
    assert np.array_equal(cv2.sepFilter2D(img, cv2.CV_8U, kf, kf, borderType=cv2.BORDER_REFLECT_101),
                          cv_orb.gaussian_blur7(img))
    assert np.array_equal(cv2.GaussianBlur(img, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101),
                          cv_orb.gaussian_blur_fixed7(img))


def test_gray_conversion_equals_cv2():
    bgr = np.random.default_rng(6).integers(0, 256, (64, 80, 3), dtype=np.uint8)
    b = bgr.astype(np.int64)
    mine = ((b[..., 0] * 3735 + b[..., 1] * 19235 + b[..., 2] * 9798 + (1 << 14)) >> 15).astype(np.uint8)
    assert np.array_equal(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), mine)


def test_describe_equals_cv2_compute_with_given_angles():
    rng = np.random.default_rng(7)
    img = synth.textured_image(240, 320, 8)
    kps = [cv2.KeyPoint(float(rng.integers(40, 280)), float(rng.integers(40, 200)), 31.0,
                        float(rng.uniform(0, 360)), 1.0, 0) for _ in range(300)]
    kps2, d = cv2.ORB_create(nfeatures=10, nlevels=1).compute(img, kps)
    assert len(kps2) == len(kps)
    mine = cv_orb.describe(cv_orb.gaussian_blur7(img), np.array([k.pt for k in kps]),
                           np.array([k.angle for k in kps], np.float32), PATTERN)
    assert np.array_equal(mine, d)


def _check(r, kp, desc):
    mine = np.stack([r["pt"][:, 0], r["pt"][:, 1], r["size"], r["angle"], r["response"],
                     r["octave"].astype(np.float32)], 1).astype(np.float32)
    assert mine.shape == kp.shape
    assert np.array_equal(mine, kp)            # bit-exact, including angle and Harris response
    assert np.array_equal(r["desc"], desc)


@pytest.mark.parametrize("name", ORBDET)
def test_detector_equals_golden(name):
    img, p, z = load(name)
    _check(cv_orb.orb_detect_and_compute(img, PATTERN, **p), z["kp"], z["desc"])


def test_detector_equals_live_cv2():
    img = synth.textured_image(360, 500, 21)
    for nf, nl in ((300, 8), (1500, 5)):
        kps, d = cv2.ORB_create(nfeatures=nf, nlevels=nl).detectAndCompute(img, None)
        kp = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps], np.float32)
        o = np.lexsort((kp[:, 0], kp[:, 1], kp[:, 5]))
        _check(cv_orb.orb_detect_and_compute(img, PATTERN, nfeatures=nf, nlevels=nl), kp[o], d[o])


def test_level_budget():
    assert cv_orb.features_per_level(500, 8) == [109, 90, 75, 63, 52, 44, 36, 31]
    assert sum(cv_orb.features_per_level(2000, 8)) == 2000


def test_library_level_geometry_equals_oracle_without_a_device():
    """vo_orb_level_info is host-side logic of the C-ABI (pyramid sizes, per-level feature budget,
    float32 layer scales as cv2 computes them): checked here without a GPU."""
    from visual_odometry_b200.orb import OrbDetector
    for (h, w, nf, nl) in ((1080, 1920, 2000, 8), (480, 640, 500, 8), (97, 131, 50, 3), (2160, 3840, 5000, 12)):
        info = OrbDetector(nfeatures=nf, nlevels=nl, device="cpu").level_info(w, h)
        assert list(zip(info["w"], info["h"])) == cv_orb.level_sizes((h, w), nl)
        assert info["nfeatures"] == cv_orb.features_per_level(nf, nl)
        assert np.array_equal(np.array(info["scale"], np.float32), cv_orb.level_scales(nl))
    from visual_odometry_b200 import _lib
    lib = _lib.load()
    assert lib.vo_orb_workspace_bytes(4, 1920, 1080, 2000, 1.2, 8, 2256) > 4 * 1920 * 1080
    assert lib.vo_orb_workspace_bytes(1, 5000, 100, 10, 1.2, 8, 100) == 0      # wider than 4095
    rc = lib.vo_orb_detect_batch(None, 1, 640, 480, 500, 1.2, 8, 31, 20, 600, None, None, None, None, None, 0, None)
    assert rc == _lib.VO_ERR_INVALID_ARG and b"null" in lib.vo_last_error_string()
