"""CPU: oracle/cv_sift.py (the restatement of cv2.SIFT_create().detectAndCompute that the GPU
tests use as checker) pinned against the cv2 binary, stage by stage and as a whole, and against
the golden vectors tests/golden/siftdet_*.npz (cv2 outputs, tests/golden/make_golden_siftdet.py).

Tolerances (float work): the Gaussian pyramid is BIT-exact; key points: same count, same order,
same octave words, |x|,|y| <= 1e-3 px, size <= 1e-4, angle <= 1e-3 deg, response <= 1e-6;
descriptors: identical integers.
"""
import os

import numpy as np
import pytest

from oracle import cv_sift as S
from visual_odometry_b200 import synth

cv2 = pytest.importorskip("cv2")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    h, w, seed = (int(v) for v in z["hws"])
    return synth.textured_image(h, w, seed), z["kp"], z["octave"], z["desc"]


@pytest.mark.parametrize("shape", [(97, 163), (50, 166), (33, 37), (128, 160), (61, 3), (6, 5)])
def test_blur_is_bit_exact_including_the_scalar_tails(shape):
    h, w = shape
    img = synth.textured_image(max(h, 16), max(w, 16), 5)[:h, :w].astype(np.float32)
    for sigma in [1.2489996] + S.layer_sigmas()[1:]:
        assert np.array_equal(cv2.GaussianBlur(img, (0, 0), sigma), S.blur(img, sigma))


def test_resizes_are_exact():
    img = synth.textured_image(121, 163, 5).astype(np.float32)
    up = cv2.resize(img, (326, 242), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(up, S.upsample2_linear(img))
    for shp in ((242, 326), (121, 163), (60, 81), (7, 9)):
        x = np.random.default_rng(0).random(shp).astype(np.float32)
        ref = cv2.resize(x, (shp[1] // 2, shp[0] // 2), interpolation=cv2.INTER_NEAREST)
        assert np.array_equal(ref, S.halve_nearest(x))


def test_fast_atan2_equals_cv2():
    rng = np.random.default_rng(1)
    x = rng.normal(size=4000).astype(np.float32)
    y = rng.normal(size=4000).astype(np.float32)
    x[:10] = 0
    y[5:15] = 0
    ref = cv2.phase(x, y, angleInDegrees=True).ravel()
    assert np.abs(S.fast_atan2(y, x) - ref).max() <= 1e-4


def test_octave_count_and_layer_sigmas():
    assert S.n_octaves(480, 640) == 8 and S.n_octaves(120, 160) == 6
    sig = S.layer_sigmas()
    assert len(sig) == 6 and abs(sig[1] - 1.2262735) < 1e-6 and abs(sig[5] - 3.0900155) < 1e-6


def _check(kp, desc, gkp, goct, gdesc):
    assert len(kp) == len(gkp)
    assert np.array_equal(kp[:, 5].astype(np.int64), goct.astype(np.int64))
    d = np.abs(kp[:, :5] - gkp.astype(np.float64)).max(0)
    assert d[0] <= 1e-3 and d[1] <= 1e-3 and d[2] <= 1e-4 and d[3] <= 1e-3 and d[4] <= 1e-6
    assert np.array_equal(desc.astype(np.int64), gdesc.astype(np.int64))


@pytest.mark.parametrize("name", ["siftdet_120x160", "siftdet_243x325"])
def test_detector_equals_golden(name):
    img, gkp, goct, gdesc = load(name)
    kp, desc = S.detect_and_compute(img)
    _check(kp, desc, gkp, goct, gdesc)


def test_detector_equals_live_cv2_on_a_bgr_image():
    g = synth.textured_image(150, 200, 21)
    img = np.stack([g, np.roll(g, 3, 0), np.roll(g, 5, 1)], -1)
    kps, rdesc = cv2.SIFT_create().detectAndCompute(img, None)
    gkp = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response) for k in kps], np.float32)
    goct = np.array([k.octave for k in kps])
    kp, desc = S.detect_and_compute(img)
    assert len(kps) > 20
    _check(kp, desc, gkp, goct, rdesc)
