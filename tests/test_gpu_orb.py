"""GPU: ORB feature extraction (csrc/orb.cu behind vo_orb_detect_batch) against the
cv2 golden vectors and the oracle restatement - key-point records (position, size, angle, Harris
response, octave) and descriptors BIT-EXACT, in (octave, y, x) order."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from visual_odometry_b200 import synth
from visual_odometry_b200.frame import FrameGenerator
from visual_odometry_b200.orb import OrbDetector

pytestmark = pytest.mark.gpu
ORBDET = ["orbdet_480x640_500", "orbdet_1080x1920_2000", "orbdet_240x320_all_1level",
          "orbdet_150x200_300_thr10"]
PATTERN = np.load(os.path.join(GOLDEN, "orb_pattern_31.npy"))


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    h, w, seed, nf, nl, edge, thr = (int(v) for v in z["params"])
    img = synth.textured_image(h, w, seed)
    assert int(img.astype(np.uint64).sum()) == int(z["image_crc"][0])
    return img, dict(nfeatures=nf, nlevels=nl, edgeThreshold=edge, fastThreshold=thr), z


def run(det, images, cap=None):
    out = det.detect_batch(images, cap=cap)
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    return [(out["kp"][f, :n[f]].cpu().numpy(), out["desc"][f, :n[f]].cpu().numpy())
            for f in range(len(n))], out["status"].cpu().numpy()


@pytest.mark.parametrize("name", ORBDET)
def test_equals_cv2_golden(cuda_dev, name):
    img, p, z = load(name)
    (kp, desc), = run(OrbDetector(**p), img[None])[0]
    assert kp.shape == z["kp"].shape
    assert np.array_equal(kp, z["kp"])
    assert np.array_equal(desc, z["desc"])


@pytest.mark.parametrize("h,w,seed,nf,nl,thr", [(200, 260, 31, 150, 8, 20), (333, 517, 32, 4000, 6, 7),
                                                (97, 131, 33, 50, 3, 20), (720, 1280, 34, 1000, 8, 20)])
def test_equals_oracle(cuda_dev, h, w, seed, nf, nl, thr):
    from oracle import cv_orb
    img = synth.textured_image(h, w, seed)
    r = cv_orb.orb_detect_and_compute(img, PATTERN, nfeatures=nf, nlevels=nl, fast_threshold=thr)
    ref = np.stack([r["pt"][:, 0], r["pt"][:, 1], r["size"], r["angle"], r["response"],
                    r["octave"].astype(np.float32)], 1).astype(np.float32)
    (kp, desc), = run(OrbDetector(nfeatures=nf, nlevels=nl, fastThreshold=thr), img[None], cap=nf + 512)[0]
    assert np.array_equal(kp, ref)
    assert np.array_equal(desc, r["desc"])


def test_level_geometry_equals_oracle():
    from oracle import cv_orb
    for (h, w, nf, nl) in ((1080, 1920, 2000, 8), (480, 640, 500, 8), (97, 131, 50, 3)):
        info = OrbDetector(nfeatures=nf, nlevels=nl).level_info(w, h)
        assert list(zip(info["w"], info["h"])) == cv_orb.level_sizes((h, w), nl)
        assert info["nfeatures"] == cv_orb.features_per_level(nf, nl)
        assert np.array_equal(np.array(info["scale"], np.float32), cv_orb.level_scales(nl))


def test_batch_equals_single_frames_and_bgr(cuda_dev):
    imgs = np.stack([synth.textured_image(240, 320, 40 + k) for k in range(5)])
    det = OrbDetector(nfeatures=300)
    batch, status = run(det, imgs)
    assert not status.any()
    for k in range(5):
        (kp, desc), = run(det, imgs[k][None])[0]
        assert np.array_equal(kp, batch[k][0]) and np.array_equal(desc, batch[k][1])
    # BGR input goes through cv2's fixed-point gray conversion on the device
    cv2 = pytest.importorskip("cv2")
    bgr = np.stack([np.roll(imgs[0], 3, 1), imgs[0], np.roll(imgs[0], 5, 0)], -1)
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    (a, da), = run(det, bgr[None])[0]
    (b, db), = run(det, gray[None])[0]
    assert np.array_equal(a, b) and np.array_equal(da, db)


def test_edge_cases(cuda_dev):
    # image smaller than twice the edge threshold: nothing survives the border filter
    (kp, desc), = run(OrbDetector(nfeatures=100), synth.textured_image(60, 60, 50)[None])[0]
    assert len(kp) == 0 and desc.shape == (0, 32)
    # flat image: no corners
    (kp, desc), = run(OrbDetector(nfeatures=100), np.full((1, 200, 200), 128, np.uint8))[0]
    assert len(kp) == 0
    # capacity overflow is reported, not silent
    img = synth.textured_image(240, 320, 51)
    _, status = run(OrbDetector(nfeatures=200), img[None], cap=50)
    assert status[0] == 1
    # bad geometry fails loudly
    from visual_odometry_b200 import _lib
    with pytest.raises(_lib.VoError):
        OrbDetector(nfeatures=10, nlevels=40).detect_batch(img[None])


def test_detect_and_compute_drop_in(cuda_dev):
    """The reference's FrameGenerator.make_frame (src/frame_generator.py:10-38) with the GPU
    detector in place of cv2.ORB_create(): same key points and descriptors as cv2 (as sets)."""
    cv2 = pytest.importorskip("cv2")
    img = synth.textured_image(480, 640, 60)
    frame = FrameGenerator(OrbDetector(nfeatures=500)).make_frame(img)
    kps, d = cv2.ORB_create(nfeatures=500).detectAndCompute(img, None)
    ref = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps], np.float32)
    o = np.lexsort((ref[:, 0], ref[:, 1], ref[:, 5]))
    mine = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in frame.keypoints],
                    np.float32)
    assert np.array_equal(mine, ref[o])
    assert np.array_equal(frame.descriptors, d[o])
    assert frame.features[3].feature_id == (0, 3)


def test_images_to_matches_through_the_reference_surface(cuda_dev):
    """Two overlapping views -> FrameGenerator(OrbDetector) -> ImagePair.match_features with the
    B200 matcher: the same matches (as point pairs with distances) as cv2.ORB_create +
    cv2.BFMatcher(NORM_HAMMING, crossCheck=True) on the same images (reference
    src/visual_slam.py:16-18,294; src/image_pair.py:233-258).  Index-based tie-breaks can differ
    because cv2 leaves an octave's key points in std::nth_element order."""
    cv2 = pytest.importorskip("cv2")
    from visual_odometry_b200.image_pair import ImagePair
    from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING
    big = synth.textured_image(700, 900, 70)
    img1, img2 = np.ascontiguousarray(big[60:540, 80:720]), np.ascontiguousarray(big[71:551, 97:737])
    K = np.array([[700.0, 0, 320], [0, 700.0, 240], [0, 0, 1]])
    gen = FrameGenerator(OrbDetector(nfeatures=800))
    f1, f2 = gen.make_frame(img1), gen.make_frame(img2)
    pair = ImagePair(f1, f2, B200Matcher(NORM_HAMMING, crossCheck=True), K)
    pair.match_features()
    mine = {(m.keypoint1, m.keypoint2, float(m.distance)) for m in pair.filtered_matches}
    orb = cv2.ORB_create(nfeatures=800)
    k1, d1 = orb.detectAndCompute(img1, None)
    k2, d2 = orb.detectAndCompute(img2, None)
    ref = {(k1[m.queryIdx].pt, k2[m.trainIdx].pt, float(m.distance))
           for m in cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d1, d2)}
    assert len(ref) > 300
    assert len(mine & ref) >= 0.995 * len(ref) and abs(len(mine) - len(ref)) <= 0.005 * len(ref) + 1
    # most of them are the true shift (17, 11) of the second crop
    shift = np.array([np.subtract(a, b) for a, b, _ in mine])
    assert (np.abs(shift - [17, 11]).max(1) < 2.0).mean() > 0.5
