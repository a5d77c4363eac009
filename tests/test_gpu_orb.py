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


def _render_two_plane_sequence(n_frames, W=640, H=480):
    """A far textured plane (Z = 14) with two nearer ones (Z = 8, Z = 5) in front of parts of it, seen by a
    camera that translates and yaws a little per frame; views rendered with cv2.warpPerspective.
    -> (images u8 [F,H,W], K, list of (R_k, C_k))."""
    import cv2
    K = np.array([[700.0, 0, W / 2], [0, 700.0, H / 2], [0, 0, 1]])
    far, mid, near = (synth.textured_image(1600, 2200, 81), synth.textured_image(520, 520, 82),
                      synth.textured_image(400, 400, 83))
    planes = [(far, 0.014, 1100, 800, 14.0, 0.0, 0.0), (mid, 0.007, 260, 260, 8.0, -1.6, 0.4),
              (near, 0.005, 200, 200, 5.0, 1.3, -0.5)]
    images, cams = [], []
    for k in range(n_frames):
        yaw = 0.012 * k
        R = np.array([[np.cos(yaw), 0, np.sin(yaw)], [0, 1, 0], [-np.sin(yaw), 0, np.cos(yaw)]])
        C = np.array([0.35 * k, 0.05 * k, 0.08 * k])
        t = -R @ C
        img = None
        for tex, s, cx, cy, Z0, ox, oy in planes:
            M = np.array([[s, 0, -cx * s + ox], [0, s, -cy * s + oy], [0, 0, Z0]])
            Hm = K @ (R @ M + np.outer(t, [0, 0, 1.0]))
            layer = cv2.warpPerspective(tex, Hm, (W, H), flags=cv2.INTER_LINEAR)
            cover = cv2.warpPerspective(np.full_like(tex, 255), Hm, (W, H), flags=cv2.INTER_NEAREST) > 0
            img = layer if img is None else np.where(cover, layer, img)
        images.append(img)
        cams.append((R, C))
    return np.stack(images), K, cams


def test_image_sequence_to_poses_on_the_device(cuda_dev):
    """Images -> ORB extraction -> Hamming matching -> E-RANSAC -> recoverPose without leaving the
    device (PairPipeline.run_images); the relative poses follow the ground truth of the rendered
    sequence, and so does cv2's own chain on the same images."""
    cv2 = pytest.importorskip("cv2")
    from visual_odometry_b200.pipeline import PairPipeline
    images, K, cams = _render_two_plane_sequence(5)
    pipe = PairPipeline(1500, 4, 5, K=K, device="cuda:0")
    b = pipe.run_images(images, OrbDetector(nfeatures=1200))
    torch.cuda.synchronize()
    assert b == 4
    Rg, tg = pipe.R[:b].cpu().numpy().reshape(b, 3, 3), pipe.t[:b].cpu().numpy().reshape(b, 3)
    n_inl = pipe.n_inl[:b].cpu().numpy()
    orb, bf = cv2.ORB_create(nfeatures=1200), cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)
    for k in range(b):
        (R1, C1), (R2, C2) = cams[k], cams[k + 1]
        R_true = R2 @ R1.T
        t_true = R2 @ (C1 - C2)
        t_true /= np.linalg.norm(t_true)
        ang = np.degrees(np.arccos(np.clip((np.trace(Rg[k].T @ R_true) - 1) / 2, -1, 1)))
        tdir = np.degrees(np.arccos(np.clip(tg[k] @ t_true, -1, 1)))
        assert n_inl[k] > 150, (k, n_inl[k])
        # (1 px RANSAC without a refit on ORB's integer pyramid coordinates: cv2's own chain is
        # within 1-3 degrees / 3-12 degrees of the truth on this scene)
        assert ang < 4.0 and tdir < 20.0, (k, ang, tdir)
        # cv2's own chain on the same two images lands on the same motion
        k1, d1 = orb.detectAndCompute(images[k], None)
        k2, d2 = orb.detectAndCompute(images[k + 1], None)
        m = bf.match(d1, d2)
        assert abs(len(m) - int(pipe.n_match[k].item())) <= 0.01 * len(m) + 2
        p1 = np.float64([k1[x.queryIdx].pt for x in m]); p2 = np.float64([k2[x.trainIdx].pt for x in m])
        E, mask = cv2.findEssentialMat(p1, p2, K, cv2.FM_RANSAC, 0.99, 1)
        _, Rc, tc, _ = cv2.recoverPose(E, p1[mask.ravel() == 1], p2[mask.ravel() == 1], K)
        ang_c = np.degrees(np.arccos(np.clip((np.trace(Rg[k].T @ Rc) - 1) / 2, -1, 1)))
        assert ang_c < 5.0, (k, ang_c)


@pytest.mark.parametrize("kind,h,w,nf,nl,thr", [("noise", 480, 640, 500, 8, 20), ("noise", 300, 401, 3000, 5, 40),
                                                ("texture", 1080, 1920, 500, 8, 20), ("texture", 2160, 3840, 5000, 8, 20),
                                                ("texture", 333, 517, 1000, 8, 7), ("blocks", 256, 256, 200, 4, 20)])
def test_equals_live_cv2(cuda_dev, kind, h, w, nf, nl, thr):
    """Against cv2 itself on the box (not a stored golden): noise images have tens of thousands of
    FAST corners and heavy score ties at the first cut, the block image has exactly repeated
    structure (tied Harris responses), 4K is the largest size the packed corner words hold."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(h * w + nf)
    if kind == "noise":
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    elif kind == "blocks":
        img = np.kron(rng.integers(0, 2, (h // 32, w // 32)) * 120 + 60, np.ones((32, 32))).astype(np.uint8)
        img[:, w // 2:] = img[:, :w // 2]          # exact repetition -> tied responses
    else:
        img = synth.textured_image(h, w, h + w)
    kps, d = cv2.ORB_create(nfeatures=nf, nlevels=nl, fastThreshold=thr).detectAndCompute(img, None)
    ref = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps],
                   np.float32).reshape(-1, 6)
    d = np.zeros((0, 32), np.uint8) if d is None else d
    o = np.lexsort((ref[:, 0], ref[:, 1], ref[:, 5]))
    res, status = run(OrbDetector(nfeatures=nf, nlevels=nl, fastThreshold=thr), img[None], cap=len(ref) + 64)
    assert status[0] == 0
    kp, desc = res[0]
    assert kp.shape == ref.shape
    assert np.array_equal(kp, ref[o])
    assert np.array_equal(desc, d[o])


def test_legacy_two_image_demo_from_files(cuda_dev, tmp_path):
    """TriangulatePointsFromTwoImages.run (reference src/triangulate_points_from_images.py:17-41,
    what main_triangulate.py calls) on two image FILES: cv2.imread -> ImageAndKeypoints("ORB")
    (GPU detector through the cv2 shim) -> legacy ImagePair("ORB").standard_pipeline -> the
    match visualisation written to disk."""
    cv2 = pytest.importorskip("cv2")
    from visual_odometry_b200.legacy import TriangulatePointsFromTwoImages
    from visual_odometry_b200.orb import OrbDetector as _Det
    images, K, cams = _render_two_plane_sequence(2)
    f1, f2, out = str(tmp_path / "a.png"), str(tmp_path / "b.png"), str(tmp_path / "m.png")
    cv2.imwrite(f1, cv2.cvtColor(images[0], cv2.COLOR_GRAY2BGR))
    cv2.imwrite(f2, cv2.cvtColor(images[1], cv2.COLOR_GRAY2BGR))
    tp = TriangulatePointsFromTwoImages()
    tp.run(f1, f2, output=out)
    pair = tp.pair
    assert isinstance(pair.image1.detector, _Det)
    assert len(pair.image1.keypoints) == 500 and pair.image1.descriptors.shape == (500, 32)
    assert pair.R.shape == (3, 3) and abs(np.linalg.det(pair.R) - 1) < 1e-9
    assert abs(np.linalg.norm(pair.t) - 1) < 1e-9
    assert pair.points3d_reconstr.shape[0] == 4 and pair.points3d_reconstr.shape[1] > 50
    assert os.path.getsize(out) > 10000


def test_slam_loop_from_images(cuda_dev):
    """DeviceSlam.run_images: rendered frames -> ORB extraction -> hot path -> map stage (PnP,
    map growth, bundle adjustment, pruning) without descriptors or key points touching the host."""
    pytest.importorskip("cv2")
    from visual_odometry_b200.slam import DeviceSlam
    images, K, cams = _render_two_plane_sequence(6)
    slam = DeviceSlam(K, 1500, 6, descriptor="orb", ba_iterations=10)
    log = slam.run_images(images, OrbDetector(nfeatures=1200))
    assert len(log) == 5
    assert log[0]["n_points"] > 100 and log[0]["n_obs"] >= 2 * log[0]["n_points"] - 5
    ok = [e["pnp_ok"] for e in log[1:]]
    assert sum(bool(x) for x in ok) >= 3, log
    assert log[-1]["n_points"] > log[0]["n_points"] // 2
    assert len(log[-1]["cameras"]) >= 4
