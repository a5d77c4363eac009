"""GPU: SIFT feature extraction (csrc/sift.cu behind vo_sift_detect_batch) against
the cv2 golden vectors, the cv2 binary on the box and the oracle restatement.

Tolerances (float work, written here): the Gaussian pyramid is BIT-exact against the oracle (which
is bit-exact against cv2.GaussianBlur, tests/test_oracle_sift.py).  Key points: the same number, in
cv2's order, identical octave words, |x|,|y| <= 2e-3 px, size <= 1e-3, angle <= 0.02 deg,
response <= 1e-6 on at least 99 % of cv2's key points (a decision next to a threshold may flip
for the rest; measured: 100 % on every scene here); descriptors: at most 1 level apart in at most
1 % of the values of the common key points (measured: identical).
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, record_parity
from oracle import cv_sift as S
from visual_odometry_b200 import synth
from visual_odometry_b200.frame import FrameGenerator
from visual_odometry_b200.sift import SiftDetector

pytestmark = pytest.mark.gpu


def run(det, images, cap=None, pyramid=False):
    out = det.detect_batch(images, cap=cap, with_pyramid=pyramid)
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    res = [(out["kp"][f, :n[f]].cpu().numpy().astype(np.float64), out["desc"][f, :n[f]].cpu().numpy())
           for f in range(len(n))]
    return res, out["status"].cpu().numpy(), out


def check(kp, desc, ref, rdesc, record=None, name=""):
    """kp [n,6] / desc against cv2's (ref [m,6] float64, rdesc [m,128]) with the tolerances above"""
    kp, ref = np.asarray(kp, np.float64).reshape(-1, 6), np.asarray(ref, np.float64).reshape(-1, 6)
    ia, ib = [], []
    for j in range(len(ref)):          # pair every cv2 key point with ours at the same place
        if not len(kp):
            break
        dang = np.abs(kp[:, 3] - ref[j, 3])
        hit = np.nonzero((np.abs(kp[:, 0] - ref[j, 0]) < 0.05) & (np.abs(kp[:, 1] - ref[j, 1]) < 0.05)
                         & (np.minimum(dang, 360 - dang) < 0.5) & (kp[:, 5] == ref[j, 5]))[0]
        if len(hit):
            ia.append(int(hit[0]))
            ib.append(j)
    ia, ib = np.array(ia, np.int64), np.array(ib, np.int64)
    common = ia
    assert len(common) >= 0.99 * len(ref) and abs(len(kp) - len(ref)) <= 0.01 * len(ref) + 1
    assert np.all(np.diff(ia) > 0) == np.all(np.diff(ib) > 0)          # the same (cv2's) order
    d = np.abs(kp[ia, :5] - ref[ib, :5])
    d[:, 3] = np.minimum(d[:, 3], 360 - d[:, 3])
    good = (d[:, 0] <= 2e-3) & (d[:, 1] <= 2e-3) & (d[:, 2] <= 1e-3) & (d[:, 3] <= 0.02) & (d[:, 4] <= 1e-6)
    assert good.mean() >= 0.99
    dd = np.abs(desc[ia][good].astype(np.int32) - np.asarray(rdesc)[ib][good].astype(np.int32))
    assert dd.max(initial=0) <= 1 and (dd > 0).mean() <= 0.01
    if record is not None:
        record("sift_" + name, key_points=len(ref), common=len(common), within_tolerance=int(good.sum()),
               identical_descriptor_rows=int((dd.max(1) == 0).sum()) if len(dd) else 0,
               identical_records=int((d.max(1) == 0).sum()))


def cv_ref(img):
    import cv2
    kps, desc = cv2.SIFT_create().detectAndCompute(img, None)
    ref = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps],
                   np.float64).reshape(-1, 6)
    return ref, (np.zeros((0, 128), np.float32) if desc is None else desc)


@pytest.mark.parametrize("name", ["siftdet_120x160", "siftdet_243x325", "siftdet_480x640"])
def test_equals_cv2_golden(cuda_dev, name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    h, w, seed = (int(v) for v in z["hws"])
    img = synth.textured_image(h, w, seed)
    res, status, _ = run(SiftDetector(), img[None])
    assert status[0] == 0
    ref = np.concatenate([z["kp"].astype(np.float64), z["octave"][:, None].astype(np.float64)], 1)
    check(res[0][0], res[0][1], ref, z["desc"], record_parity, name)


@pytest.mark.parametrize("h,w,seed", [(96, 128, 1), (243, 325, 9), (150, 201, 4)])
def test_pyramid_is_bit_exact_and_key_points_equal_the_oracle(cuda_dev, h, w, seed):
    img = synth.textured_image(h, w, seed)
    res, status, out = run(SiftDetector(), img[None], pyramid=True)
    gp = S.gaussian_pyramid(S.initial_image(img), S.n_octaves(h, w))
    assert [g.shape[::-1] for g in gp[::6]] == SiftDetector.octaves(w, h)
    assert np.array_equal(out["pyramid"][0].cpu().numpy(), np.concatenate([g.ravel() for g in gp]))
    okp, odesc = S.detect_and_compute(img)
    check(res[0][0], res[0][1], okp, odesc)


@pytest.mark.parametrize("kind,h,w", [("texture", 375, 1242), ("texture", 1080, 1920), ("texture", 2160, 3840),
                                      ("noise", 200, 300),
                                      ("blocks", 256, 256), ("flat", 64, 80), ("noise", 24, 20),
                                      ("noise", 17, 33), ("smooth", 40, 56)])
def test_equals_live_cv2(cuda_dev, kind, h, w):
    """Against cv2 itself on the box: noise has ~10 key points per 100 pixels of the doubled image
    (the capacity path), the block image exactly repeated structure, a flat image no key point."""
    pytest.importorskip("cv2")
    rng = np.random.default_rng(h * w)
    if kind == "noise":
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    elif kind == "blocks":
        img = np.kron(rng.integers(0, 2, (h // 32, w // 32)) * 120 + 60, np.ones((32, 32))).astype(np.uint8)
        img[:, w // 2:] = img[:, :w // 2]
    elif kind == "flat":
        img = np.full((h, w), 77, np.uint8)
    elif kind == "smooth":       # images smaller than the widest blur's halo: the general reflection
        import cv2
        img = cv2.resize(rng.integers(0, 256, (5, 7), dtype=np.uint8), (w, h), interpolation=cv2.INTER_CUBIC)
    else:
        img = synth.textured_image(h, w, h + w)
    ref, rdesc = cv_ref(img)
    res, status, _ = run(SiftDetector(), img[None], cap=max(64, len(ref) + len(ref) // 8 + 64))
    assert status[0] == 0
    if len(ref) == 0:
        assert len(res[0][0]) == 0
        return
    check(res[0][0], res[0][1], ref, rdesc, record_parity, "%s_%dx%d" % (kind, h, w))


def test_batch_equals_single_frames_and_bgr(cuda_dev):
    imgs = np.stack([synth.textured_image(120, 168, 30 + i) for i in range(5)])
    det = SiftDetector()
    batch, status, _ = run(det, imgs)
    assert not status.any()
    for f in range(5):
        single, _, _ = run(det, imgs[f][None])
        assert np.array_equal(batch[f][0], single[0][0]) and np.array_equal(batch[f][1], single[0][1])
    # the same call twice: bit-identical (ordered sums, no atomics on floats)
    again, _, _ = run(det, imgs)
    assert all(np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) for a, b in zip(batch, again))
    # BGR input = cv2's gray conversion first
    g = imgs[0]
    bgr = np.stack([g, np.roll(g, 3, 0), np.roll(g, 5, 1)], -1)
    res, _, _ = run(det, bgr[None])
    ref, rdesc = cv_ref(bgr)
    check(res[0][0], res[0][1], ref, rdesc)


def test_capacity_overflow_is_reported_and_detect_and_compute_retries(cuda_dev):
    pytest.importorskip("cv2")
    img = synth.textured_image(240, 320, 8)
    ref, rdesc = cv_ref(img)
    assert len(ref) > 100
    _, status, _ = run(SiftDetector(), img[None], cap=50)
    assert status[0] == 1
    kps, desc = SiftDetector(cap=40).detectAndCompute(img)
    mine = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps], np.float64)
    assert desc.dtype == np.float32 and desc.shape == (len(kps), 128)
    check(mine, desc, ref, rdesc)
    with pytest.raises(NotImplementedError):
        SiftDetector(nOctaveLayers=4)


def test_images_to_matches_through_the_reference_surface(cuda_dev):
    """The reference's live configuration (src/visual_slam.py:17-19): FrameGenerator with the SIFT
    detector -> ImagePair.match_features with BFMatcher(NORM_L2, crossCheck=True): the same matches
    as cv2.SIFT_create + cv2.BFMatcher on the same images."""
    cv2 = pytest.importorskip("cv2")
    from visual_odometry_b200.image_pair import ImagePair
    from visual_odometry_b200.matcher import B200Matcher, NORM_L2
    big = synth.textured_image(700, 900, 70)
    img1, img2 = np.ascontiguousarray(big[60:540, 80:720]), np.ascontiguousarray(big[71:551, 97:737])
    K = np.array([[700.0, 0, 320], [0, 700.0, 240], [0, 0, 1]])
    gen = FrameGenerator(SiftDetector())
    f1, f2 = gen.make_frame(img1), gen.make_frame(img2)
    pair = ImagePair(f1, f2, B200Matcher(NORM_L2, crossCheck=True), K)
    pair.match_features()
    mine = {(tuple(np.round(m.keypoint1, 2)), tuple(np.round(m.keypoint2, 2))) for m in pair.filtered_matches}
    sift = cv2.SIFT_create()
    k1, d1 = sift.detectAndCompute(img1, None)
    k2, d2 = sift.detectAndCompute(img2, None)
    ref = {(tuple(np.round(k1[m.queryIdx].pt, 2)), tuple(np.round(k2[m.trainIdx].pt, 2)))
           for m in cv2.BFMatcher(cv2.NORM_L2, crossCheck=True).match(d1, d2)}
    assert len(ref) > 200
    assert len(mine & ref) >= 0.99 * len(ref) and abs(len(mine) - len(ref)) <= 0.01 * len(ref) + 1
    shift = np.array([np.subtract(a, b) for a, b in mine])
    assert (np.abs(shift - [17, 11]).max(1) < 1.0).mean() > 0.5


def test_image_sequence_to_poses_on_the_device_with_sift(cuda_dev):
    """Images -> SIFT -> L2 cross-check matching -> E-RANSAC -> recoverPose without leaving the
    device (PairPipeline.run_images on a "sift" pipeline): poses follow the rendered ground truth."""
    pytest.importorskip("cv2")
    from test_gpu_orb import _render_two_plane_sequence
    from visual_odometry_b200.pipeline import PairPipeline
    images, K, cams = _render_two_plane_sequence(4)
    pipe = PairPipeline(4096, 3, 4, descriptor="sift", K=K, device="cuda:0")
    b = pipe.run_images(images, SiftDetector())
    torch.cuda.synchronize()
    assert b == 3
    Rg, tg = pipe.R[:b].cpu().numpy().reshape(b, 3, 3), pipe.t[:b].cpu().numpy().reshape(b, 3)
    for k in range(b):
        (R1, C1), (R2, C2) = cams[k], cams[k + 1]
        R_true = R2 @ R1.T
        t_true = R2 @ (C1 - C2)
        t_true /= np.linalg.norm(t_true)
        ang = np.degrees(np.arccos(np.clip((np.trace(Rg[k].T @ R_true) - 1) / 2, -1, 1)))
        tdir = np.degrees(np.arccos(np.clip(tg[k] @ t_true, -1, 1)))
        assert int(pipe.n_inl[k].item()) > 100, (k, int(pipe.n_inl[k].item()))
        assert ang < 2.0 and tdir < 15.0, (k, ang, tdir)


def test_slam_loop_from_images_with_sift(cuda_dev):
    """DeviceSlam.run_images on the reference's live configuration: rendered frames -> SIFT
    extraction -> L2 matching -> E-RANSAC / pose / triangulation -> map stage (PnP, map growth,
    bundle adjustment, pruning) without descriptors or key points touching the host."""
    pytest.importorskip("cv2")
    from test_gpu_orb import _render_two_plane_sequence
    from visual_odometry_b200.slam import DeviceSlam
    images, K, cams = _render_two_plane_sequence(6)
    slam = DeviceSlam(K, 4096, 6, descriptor="sift", ba_iterations=10)
    log = slam.run_images(images, SiftDetector())
    assert len(log) == 5
    assert log[0]["n_points"] > 300 and log[0]["n_obs"] >= 2 * log[0]["n_points"] - 5
    assert sum(bool(e["pnp_ok"]) for e in log[1:]) >= 3, log
    assert log[-1]["n_points"] > log[0]["n_points"] // 2
    assert len(log[-1]["cameras"]) >= 4


def test_c_abi_argument_errors(cuda_dev):
    """vo_sift_detect_batch / vo_sift_workspace_bytes reject what they cannot run (error codes, no
    exception across the ABI, nothing launched)."""
    from visual_odometry_b200 import _lib
    lib = _lib.load()
    assert lib.vo_sift_workspace_bytes(0, 64, 64, 100) == 0
    assert lib.vo_sift_workspace_bytes(1, 1, 64, 100) == 0
    assert lib.vo_sift_octaves(64, 48, None, None) == 5
    img = torch.zeros((1, 48, 64), dtype=torch.uint8, device=cuda_dev)
    cap = 64
    kp = torch.zeros((1, cap, 6), dtype=torch.float32, device=cuda_dev)
    desc = torch.zeros((1, cap, 128), dtype=torch.uint8, device=cuda_dev)
    n = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    st = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    need = lib.vo_sift_workspace_bytes(1, 64, 48, cap)
    ws = torch.empty(need, dtype=torch.uint8, device=cuda_dev)
    args = lambda wsb, c=cap, w=64, nf=0: (_lib.ptr(img), 1, w, 48, c, nf, _lib.ptr(kp), _lib.ptr(desc), _lib.ptr(n),
                                     _lib.ptr(st), None, _lib.ptr(ws), wsb, _lib.cur_stream())
    assert lib.vo_sift_detect_batch(*args(need)) == 0
    assert lib.vo_sift_detect_batch(*args(need - 1)) == _lib.VO_ERR_WORKSPACE
    assert b"workspace" in lib.vo_last_error_string()
    assert lib.vo_sift_detect_batch(*args(need, c=0)) == _lib.VO_ERR_INVALID_ARG
    assert lib.vo_sift_detect_batch(*args(need, w=1)) == _lib.VO_ERR_INVALID_ARG
    assert lib.vo_sift_detect_batch(*args(need, nf=-1)) == _lib.VO_ERR_INVALID_ARG
    assert lib.vo_sift_detect_batch(None, 1, 64, 48, cap, 0, _lib.ptr(kp), _lib.ptr(desc), _lib.ptr(n),
                                    _lib.ptr(st), None, _lib.ptr(ws), need, _lib.cur_stream()) == _lib.VO_ERR_INVALID_ARG
    torch.cuda.synchronize()
    assert int(n.item()) == 0 and int(st.item()) == 0


@pytest.mark.parametrize("nfeatures", [1, 50, 300, 100000])
def test_nfeatures_equals_cv2_retain_best_as_a_set(cuda_dev, nfeatures):
    """cv2.SIFT_create(nfeatures): KeyPointsFilter::retainBest keeps every key point whose response
    is >= the n-th largest one (ties kept) and leaves them in std::nth_element's order; here they
    stay in the sorted order: equal as sets, descriptors row by row."""
    cv2 = pytest.importorskip("cv2")
    img = synth.textured_image(300, 400, 77)
    kps, rdesc = cv2.SIFT_create(nfeatures=nfeatures).detectAndCompute(img, None)
    ref = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps],
                   np.float64).reshape(-1, 6)
    res, status, _ = run(SiftDetector(nfeatures=nfeatures), img[None])
    assert status[0] == 0
    kp, desc = res[0]
    assert len(kp) == len(ref) and (nfeatures > 1000 or len(kp) <= nfeatures + 3)
    # pair as sets: sort both by (x, y, size, angle)
    oa = np.lexsort((kp[:, 3], kp[:, 2], kp[:, 1], kp[:, 0]))
    ob = np.lexsort((ref[:, 3], ref[:, 2], ref[:, 1], ref[:, 0]))
    d = np.abs(kp[oa][:, :5] - ref[ob][:, :5])
    assert d[:, :2].max() <= 2e-3 and d[:, 2].max() <= 1e-3 and d[:, 4].max() <= 1e-6
    assert np.array_equal(kp[oa][:, 5], ref[ob][:, 5])
    dd = np.abs(desc[oa].astype(np.int32) - rdesc[ob].astype(np.int32))
    assert dd.max() <= 1 and (dd > 0).mean() <= 0.01
    # and the survivors are in cv2's sorted order among themselves
    assert np.all(np.diff(kp[:, 0]) >= 0)
