"""GPU: recoverPose, DLT triangulation, chaining, the ImagePair drop-in and the batched pipeline."""
import contextlib
import io

import numpy as np
import pytest
import torch

import cv2
from oracle import cv_geometry as CG
from oracle import reference_path as RP
from visual_odometry_b200 import geometry as G
from visual_odometry_b200 import image_pair as IP
from visual_odometry_b200.frame import FrameGenerator, SyntheticDetector
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING, NORM_L2
from visual_odometry_b200.pipeline import PairPipeline
from visual_odometry_b200.synth import make_sequence
from conftest import load_golden, GOLDEN_CASES, record_parity

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-4    # rad   (north_star)
T_TOL = 1e-4      # unit translation
X_TOL = 1e-4      # relative, triangulated points


def rot_angle(Ra, Rb):
    return float(np.arccos(np.clip((np.trace(Ra.T @ Rb) - 1) / 2, -1, 1)))


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_recover_pose_and_triangulation_equal_golden(cuda_dev, name):
    g = load_golden(name)
    inl = np.isin(g["q"], g["inlier_q"])
    p1 = g["kp"][0][g["q"]][inl].astype(np.float64)
    p2 = g["kp"][1][g["t"]][inl].astype(np.float64)
    n, R, t, mask = IP.recover_pose(g["E"], p1, p2, g["K"])
    ncv, Rcv, tcv, mcv = cv2.recoverPose(g["E"], p1, p2, g["K"])
    assert t.shape == (3, 1) and mask.shape == (len(p1), 1) and set(np.unique(mask)) <= {0, 255}
    assert rot_angle(R, g["R"]) < ROT_TOL and np.abs(t - g["tvec"]).max() < T_TOL
    assert n == ncv and np.array_equal(mask, mcv)
    X = IP.triangulate_points(g["projection_matrix"], g["null_projection_matrix"], p1.T, p2.T)
    assert X.shape == g["X"].shape and abs(np.linalg.norm(X[:, 0]) - 1) < 1e-12  # unit norm
    X = X / X[3]
    rel = np.abs(X - g["X"]).max(0) / np.abs(g["X"]).max(0)
    assert rel.max() < X_TOL


def test_triangulate_float32_points_give_float32(cuda_dev):
    P1 = np.hstack([cv2.Rodrigues(np.array([0.01, 0.02, -0.01]))[0], [[0.4], [0.1], [-0.3]]])
    x1 = np.random.default_rng(0).normal(0, 0.3, (2, 64)).astype(np.float32)
    X = IP.triangulate_points(P1, np.eye(3, 4), x1, x1 + np.float32(0.01))
    Xcv = cv2.triangulatePoints(P1, np.eye(3, 4), x1, x1 + np.float32(0.01))
    assert X.dtype == Xcv.dtype == np.float32
    s = np.sign((X * Xcv).sum(0))
    assert np.abs(X * s - Xcv).max() < 1e-5


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_image_pair_dropin_attribute_for_attribute(cuda_dev, name):
    """Our ImagePair driven exactly like src/visual_slam.py:294-298 / :165-172 drives the
    reference's, compared with the real reference's outputs (golden)."""
    g = load_golden(name)
    seq = type("S", (), {})()
    seq.desc, seq.kp, seq.nkp = g["desc"], g["kp"], np.array([len(g["kp"][0])] * 2)
    fg = FrameGenerator(SyntheticDetector(seq))
    f1, f2 = fg.make_frame(np.zeros((4, 4, 3), np.uint8)), fg.make_frame(np.zeros((4, 4, 3), np.uint8))
    assert f1.id == 0 and f2.id == 1 and f1.features[3].feature_id == (0, 3)
    norm = NORM_HAMMING if g["descriptor"] == "orb" else NORM_L2
    ip = IP.ImagePair(f1, f2, B200Matcher(norm, crossCheck=True), g["K"])
    with contextlib.redirect_stdout(io.StringIO()):
        ip.match_features()
        ess = ip.determine_essential_matrix(ip.filtered_matches)
        ip.estimate_camera_movement(ess)
        ip.reconstruct_3d_points(ess)
    assert [m.featureid1 for m in ip.raw_matches] == [(0, int(i)) for i in g["q"]]
    assert [m.featureid2 for m in ip.raw_matches] == [(1, int(i)) for i in g["t"]]
    assert [m.distance for m in ip.filtered_matches] == g["dist"].tolist()
    assert [m.featureid1[1] for m in ess] == g["inlier_q"].tolist()
    assert ip.t.shape == (3, 1) and ip.relative_pose.shape == (4, 4)
    assert rot_angle(ip.R, g["R"]) < ROT_TOL and np.abs(ip.t - g["tvec"]).max() < T_TOL
    assert np.abs(ip.projection_matrix - g["projection_matrix"]).max() < 1e-3 * ROT_TOL * 1e4
    rel = np.abs(ip.points3d_reconstr - g["X"]).max(0) / np.abs(g["X"]).max(0)
    assert rel.max() < X_TOL
    assert len(ip.matches_with_3d_information) == len(ess)
    assert ip.matches_with_3d_information[0].point == tuple(ip.points3d_reconstr[:3, 0])


def test_pipeline_sequence_vs_reference_path_and_ground_truth(cuda_dev):
    F, N = 17, 1500
    seq = make_sequence(F, N, 8)
    pipe = PairPipeline(N, F - 1, K=seq.K)
    out = pipe.run_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
    Rs, ts = [], []
    exact, off_pairs = 0, []
    for b in range(F - 1):
        ref = RP.pair_cv2(seq.desc[b], seq.kp[b], seq.desc[b + 1], seq.kp[b + 1], seq.K)
        assert out["n_match"][b] == len(ref["q"])
        n = int(out["n_match"][b])
        if not np.array_equal(pipe.mask[b, :n].cpu().numpy(), ref["mask"].ravel()):
            # a different consensus set (an inlier-count tie between two hypotheses, or a
            # correspondence on the threshold): how far the pose then is from cv2's
            off_pairs.append({"pair": b, "inliers": [int(out["n_inl"][b]), int(ref["mask"].sum())],
                              "rot_rad_vs_cv2": float(rot_angle(out["R"][b], ref["R"])),
                              "t_abs_vs_cv2": float(np.abs(out["t"][b] - ref["tvec"].ravel()).max())})
        if out["n_inl"][b] == int(ref["mask"].sum()):
            exact += 1
            assert rot_angle(out["R"][b], ref["R"]) < ROT_TOL
            assert np.abs(out["t"][b] - ref["tvec"].ravel()).max() < T_TOL
            n = int(pipe.n_x[b])
            X = pipe.X[b, :n].cpu().numpy().T
            rel = np.abs(X - ref["X"]).max(0) / np.abs(ref["X"]).max(0)
            assert np.median(rel) < X_TOL / 4 and (rel < X_TOL).mean() > 0.99
        Rg, tg = seq.relative_pose(b, b + 1)
        assert rot_angle(out["R"][b], Rg) < 0.02          # and close to ground truth
        Rs.append(ref["R"]); ts.append(ref["tvec"])
    record_parity("pipeline_sequence_vs_reference_path", pairs=F - 1, inlier_counts_equal=exact,
                  pairs_with_different_mask=off_pairs)
    # measured (profiles/parity_r2.json): 15 of 16 pairs identical; on pair 5 ONE correspondence on
    # the threshold flips (911 vs 912 inliers: the five-point roots differ from cv2's Durand-Kerner
    # roots at 1e-8) and the pose stays within 2e-6 rad / 1.3e-4 of cv2's
    assert exact >= F - 2 and len(off_pairs) <= 1
    for o in off_pairs:
        assert abs(o["inliers"][0] - o["inliers"][1]) <= 2
        assert o["rot_rad_vs_cv2"] < 1e-4 and o["t_abs_vs_cv2"] < 5e-4
    if exact == F - 1:
        assert np.abs(out["poses"] - RP.chain_poses(Rs, ts)).max() < 1e-3


def test_chain_is_product_of_relative_poses(cuda_dev):
    rng = np.random.default_rng(0)
    R = np.stack([cv2.Rodrigues(rng.normal(0, 0.1, 3))[0] for _ in range(50)])
    t = rng.normal(0, 1, (50, 3))
    poses = G.chain_poses(torch.from_numpy(R).cuda(), torch.from_numpy(t).cuda()).cpu().numpy()
    assert np.abs(poses - RP.chain_poses(R, t)).max() < 1e-12


def test_full_size_c2_property(cuda_dev):
    """configs[1] size (ORB 2000 kp; 120 consecutive pairs here): every pair solves, inlier
    ratios and poses are sane against ground truth, the chain stays orthonormal."""
    F, N = 121, 2000
    seq = make_sequence(F, N, 20261105)
    pipe = PairPipeline(N, F - 1, K=seq.K)
    out = pipe.run_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
    assert np.all(out["status"] == 0)
    assert np.all(out["n_match"] > 0.6 * N) and np.all(out["n_inl"] > 0.5 * out["n_match"])
    errs = [rot_angle(out["R"][b], seq.relative_pose(b, b + 1)[0]) for b in range(F - 1)]
    assert np.median(errs) < 5e-3
    P = out["poses"]
    assert np.abs(P[:, :3, :3] @ P[:, :3, :3].transpose(0, 2, 1) - np.eye(3)).max() < 1e-9


def test_full_size_c4_batch_vs_reference_path(cuda_dev):
    """config C4 size (ORB 4000 kp; 24 pairs of the 4096 here): the batched pipeline equals the
    reference's cv2 chain pair by pair (matches and masks exact, pose within 1e-4) on the pairs
    checked, and every pair of the batch solves."""
    F, N = 25, 4000
    seq = make_sequence(F, N, 20261104 + 3)
    pipe = PairPipeline(N, F - 1, K=seq.K)
    out = pipe.run_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
    assert np.all(out["status"] == 0) and np.all(out["n_match"] > 0.6 * N)
    exact, off_pairs = 0, []
    for b in (0, 11, 23):
        ref = RP.pair_cv2(seq.desc[b], seq.kp[b], seq.desc[b + 1], seq.kp[b + 1], seq.K,
                          cv2.NORM_HAMMING)
        n = int(out["n_match"][b])
        assert n == len(ref["q"])
        assert np.array_equal(pipe.q[b, :n].cpu().numpy(), ref["q"])
        assert np.array_equal(pipe.t_idx[b, :n].cpu().numpy(), ref["t"])
        assert np.array_equal(pipe.dist[b, :n].cpu().numpy(), ref["dist"])
        if np.array_equal(pipe.mask[b, :n].cpu().numpy(), ref["mask"].ravel()):
            exact += 1
            assert rot_angle(out["R"][b], ref["R"]) < 1e-4
            assert np.abs(out["t"][b] - ref["tvec"].ravel()).max() < 1e-4
        else:
            off_pairs.append({"pair": b, "mask_bits": int((pipe.mask[b, :n].cpu().numpy() != ref["mask"].ravel()).sum()),
                              "rot_rad_vs_cv2": float(rot_angle(out["R"][b], ref["R"])),
                              "t_abs_vs_cv2": float(np.abs(out["t"][b] - ref["tvec"].ravel()).max())})
    record_parity("c4_batch_vs_reference_path", pairs_checked=3, masks_identical=exact,
                  pairs_with_different_mask=off_pairs)
    assert exact == 3                         # measured: masks identical on 3 / 3


@pytest.mark.parametrize("use_fill", [True, False])
def test_run_stream_equals_the_one_call_path(cuda_dev, use_fill):
    """PairPipeline.run_stream (batch 0 uploaded in two pieces with the chain started on the first,
    every later batch uploaded whole while its predecessor computes, two input slots) must give,
    batch by batch, exactly what run_host gives for the same frames."""
    F, N = 81, 640
    seqs = [make_sequence(F, N, 300 + s) for s in range(3)]
    pairs = seqs[0].consecutive_pairs()
    pipe = PairPipeline(N, F - 1, K=seqs[0].K)
    want = [pipe.run_host(s.desc, s.kp, s.nkp, pairs) for s in seqs]
    got = {}

    def fill(k, pin):
        s = seqs[k % 3]
        pin["desc"][:F].copy_(torch.from_numpy(s.desc))
        pin["kp"][:F].copy_(torch.from_numpy(s.kp.astype(np.float32)))
        pin["nkp"][:F].copy_(torch.from_numpy(np.asarray(s.nkp, np.int32)))
        pin["pairs"][:F - 1].copy_(torch.from_numpy(np.asarray(pairs, np.int32)))

    def consume(k, res):
        got[k] = {n: v.numpy().copy() for n, v in res.items()}

    pipe.stage_host(seqs[0].desc, seqs[0].kp, seqs[0].nkp, pairs)
    pipe.run_stream(F, F - 1, 5, fill=fill if use_fill else None, consume=consume)
    torch.cuda.synchronize()
    assert sorted(got) == [0, 1, 2, 3, 4]
    for k in range(5):
        w = want[k % 3 if use_fill else 0]
        assert (w["n_match"] > 0.5 * N).all()
        for name in ("n_match", "n_inl", "status", "R", "t", "poses"):
            assert np.array_equal(got[k][name], w[name]), (k, name)
