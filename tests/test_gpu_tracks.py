"""GPU: N1 feature-track bookkeeping (csrc/tracks.cu) vs the dict-based restatement of the
reference's VisualSlam logic on a synthetic sequence: identical roots, identical MatchWithMap
lists (order, feature ids, coordinates bit-exact), and PnP fed straight from the device arrays."""
import numpy as np
import pytest
import torch

from oracle.tracks_restate import TracksOracle
from visual_odometry_b200.pipeline import PairPipeline
from visual_odometry_b200.synth import make_sequence
from visual_odometry_b200.tracks import FeatureTracks

pytestmark = pytest.mark.gpu


def test_tracks_equal_reference_bookkeeping(cuda_dev):
    F, N = 8, 600
    seq = make_sequence(F, N, 77)
    pipe = PairPipeline(N, F - 1, K=seq.K)
    pipe.run_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
    tr, orc = FeatureTracks(F, N), TracksOracle()
    q, t, mask = pipe.q.cpu().numpy(), pipe.t_idx.cpu().numpy(), pipe.mask.cpu().numpy()
    nm = pipe.n_match.cpu().numpy()

    def pair_matches(b):
        return [((b, int(q[b, i])), (b + 1, int(t[b, i])), tuple(seq.kp[b + 1][t[b, i]].astype(np.float64)))
                for i in range(nm[b]) if mask[b, i]]

    # map initialised from pair 0 (visual_slam.py:43-87): one point per inlier match, keyed by
    # featureid1; coordinates are arbitrary here (bookkeeping is what is under test)
    m0 = pair_matches(0)
    rng = np.random.default_rng(0)
    map_xyz = rng.uniform(-10, 10, (len(m0), 3))
    orc.update_feature_mapper(m0)
    orc.set_map_points([(f1, map_xyz[i]) for i, (f1, _, _) in enumerate(m0)])
    tr.update(pipe.q[0], pipe.t_idx[0], pipe.mask[0], pipe.n_match[0:1], 0, 1)
    tr.set_map_points([tr.feature_id(*f1) for f1, _, _ in m0])
    map_dev = torch.from_numpy(map_xyz).cuda()
    total = 0
    for b in range(1, F - 1):
        mb = pair_matches(b)
        orc.update_feature_mapper(mb)
        tr.update(pipe.q[b], pipe.t_idx[b], pipe.mask[b], pipe.n_match[b:b + 1], b, b + 1)
        # roots of every feature of the newest frame
        ids = [tr.feature_id(b + 1, i) for i in range(N)]
        roots = tr.root_of(ids).cpu().numpy()
        want = [tr.feature_id(*orc.track_feature_back_in_time((b + 1, i))) for i in range(N)]
        assert roots.tolist() == want
        img, xyz, feat2, prow = tr.matches_with_map(pipe.q[b], pipe.t_idx[b], pipe.mask[b],
                                                    pipe.n_match[b:b + 1], b, b + 1, pipe.kp, map_dev)
        ref = orc.matches_with_map(mb)
        assert len(ref) == img.shape[0]
        assert feat2.cpu().tolist() == [tr.feature_id(*r[0]) for r in ref]
        assert prow.cpu().tolist() == [r[4] for r in ref]
        assert np.array_equal(img.cpu().numpy(), np.array([r[2] for r in ref]).reshape(-1, 2))
        assert np.array_equal(xyz.cpu().numpy(), np.array([r[3] for r in ref]).reshape(-1, 3))
        total += len(ref)
    assert total > 200          # tracks really do survive several frames


def test_tracks_feed_pnp_on_device(cuda_dev):
    """map tracking step (C5 flavour): triangulated points of pair 0 become the map; frame 2 is
    localised by PnP-RANSAC from the device-side MatchWithMap arrays; pose = chained ground truth."""
    import cv2
    from visual_odometry_b200 import pnp as PNP
    F, N = 3, 2000
    seq = make_sequence(F, N, 20261104 + 4)
    pipe = PairPipeline(N, F - 1, K=seq.K)
    pipe.run_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
    tr = FeatureTracks(F, N)
    n0 = int(pipe.n_match[0])
    inl = pipe.mask[0, :n0].bool()
    # reference flavour: points in frame-2 coordinates of pair 0 (image_pair.py:316-339), compacted
    X = pipe.X[0, :int(pipe.n_x[0]), :3].contiguous()
    fid = (0 * N + pipe.q[0, :n0][inl]).to(torch.int32)
    tr.update(pipe.q[0], pipe.t_idx[0], pipe.mask[0], pipe.n_match[0:1], 0, 1)
    tr.set_map_points(fid)
    tr.update(pipe.q[1], pipe.t_idx[1], pipe.mask[1], pipe.n_match[1:2], 1, 2)
    img, xyz, _, _ = tr.matches_with_map(pipe.q[1], pipe.t_idx[1], pipe.mask[1], pipe.n_match[1:2],
                                         1, 2, pipe.kp, X)
    assert img.shape[0] > 300
    ok, rvec, tvec, inliers = PNP.solve_pnp_ransac(xyz.cpu().numpy(), img.cpu().numpy(), seq.K, np.zeros(4))
    assert ok and len(inliers) > 0.8 * img.shape[0]
    okr, rr, tr_, inr = cv2.solvePnPRansac(xyz.cpu().numpy(), img.cpu().numpy(), seq.K, np.zeros(4))
    ang = lambda A, B: np.arccos(np.clip((np.trace(A.T @ B) - 1) / 2, -1, 1))
    assert ang(cv2.Rodrigues(rvec)[0], cv2.Rodrigues(rr)[0]) < 1e-3


def test_tracks_equal_reference_golden(cuda_dev):
    """tests/golden/tracks.npz (made by the reference's own VisualSlam): the device bookkeeping
    returns the same roots and the same solvePnPRansac argument arrays, bit for bit."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tracks.npz"))
    F, N = int(g["F"]), int(g["N"])
    cap = N + 8                                   # some map points carry idx >= N
    tr = FeatureTracks(F, cap)
    kp = np.zeros((F, cap, 2), np.float32)
    for b in range(F - 1):
        kp[b + 1, g[f"f2_{b}"][:, 1]] = g[f"kp2_{b}"]
    kp_d = torch.from_numpy(kp).cuda()
    n_calls = 0
    for b in range(F - 1):
        f1, f2 = g[f"f1_{b}"], g[f"f2_{b}"]
        assert (f1[:, 0] == b).all() and (f2[:, 0] == b + 1).all()
        q = torch.from_numpy(f1[:, 1].astype(np.int32)).cuda()
        t = torch.from_numpy(f2[:, 1].astype(np.int32)).cuda()
        n = torch.tensor([len(f1)], dtype=torch.int32, device="cuda")
        tr.update(q, t, None, n, b, b + 1)
        n_map = int(g[f"n_map_{b}"])
        ids = g["map_ids"][:n_map]
        tr.set_map_points(ids[:, 0] * cap + ids[:, 1])
        roots = tr.root_of([tr.feature_id(b + 1, i) for i in range(N)]).cpu().numpy()
        want = g[f"roots_{b}"]
        assert np.array_equal(roots, want[:, 0] * cap + want[:, 1])
        if int(g[f"called_{b}"]):
            xyz_d = torch.from_numpy(g["map_xyz"][:n_map]).cuda()
            img, xyz, _, _ = tr.matches_with_map(q, t, None, n, b, b + 1, kp_d, xyz_d)
            assert np.array_equal(img.cpu().numpy(), g[f"img_{b}"])
            assert np.array_equal(xyz.cpu().numpy(), g[f"obj_{b}"])
            n_calls += 1
    assert n_calls == F - 2


def test_duplicate_train_features_last_match_wins_like_the_reference(cuda_dev):
    """knn + ratio matching can pair several queries with ONE train feature.  The reference
    inserts feature_mapper[featureid2] = featureid1 in match order, so the last match wins; the
    device update elects the highest match index per train feature (deterministic, twice the
    same answer)."""
    N = 64
    tr = FeatureTracks(3, N)
    q = torch.tensor([3, 5, 7, 9, 11, 13], dtype=torch.int32, device=cuda_dev)
    t = torch.tensor([2, 4, 2, 6, 4, 2], dtype=torch.int32, device=cuda_dev)     # 2 three times, 4 twice
    n = torch.tensor([6], dtype=torch.int32, device=cuda_dev)
    orc = TracksOracle()
    orc.update_feature_mapper([((0, int(a)), (1, int(b)), (0.0, 0.0)) for a, b in zip(q.tolist(), t.tolist())])
    for _ in range(2):
        tr.update(q, t, None, n, 0, 1)
        ids = [tr.feature_id(1, i) for i in (2, 4, 6, 8)]
        roots = tr.root_of(ids).cpu().tolist()
        want = [tr.feature_id(*orc.track_feature_back_in_time((1, i))) for i in (2, 4, 6, 8)]
        assert roots == want == [tr.feature_id(0, 13), tr.feature_id(0, 11), tr.feature_id(0, 9),
                                 tr.feature_id(1, 8)]
