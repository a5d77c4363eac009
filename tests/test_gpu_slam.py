"""GPU: the whole SLAM frame loop on the device path (DeviceSlam: batched hot path + device map +
PnP + bundle adjustment + pruning + culling) against tests/golden/slam_run.npz, which the
reference's OWN VisualSlam.match_current_and_previous_frame produced on the CPU
(tests/golden/make_golden_slam_run.py: cv2 for the hot path, oracle/ba_restate.py in place of
g2o), and a functional check against the ground-truth trajectory."""
import os

import numpy as np
import pytest

from visual_odometry_b200.slam import DeviceSlam
from visual_odometry_b200.synth import make_sequence

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_device_slam_equals_reference_frame_loop(cuda_dev):
    """Frame 1 (initialize_map + first bundle adjustment) is compared tightly.  From frame 2 on
    the loop is chaotic — PnP-RANSAC picks among near-equal hypotheses and the map's scale is a
    free gauge of the adjustment, so last-bit input differences grow to percents — and only what
    is not amplified is compared: the hot-path results per pair (exact), the camera list, the
    fixed flags and the culling schedule."""
    z = np.load(os.path.join(GOLDEN, "slam_run.npz"))
    g = {k: z[k] for k in z.files}
    F, N = int(g["F"]), int(g["N"])
    seq = make_sequence(F, N, int(g["seed"]), descriptor="sift", K=g["K"])
    slam = DeviceSlam(g["K"], N, F, descriptor="sift", max_cameras=int(g["max_cams"]))
    p = slam.pipe
    pairs = np.stack([np.arange(F - 1), np.arange(1, F)], 1).astype(np.int32)
    p.run_host(seq.desc, seq.kp, seq.nkp, pairs)
    for b in range(F - 1):
        info = slam.step(b)
        tag = f"f{b + 1}"
        m = slam.map
        assert int(p.n_match[b]) == int(g[tag + "_n_matches"])
        assert int(p.mask[b, :int(p.n_match[b])].sum()) == int(g[tag + "_n_3d"])
        n_cam = len(m.cam_frame)
        if b <= 4:      # later the culled map holds a few dozen points and localisation is a coin toss
            assert info["cameras"] == g[tag + "_cam_frame"].tolist()
            assert m.fixed[:n_cam].cpu().numpy().astype(bool).tolist() == g[tag + "_cam_fixed"].tolist()
        if b == 0:
            n_p, n_o = m.counts()
            assert (n_p, n_o) == (len(g[tag + "_pt_xyz"]), len(g[tag + "_obs_pt"]))
            fid = g[tag + "_pt_fid"]
            assert np.array_equal(m.point_fid[:n_p].cpu().numpy(), fid[:, 0] * N + fid[:, 1])
            assert np.array_equal(m.obs_pt[:n_o].cpu().numpy(), g[tag + "_obs_pt"])
            assert np.array_equal(m.obs_cam[:n_o].cpu().numpy(), g[tag + "_obs_cam"])
            assert np.array_equal(m.obs_xy[:n_o].cpu().numpy(), g[tag + "_obs_xy"])
            ba, ref = info["ba_init"], g[tag + "_ba_log"][0]
            assert abs(ba["chi2_initial"] - ref[3]) <= 1e-4 * ref[3]      # same problem going in
            assert abs(ba["chi2"][-1] - ref[4]) <= 1e-3 * ref[4]          # same cost coming out
            poses = m.poses[:n_cam].cpu().numpy()
            assert np.abs(poses - g[tag + "_cam_pose"]).max() < 1e-4
            X, Xr = m.points[:n_p].cpu().numpy(), g[tag + "_pt_xyz"]
            assert np.median(np.abs(X - Xr).max(1) / np.abs(Xr).max(1)) < 1e-4
        elif b <= 3:
            # the reference localised these frames too (before culling empties the map)
            assert info["pnp_ok"] and int(g[tag + "_pnp_log"][0][1]) == 1
            assert info["pnp_points"] == int(g[tag + "_pnp_log"][0][0]) or b > 1


def test_device_slam_tracks_the_true_trajectory(cuda_dev):
    """functional check on an ORB sequence: after Procrustes alignment (the map's scale and frame
    are free) the camera centres follow the ground truth."""
    F, N = 10, 2000
    seq = make_sequence(F, N, 5, descriptor="orb")
    slam = DeviceSlam(seq.K, N, F, descriptor="orb")
    log = slam.run(seq.desc, seq.kp, seq.nkp)
    m = slam.map
    assert log[-1]["cameras"] == list(range(F))
    assert all(l["pnp_ok"] and l["pnp_inliers"] > 0.9 * l["pnp_points"] for l in log[1:])
    assert log[-1]["n_points"] > 2 * log[0]["n_points"]            # the map keeps growing
    poses = m.poses[:F].cpu().numpy().reshape(F, 4, 4)
    C = np.array([-P[:3, :3].T @ P[:3, 3] for P in poses])
    Cg = np.array([-seq.R_w2c[f].T @ seq.t_w2c[f] for f in range(F)])
    A, B = C - C.mean(0), Cg - Cg.mean(0)
    U, S, Vt = np.linalg.svd(A.T @ B)
    D = np.diag([1, 1, np.sign(np.linalg.det(U @ Vt))])
    Rr = (U @ D @ Vt).T
    s = (S * np.diag(D)).sum() / (A ** 2).sum()
    err = np.linalg.norm(s * (Rr @ A.T).T - B, axis=1)
    assert err.max() < 0.03 * np.linalg.norm(B, axis=1).max()


def test_streaming_equals_batch(cuda_dev):
    """add_frame (one pair at a time, as the reference consumes a video) gives the same map as
    the batched pass over the whole sequence, bit for bit."""
    F, N = 6, 1200
    seq = make_sequence(F, N, 5, descriptor="orb")
    a = DeviceSlam(seq.K, N, F, descriptor="orb")
    a.run(seq.desc, seq.kp, seq.nkp)
    b = DeviceSlam(seq.K, N, F, descriptor="orb")
    for f in range(F):
        b.add_frame(seq.desc[f, :seq.nkp[f]], seq.kp[f, :seq.nkp[f]])
    (pa, oa), (pb, ob) = a.map.counts(), b.map.counts()
    assert (pa, oa) == (pb, ob) and a.map.cam_frame == b.map.cam_frame
    assert np.array_equal(a.map.poses[:F].cpu().numpy(), b.map.poses[:F].cpu().numpy())
    assert np.array_equal(a.map.points[:pa].cpu().numpy(), b.map.points[:pb].cpu().numpy())
    assert np.array_equal(a.map.obs_pt[:oa].cpu().numpy(), b.map.obs_pt[:ob].cpu().numpy())
