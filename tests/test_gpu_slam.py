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


def _load_golden_state(slam, g, tag, N, b):
    """put the reference's map after frame b (golden `tag`) into the device map and replay the
    feature_mapper links of pairs 0..b-1 from the hot-path results"""
    import torch
    m, p = slam.map, slam.pipe
    m.tracks.reset()
    for k in range(b):
        m.update_tracks(p.q[k], p.t_idx[k], p.mask[k], p.n_match[k:k + 1], k, k + 1)
    fid = g[tag + "_pt_fid"]
    n_p, n_o, n_c = len(fid), len(g[tag + "_obs_pt"]), len(g[tag + "_cam_frame"])
    dev = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dt).cuda()
    m.points[:n_p].copy_(dev(g[tag + "_pt_xyz"], torch.float64))
    m.point_fid[:n_p].copy_(dev(fid[:, 0] * N + fid[:, 1], torch.int32))
    m.obs_cam[:n_o].copy_(dev(g[tag + "_obs_cam"], torch.int32))
    m.obs_pt[:n_o].copy_(dev(g[tag + "_obs_pt"], torch.int32))
    m.obs_xy[:n_o].copy_(dev(g[tag + "_obs_xy"], torch.float64))
    m.poses[:n_c].copy_(dev(g[tag + "_cam_pose"], torch.float64))
    m.fixed[:n_c].copy_(dev(g[tag + "_cam_fixed"], torch.uint8))
    m.cam_frame = g[tag + "_cam_frame"].tolist()
    m.n_points.fill_(n_p); m.n_obs.fill_(n_o); m.status.zero_()
    m.tracks.set_map_points(m.point_fid[:n_p])


def test_each_frame_from_the_reference_state(cuda_dev):
    """Teacher-forced: every frame's map stage starts from the REFERENCE's map after the previous
    frame (golden), so rounding differences cannot accumulate.  Where PnP-RANSAC settles on cv2's
    consensus set (pose equal to 1e-6) the whole stage is compared tightly: adjustment problem going
    in (robust chi2 1e-8), cost coming out (1e-6), map sizes after pruning and culling (exact),
    points and poses (1e-6).  On this poor map (the reference's own two-view initialisation) many
    points sit near the 8 px threshold and cv2 / the kernel pick different, equally supported sets on
    some frames; there the adjusted cost and the map sizes must still agree to a few percent."""
    import cv2
    z = np.load(os.path.join(GOLDEN, "slam_run.npz"))
    g = {k: z[k] for k in z.files}
    F, N = int(g["F"]), int(g["N"])
    seq = make_sequence(F, N, int(g["seed"]), descriptor="sift", K=g["K"])
    slam = DeviceSlam(g["K"], N, F, descriptor="sift", max_cameras=int(g["max_cams"]))
    pairs = np.stack([np.arange(F - 1), np.arange(1, F)], 1).astype(np.int32)
    slam.pipe.run_host(seq.desc, seq.kp, seq.nkp, pairs)
    checked = 0
    for b in range(1, F - 1):
        prev, tag = f"f{b}", f"f{b + 1}"
        pnp_ref = g[tag + "_pnp_log"]
        if len(pnp_ref) == 0 or pnp_ref[0][1] != 1 or pnp_ref[0][0] < 20:
            continue                       # the reference did not localise this frame (tiny map)
        _load_golden_state(slam, g, prev, N, b)
        info = slam.step(b)
        assert info["pnp_points"] == int(pnp_ref[0][0])
        assert info["pnp_ok"]
        d_pose = np.abs(info["pnp_pose"] - pnp_ref[0][2:8]).max()
        ba_ref = g[tag + "_ba_log"][-1]
        n_p, n_o = slam.map.counts()
        ref_p, ref_o = len(g[tag + "_pt_xyz"]), len(g[tag + "_obs_pt"])
        assert info["cameras"] == g[tag + "_cam_frame"].tolist()
        if d_pose < 1e-6:
            # same consensus set as cv2: everything downstream is the same problem
            assert abs(info["ba"]["chi2_initial"] - ba_ref[3]) <= 1e-8 * ba_ref[3]
            assert abs(info["ba"]["chi2"][-1] - ba_ref[4]) <= 1e-6 * ba_ref[4]
            assert (n_p, n_o) == (ref_p, ref_o)
            X, Xr = slam.map.points[:n_p].cpu().numpy(), g[tag + "_pt_xyz"]
            assert np.abs(X - Xr).max() <= 1e-6 * np.abs(Xr).max()
            poses = slam.map.poses[:len(slam.map.cam_frame)].cpu().numpy()
            assert np.abs(poses - g[tag + "_cam_pose"]).max() < 1e-6
            checked += 1
        else:
            # a different (equally supported) consensus set: the adjusted cost and the map sizes
            # must still be close
            assert abs(info["ba"]["chi2"][-1] - ba_ref[4]) <= 5e-2 * ba_ref[4]
            assert abs(n_p - ref_p) <= max(2, ref_p // 20) and abs(n_o - ref_o) <= max(4, ref_o // 20)
        print(tag, "pnp pose diff", d_pose, "chi2 in", info["ba"]["chi2_initial"], ba_ref[3], "out",
              float(info["ba"]["chi2"][-1]), ba_ref[4], "sizes", (n_p, n_o), (ref_p, ref_o))
    assert checked >= 1
