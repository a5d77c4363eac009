"""GPU: device-resident map bookkeeping (DeviceMap: csrc/tracks.cu + csrc/mapops.cu) through the
C-ABI against (1) the golden produced by the reference's own VisualSlam + Map
(tests/golden/slam_map.npz) and (2) the oracle restatement on a larger scene whose pairs span
several 1024-match tiles.  Bar: integer arrays and copied coordinates bit-exact, order included;
total reprojection error 1e-12 relative (fixed-order tree sum vs sequential sum)."""
import os

import numpy as np
import pytest
import torch

from oracle import map_restate as MR
from visual_odometry_b200.device_map import DeviceMap
from slam_scene import make_scene

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _dev(a, dt):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dt).cuda()


def _same(dm, pts_fid, pts_xyz, obs_cam, obs_pt, obs_xy):
    n_p, n_o = dm.counts()
    assert (n_p, n_o) == (len(pts_xyz), len(obs_pt))
    assert np.array_equal(dm.point_fid[:n_p].cpu().numpy(), pts_fid)
    assert np.array_equal(dm.points[:n_p].cpu().numpy(), pts_xyz)
    assert np.array_equal(dm.obs_cam[:n_o].cpu().numpy(), obs_cam)
    assert np.array_equal(dm.obs_pt[:n_o].cpu().numpy(), obs_pt)
    assert np.array_equal(dm.obs_xy[:n_o].cpu().numpy(), obs_xy)


def test_device_map_equals_reference_golden(cuda_dev):
    z = np.load(os.path.join(GOLDEN, "slam_map.npz"))
    g = {k: z[k] for k in z.files}
    F, N = int(g["F"]), int(g["N"])
    dm = DeviceMap(g["K"], F, N, cap_points=4096, cap_obs=8192)
    kp = _dev(g["kp"], torch.float32)

    def same(tag):
        fid = g[tag + "_pt_fid"]
        _same(dm, fid[:, 0] * N + fid[:, 1], g[tag + "_pt_xyz"], g[tag + "_obs_cam"],
              g[tag + "_obs_pt"], g[tag + "_obs_xy"])

    for b in range(F - 1):
        q, t = _dev(g[f"q_{b}"], torch.int32), _dev(g[f"t_{b}"], torch.int32)
        n = torch.tensor([q.numel()], dtype=torch.int32, device="cuda")
        Xw = _dev(g[f"Xworld_{b}"], torch.float64)
        if b == 0:
            pose1 = g["s0_init_cam_pose"][1].reshape(4, 4)
            dm.initialize(q, t, None, n, 0, 1, kp, Xw, pose1[:3, :3], pose1[:3, 3])
            same("s0_init")
        dm.update_tracks(q, t, None, n, b, b + 1)
        if b + 2 >= 3:
            img, xyz, _, _ = dm.matches_with_map(q, t, None, n, b, b + 1, kp)
            assert np.array_equal(xyz.cpu().numpy(), g[f"s{b}_pnp_obj"])
            assert np.array_equal(img.cpu().numpy(), g[f"s{b}_pnp_img"])
            pose = g[f"s{b}_end_cam_pose"][-1].reshape(4, 4)
            dm.add_camera(pose[:3, :3], pose[:3, 3], b + 1)
            dm.extend(q, t, None, n, b, b + 1, kp, Xw)
            tot = dm.reprojection_error()
            assert abs(tot - g[f"s{b}_totals"][0]) <= 1e-12 * tot
            same(f"s{b}_before_prune")
            dm.prune(float(g[f"s{b}_prune_thr"]))
        same(f"s{b}_end")
        assert np.array_equal(dm.poses[:len(dm.cam_frame)].cpu().numpy(), g[f"s{b}_end_cam_pose"])
    # map culling by the reference's own limit_number_of_camera_in_map, twice
    for k in range(2):
        dm.limit_number_of_cameras(int(g[f"cull{k}_limit"]))
        same(f"cull{k}")
        assert dm.cam_frame == g[f"cull{k}_cam_frame"].tolist()
        assert np.array_equal(dm.poses[:len(dm.cam_frame)].cpu().numpy(), g[f"cull{k}_cam_pose"])
    # mappointdict after culling: every surviving point is found through its feature id
    n_p, _ = dm.counts()
    rp = dm.tracks.root_point[dm.point_fid[:n_p].long()].cpu().numpy()
    assert np.array_equal(rp, np.arange(n_p))


def test_device_map_equals_oracle_large_scene(cuda_dev):
    F, N, NW = 9, 4000, 3600
    K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
    sc = make_scene(F, N, NW, 20261104 + 8, K)
    dm = DeviceMap(K, F, N, cap_points=40000, cap_obs=120000)
    o = MR.SlamMapOracle(K)
    kp = _dev(sc["kp"], torch.float32)
    kp64 = sc["kp"].astype(np.float64)
    n_pnp = 0
    for b, pr in enumerate(sc["pairs"]):
        assert len(pr["q"]) > 2048                      # three tiles of the single-block kernels
        inl = pr["mask"].astype(bool)
        ms = [((b, int(qi)), (b + 1, int(ti)), tuple(kp64[b, qi]), tuple(kp64[b + 1, ti]), x)
              for qi, ti, x in zip(pr["q"][inl], pr["t"][inl], pr["X"])]
        q, t = _dev(pr["q"], torch.int32), _dev(pr["t"], torch.int32)
        mask = _dev(pr["mask"], torch.uint8)
        n = torch.tensor([q.numel()], dtype=torch.int32, device="cuda")
        X = _dev(np.c_[pr["X"], np.ones(len(pr["X"]))], torch.float64)     # [k,4] rows, stride 4
        T = sc["poses"][b + 1]
        if b == 0:
            o.initialize_map(ms, 0, 1, T[:3, :3], T[:3, 3])
            dm.initialize(q, t, mask, n, 0, 1, kp, X, T[:3, :3], T[:3, 3])
        o.update_feature_mapper(ms)
        o.rebuild_mappointdict()
        dm.update_tracks(q, t, mask, n, b, b + 1)
        if b >= 1:
            obj = [o.points[o.mappointdict[r]][1] for r in
                   (o.track_feature_back_in_time(m[1]) for m in ms) if r in o.mappointdict]
            img, xyz, _, _ = dm.matches_with_map(q, t, mask, n, b, b + 1, kp)
            assert np.array_equal(xyz.cpu().numpy(), np.array(obj).reshape(-1, 3))
            n_pnp += len(obj)
            o.add_camera(b + 1, T[:3, :3], T[:3, 3])
            dm.add_camera(T[:3, :3], T[:3, 3], b + 1)
            o.add_information_to_map(ms, b, b + 1)
            dm.extend(q, t, mask, n, b, b + 1, kp, X)
            tot = o.total_reprojection_error()
            assert abs(dm.reprojection_error() - tot) <= 1e-12 * tot
            o.remove_observations_above(1.0)
            kept = dm.prune(1.0)
            assert kept == len(o.observations)
        pts, _poses, oc, op, xy = o.flat()
        _same(dm, np.array([f[0] * N + f[1] for f, _x in o.points]), pts, oc, op, xy)
    assert n_pnp > 5000
    for row in (0, 2, -1):                               # oldest, a middle one, the newest
        row = row % len(o.cameras)
        o.remove_camera_from_map(row)
        dm.remove_camera(row)
        pts, poses, oc, op, xy = o.flat()
        _same(dm, np.array([f[0] * N + f[1] for f, _x in o.points]), pts, oc, op, xy)
        assert np.array_equal(dm.poses[:len(dm.cam_frame)].cpu().numpy(), poses)
        assert dm.cam_frame == [f for f, _T in o.cameras]


def test_device_map_overflow_is_reported_not_written(cuda_dev):
    K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
    sc = make_scene(3, 300, 280, 5, K)
    pr = sc["pairs"][0]
    dm = DeviceMap(K, 3, 300, cap_points=10, cap_obs=50)
    q, t = _dev(pr["q"], torch.int32), _dev(pr["t"], torch.int32)
    n = torch.tensor([q.numel()], dtype=torch.int32, device="cuda")
    X = _dev(pr["X"], torch.float64)
    dm.initialize(q, t, _dev(pr["mask"], torch.uint8), n, 0, 1, _dev(sc["kp"], torch.float32), X,
                  np.eye(3), np.zeros(3))
    with pytest.raises(RuntimeError):
        dm.counts()
    assert int(dm.n_points.item()) == 0 and int(dm.n_obs.item()) == 0
