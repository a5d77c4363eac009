"""GPU: N3 bundle adjustment (csrc/ba.cu) through the C-ABI vs the oracle restatement of g2o's
Levenberg / Schur / Huber (oracle/ba_restate.py).  Floating point, fp64: tolerance 1e-8 relative on
the robust chi2 of the first iterations (same LM decisions), 1e-6 on the final state; bit-identical
between two runs (fixed summation order)."""
import numpy as np
import pytest
import torch

from oracle import ba_restate as BA
from visual_odometry_b200 import _lib
from visual_odometry_b200.ba import bundle_adjust
from ba_problem import make_ba

pytestmark = pytest.mark.gpu


def _to_dev(d):
    C = len(d["R"])
    poses = np.tile(np.eye(4), (C, 1, 1))
    poses[:, :3, :3], poses[:, :3, 3] = d["R"], d["t"]
    K = np.array([[d["f"], 0, d["cx"]], [0, d["f"], d["cy"]], [0, 0, 1]])
    g = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dt).cuda()
    return dict(poses=g(poses.reshape(C, 16), torch.float64), fixed=g(d["fixed"], torch.uint8),
                points=g(d["points"], torch.float64), obs_cam=g(d["obs_cam"], torch.int32),
                obs_pt=g(d["obs_pt"], torch.int32), obs_xy=g(d["obs_xy"], torch.float64), K=K)


def _run(d, iterations=40):
    t = _to_dev(d)
    out = bundle_adjust(t["poses"], t["fixed"], t["points"], t["obs_cam"], t["obs_pt"], t["obs_xy"],
                        t["K"], iterations)
    return out, t["poses"].cpu().numpy().reshape(-1, 4, 4), t["points"].cpu().numpy()


@pytest.mark.parametrize("C,P,seed,n_fixed", [(2, 40, 1, 1), (5, 200, 2, 1), (8, 600, 3, 3),
                                               (18, 1500, 4, 16)])
def test_lm_trajectory_equals_oracle(cuda_dev, C, P, seed, n_fixed):
    d = make_ba(C, P, seed, n_fixed=n_fixed)
    pr = BA.BAProblem(**d)
    chi0 = pr.robust_chi2(pr.state)
    n_it = 12
    with np.errstate(all="ignore"):
        log = pr.optimize(n_it)
    out, poses, points = _run(d, n_it)
    assert abs(out["chi2_initial"] - chi0) <= 1e-12 * chi0
    assert out["iterations"] == len(log)
    assert out["trials"].tolist() == [l[3] for l in log]
    ref_chi = np.array([l[1] for l in log])
    assert np.abs(out["chi2"] - ref_chi).max() <= 1e-8 * ref_chi.max()
    assert np.allclose(out["lam"], [l[2] for l in log], rtol=1e-6)
    s = pr.state
    assert np.abs(poses[:, :3, :3] - s.R).max() < 1e-6
    assert np.abs(poses[:, :3, 3] - s.t).max() < 1e-6 * max(1.0, np.abs(s.t).max())
    assert np.abs(points - s.points).max() < 1e-6 * np.abs(s.points).max()
    # fixed cameras are untouched, bit for bit
    assert np.array_equal(poses[d["fixed"], :3, :3], d["R"][d["fixed"]])
    assert np.array_equal(poses[d["fixed"], :3, 3], d["t"][d["fixed"]])


def test_forty_iterations_reduce_the_cost_like_the_oracle_and_are_reproducible(cuda_dev):
    d = make_ba(6, 400, 9)
    pr = BA.BAProblem(**d)
    with np.errstate(all="ignore"):
        log = pr.optimize(40)
    out1, poses1, points1 = _run(d, 40)
    out2, poses2, points2 = _run(d, 40)
    assert np.array_equal(poses1, poses2) and np.array_equal(points1, points2)
    assert np.array_equal(out1["chi2"], out2["chi2"])
    # LM decisions may part ways late in a flat valley; the cost reached must agree closely
    assert abs(out1["chi2"][-1] - log[-1][1]) <= 1e-4 * log[-1][1]
    assert np.all(np.diff(np.r_[out1["chi2_initial"], out1["chi2"]]) <= 0)


def test_inactive_vertices_and_no_robust_kernel(cuda_dev):
    d = make_ba(4, 80, 13)
    # point 5 and camera 3 lose every observation: they must come back unchanged
    keep = (d["obs_pt"] != 5) & (d["obs_cam"] != 3)
    for k in ("obs_cam", "obs_pt", "obs_xy"):
        d[k] = d[k][keep]
    pr = BA.BAProblem(**d)
    with np.errstate(all="ignore"):
        log = pr.optimize(8)
    out, poses, points = _run(d, 8)
    assert np.array_equal(points[5], d["points"][5])
    assert np.array_equal(poses[3, :3, :3], d["R"][3]) and np.array_equal(poses[3, :3, 3], d["t"][3])
    assert np.abs(out["chi2"] - [l[1] for l in log]).max() <= 1e-8 * log[0][1]
    # huber_delta <= 0: plain least squares; cost = sum of squared errors
    t = _to_dev(d)
    o2 = bundle_adjust(t["poses"], t["fixed"], t["points"], t["obs_cam"], t["obs_pt"], t["obs_xy"],
                       t["K"], 5, huber_delta=0.0)
    e = pr.errors(BA.BAProblem(**d).state)
    assert abs(o2["chi2_initial"] - float((e ** 2).sum())) <= 1e-12 * float((e ** 2).sum())
    assert o2["chi2"][-1] < o2["chi2_initial"]


def test_argument_errors(cuda_dev):
    d = make_ba(3, 20, 1)
    t = _to_dev(d)
    bad = t["obs_pt"].clone()
    bad[3] = 999
    with pytest.raises(_lib.VoError):
        bundle_adjust(t["poses"], t["fixed"], t["points"], t["obs_cam"], bad, t["obs_xy"], t["K"], 3)
    with pytest.raises(_lib.VoError):
        bundle_adjust(torch.zeros(25, 16, dtype=torch.float64, device="cuda"),
                      torch.zeros(25, dtype=torch.uint8, device="cuda"), t["points"], t["obs_cam"],
                      t["obs_pt"], t["obs_xy"], t["K"], 3)


def test_drop_in_optimize_map_and_device_map(cuda_dev):
    """Map.optimize_map drop-in on objects shaped like the reference's (cameras with R, t, fixed;
    points with .point) and DeviceMap.optimize on the flat map: reprojection error drops, fixed
    camera untouched."""
    from visual_odometry_b200.ba import optimize_map
    from visual_odometry_b200.map_ops import reprojection_errors, flatten_map

    class Cam:
        def __init__(s, R, t, fixed, i):
            s.R, s.t, s.fixed, s.camera_id, s.frame_id = R.copy(), t.copy(), fixed, 100 + i, i

        def pose(s):
            T = np.eye(4); T[:3, :3] = s.R; T[:3, 3] = s.t
            return T

    class Pt:
        def __init__(s, p, i):
            s.point, s.point_id = p.copy(), 500 + i

    class Obs:
        def __init__(s, p, c, xy):
            s.point_id, s.camera_id, s.image_coordinates = 500 + int(p), 100 + int(c), xy

    class M:
        pass
    d = make_ba(5, 150, 21)
    m = M()
    m.cameras = [Cam(d["R"][i], d["t"][i], bool(d["fixed"][i]), i) for i in range(5)]
    m.points = [Pt(p, i) for i, p in enumerate(d["points"])]
    m.observations = [Obs(p, c, xy) for p, c, xy in zip(d["obs_pt"], d["obs_cam"], d["obs_xy"])]
    m.camera_matrix = np.array([[d["f"], 0, d["cx"]], [0, d["f"], d["cy"]], [0, 0, 1]])
    before = reprojection_errors(*flatten_map(m), 10)["total"]
    out = optimize_map(m)
    after = reprojection_errors(*flatten_map(m), 10)["total"]
    assert out["iterations"] >= 1 and after < 0.5 * before
    assert np.array_equal(m.cameras[0].R, d["R"][0]) and np.array_equal(m.cameras[0].t, d["t"][0])
