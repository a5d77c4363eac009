"""GPU: N2 reprojection-error kernel through the C-ABI vs the oracle and the reference golden.
Tolerance: 1e-12 relative on every squared error and on the total (fp64; the reference's numpy
matmul order is restated without FMA, the total is a tree sum); kept sets identical."""
import os

import numpy as np
import pytest

from oracle import map_restate as MR
from visual_odometry_b200 import map_ops as MO
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _g():
    z = np.load(os.path.join(GOLDEN, "map_reproj.npz"))
    return {k: z[k] for k in z.files}


def test_reprojection_equals_reference_golden(cuda_dev):
    g = _g()
    out = MO.reprojection_errors(g["points"], g["poses"], g["obs_cam"], g["obs_pt"], g["obs_xy"],
                                 g["K"], 1.0)
    sq = out["sqerr"].cpu().numpy()
    ref = MR.reprojection_sq_errors(g["points"], g["poses"], g["obs_cam"], g["obs_pt"], g["obs_xy"], g["K"])
    assert np.abs(sq - ref).max() <= 1e-12 * ref.max()
    assert abs(out["total"] - float(g["total"])) <= 1e-12 * float(g["total"])
    assert np.array_equal(out["keep"].cpu().numpy(), g["keep1"])
    assert out["n_keep"] == int(g["keep1"].sum())


def test_map_object_dropin_methods(cuda_dev):
    """the patched methods on a duck-typed Map (same attributes as the reference's classes)."""
    g = _g()

    class Cam:
        def __init__(self, T, cid):
            self.T, self.camera_id = T, cid

        def pose(self):
            return self.T

    class Pt:
        def __init__(self, p, pid):
            self.point, self.point_id = p, pid

    class Obs:
        def __init__(self, pid, cid, xy):
            self.point_id, self.camera_id, self.image_coordinates = pid, cid, xy

    class Map:
        pass

    MO.patch_map(Map)
    m = Map()
    m.camera_matrix = g["K"]
    m.cameras = [Cam(T.reshape(4, 4), 100 + i) for i, T in enumerate(g["poses"])]
    m.points = [Pt(p, 5000 + i) for i, p in enumerate(g["points"])]
    m.observations = [Obs(5000 + p, 100 + c, xy) for c, p, xy in zip(g["obs_cam"], g["obs_pt"], g["obs_xy"])]
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()) as buf:
        total = m.calculate_reprojection_error(10)
    assert total.shape == (1,) and abs(total[0] - float(g["total"])) <= 1e-12 * float(g["total"])
    ref = MR.reprojection_sq_errors(g["points"], g["poses"], g["obs_cam"], g["obs_pt"], g["obs_xy"], g["K"])
    assert buf.getvalue().count("high reprojection error") == int((ref > 10).sum())
    m.remove_observations_with_reprojection_errors_above_threshold(1)
    assert len(m.observations) == int(g["keep1"].sum())
    kept_xy = np.array([o.image_coordinates for o in m.observations])
    assert np.array_equal(kept_xy, g["obs_xy"][g["keep1"]])
    e = MO.calculate_reprojection_error_for_point(m, np.r_[g["points"][g["obs_pt"][3]], 1.0],
                                                  g["poses"][g["obs_cam"][3]].reshape(4, 4), g["obs_xy"][3])
    assert abs(e - g["sq_single"][3]) <= 1e-12 * max(g["sq_single"][3], 1.0)


def test_large_map_and_edge_cases(cuda_dev):
    """C5 scale: 100 k points, 18 cameras, 1 M observations (one launch); empty map; bad index."""
    rng = np.random.default_rng(3)
    P, C, O = 100_000, 18, 1_000_000
    K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
    pts = np.c_[rng.uniform(-12, 12, P), rng.uniform(-7, 7, P), rng.uniform(8, 40, P)]
    poses = np.tile(np.eye(4), (C, 1, 1))
    poses[:, :3, 3] = rng.normal(0, 0.3, (C, 3))
    oc, op = rng.integers(0, C, O).astype(np.int32), rng.integers(0, P, O).astype(np.int32)
    xy = rng.uniform(0, 1920, (O, 2))
    out = MO.reprojection_errors(pts, poses, oc, op, xy, K, 1e5)
    ref = MR.reprojection_sq_errors(pts, poses, oc, op, xy, K)
    sq = out["sqerr"].cpu().numpy()
    assert np.abs(sq - ref).max() <= 1e-12 * ref.max()
    assert abs(out["total"] - ref.sum()) <= 1e-11 * ref.sum()
    assert np.array_equal(out["keep"].cpu().numpy(), ref < 1e5)
    empty = MO.reprojection_errors(pts, poses, [], [], np.zeros((0, 2)), K, 1.0)
    assert empty["total"] == 0.0 and empty["n_keep"] == 0 and empty["sqerr"].numel() == 0
    bad = MO.reprojection_errors(pts, poses, [0, 99], [0, 5], np.zeros((2, 2)), K, 1.0)
    assert np.isnan(bad["sqerr"][1].item()) and not bool(bad["keep"][1])
