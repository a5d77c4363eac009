"""Generates tests/golden/map_reproj.npz by running the REAL reference Map class
(/root/reference/src/map.py, imported with the GUI/g2o stubs of oracle/load_reference.py) on a
seeded synthetic map.  Run in the build container only (the GPU box has no /root/reference)."""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import load_reference as LR   # noqa: E402


def main():
    mods = LR.load(extra=("map", "tracked_camera", "tracked_point", "observation"))
    ref_map, tracked_camera = mods["map"], mods["tracked_camera"]
    tracked_point, observation = mods["tracked_point"], mods["observation"]
    import cv2
    rng = np.random.default_rng(20261104 + 5)
    K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
    m = ref_map.Map()
    m.camera_matrix = K
    n_cam, n_pt = 7, 400
    cams = []
    for c in range(n_cam):
        R = cv2.Rodrigues(rng.normal(0, 0.05, 3))[0]
        t = rng.normal(0, 0.5, 3) if c % 2 else rng.normal(0, 0.5, (3, 1))   # both shapes occur
        cams.append(m.add_camera(tracked_camera.TrackedCamera(R, t, c, None)))
    pts = [m.add_point(tracked_point.TrackedPoint(
        np.array([rng.uniform(-12, 12), rng.uniform(-7, 7), rng.uniform(8, 40)]), None, None, (0, i)))
        for i in range(n_pt)]
    for c in cams:
        T = c.pose()
        for p in pts:
            if rng.random() < 0.6:
                x = K @ (T @ np.r_[p.point, 1.0])[:3]
                xy = x[:2] / x[2] + rng.normal(0, 0.7, 2)
                if rng.random() < 0.1:
                    xy += rng.normal(0, 15, 2)            # gross outliers
                m.observations.append(observation.Observation(p.point_id, c.camera_id, xy))
    from visual_odometry_b200.map_ops import flatten_map
    points, poses, obs_cam, obs_pt, obs_xy, Kf = flatten_map(m)
    with contextlib.redirect_stdout(io.StringIO()):
        total = np.asarray(m.calculate_reprojection_error(10)).ravel()[0]
    sq_single = np.array([m.calculate_reprojection_error_for_point(
        np.r_[points[obs_pt[i]], 1.0], poses[obs_cam[i]].reshape(4, 4), obs_xy[i]) for i in range(50)])
    n_before = len(m.observations)
    ids_before = [(o.point_id, o.camera_id) for o in m.observations]
    m.remove_observations_with_reprojection_errors_above_threshold(1)
    kept = set((o.point_id, o.camera_id) for o in m.observations)
    keep1 = np.array([k in kept for k in ids_before])
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "map_reproj.npz")
    np.savez_compressed(out, points=points, poses=poses, obs_cam=obs_cam, obs_pt=obs_pt,
                        obs_xy=obs_xy, K=Kf, total=total, sq_single=sq_single, keep1=keep1,
                        n_before=n_before)
    print("wrote", out, "obs", n_before, "kept(thr=1)", keep1.sum(), "total", total)


if __name__ == "__main__":
    main()
