"""Generates tests/golden/slam_run.npz: the reference's OWN frame loop
(/root/reference/src/visual_slam.py VisualSlam.add_to_list_of_frames +
match_current_and_previous_frame, unmodified, with its own ImagePair / FrameGenerator / Map /
Tracked* classes and cv2 on the CPU) run on a seeded synthetic SIFT sequence.

Replaced, because they cannot run here: the visualisation calls (ImagePair.visualize_matches,
cv2.imshow / waitKey, ThreeDimViewer: no GUI, no graphics stack) and Map.optimize_map (g2o is not
installable: the numpy restatement oracle/ba_restate.py takes its place, same arguments as
src/map.py:104-186 builds).  Everything else — matcher, findEssentialMat, recoverPose,
triangulatePoints, solvePnPRansac, the map bookkeeping and culling — is the reference's code.
The fixture stores the sequence parameters (the sequence itself is regenerated from the seed) and,
after every frame, cameras, points and observations.  Build container only."""
import contextlib
import io
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ba_restate as BA           # noqa: E402
from oracle import load_reference as LR       # noqa: E402
from visual_odometry_b200.frame import SyntheticDetector   # noqa: E402
from visual_odometry_b200.synth import K_REFERENCE, make_sequence  # noqa: E402

F, N, SEED, MAX_CAMS = 8, 400, 20261104 + 9, 5
BA_LOG, PNP_LOG, PNP_ARGS = [], [], []


def oracle_optimize_map(self, postfix=""):
    """stands in for Map.optimize_map (src/map.py:104-186): same vertices, edges, camera
    parameters and fixed flags, optimised by oracle/ba_restate.py instead of g2o"""
    cam_row = {c.camera_id: i for i, c in enumerate(self.cameras)}
    pt_row = {p.point_id: i for i, p in enumerate(self.points)}
    if not self.observations:
        return
    R = np.array([c.R for c in self.cameras], np.float64)
    t = np.array([np.asarray(c.t, np.float64).ravel() for c in self.cameras])
    pr = BA.BAProblem(R, t, [bool(c.fixed) for c in self.cameras],
                      np.array([p.point for p in self.points], np.float64),
                      [cam_row[o.camera_id] for o in self.observations],
                      [pt_row[o.point_id] for o in self.observations],
                      np.array([o.image_coordinates for o in self.observations], np.float64),
                      self.camera_matrix[0, 0], self.camera_matrix[0, 2], self.camera_matrix[1, 2])
    chi0 = pr.robust_chi2(pr.state)
    with np.errstate(all="ignore"):
        log = pr.optimize(40)
    BA_LOG.append((len(self.cameras), len(self.points), len(self.observations), chi0,
                   log[-1][1] if log else chi0, len(log)))
    for i, c in enumerate(self.cameras):
        c.R, c.t = pr.state.R[i].copy(), pr.state.t[i].copy()
    for i, p in enumerate(self.points):
        p.point = pr.state.points[i].copy()


def main():
    import cv2
    mods = LR.load(extra=("visual_slam", "map", "tracked_camera", "tracked_point", "observation"))
    vs, ref_map, ip_mod = mods["visual_slam"], mods["map"], mods["image_pair"]
    seq = make_sequence(F, N, SEED, descriptor="sift", K=K_REFERENCE)
    slam = vs.VisualSlam("/nonexistent")
    slam.set_camera_matrix()
    assert np.allclose(slam.camera_matrix, K_REFERENCE)
    slam.frame_generator = mods["frame_generator"].FrameGenerator(SyntheticDetector(seq))
    ref_map.Map.optimize_map = oracle_optimize_map
    ip_mod.ImagePair.visualize_matches = lambda self, matches: None

    class _cv2:
        def __getattr__(self, k):
            return getattr(cv2, k)

        @staticmethod
        def imshow(*a):
            pass

        @staticmethod
        def solvePnPRansac(obj, img, Kc, dist):
            r = cv2.solvePnPRansac(obj, img, Kc, dist)
            PNP_LOG.append(np.r_[len(obj), float(r[0]), r[1].ravel(), r[2].ravel(),
                                 0 if r[3] is None else len(r[3])])
            PNP_ARGS[:] = [np.array(obj, np.float64), np.array(img, np.float64),
                           np.zeros(0, np.int32) if r[3] is None else r[3].ravel().astype(np.int32)]
            return r

        @staticmethod
        def waitKey(*a):
            return -1

    class _Viewer:
        def main(self):
            pass

    vs.cv2 = _cv2()
    vs.ThreeDimViewer = types.SimpleNamespace(ThreeDimViewer=_Viewer)
    # the reference culls at 18 cameras (src/visual_slam.py:307); a lower limit exercises the same
    # code within a short sequence
    orig_limit = ref_map.Map.limit_number_of_camera_in_map
    ref_map.Map.limit_number_of_camera_in_map = lambda self, n: orig_limit(self, MAX_CAMS)

    save = dict(F=F, N=N, seed=SEED, max_cams=MAX_CAMS, K=K_REFERENCE)
    for f in range(F):
        BA_LOG.clear()
        PNP_LOG.clear()
        PNP_ARGS.clear()
        with contextlib.redirect_stdout(io.StringIO()):
            slam.add_to_list_of_frames(None)
            slam.match_current_and_previous_frame()
        if f == 0:
            continue
        m = slam.map
        cid = {c.camera_id: i for i, c in enumerate(m.cameras)}
        pid = {p.point_id: i for i, p in enumerate(m.points)}
        tag = f"f{f}"
        save[tag + "_cam_frame"] = np.array([c.frame_id for c in m.cameras], np.int32)
        save[tag + "_cam_pose"] = np.array([c.pose() for c in m.cameras]).reshape(-1, 16)
        save[tag + "_cam_fixed"] = np.array([bool(c.fixed) for c in m.cameras])
        save[tag + "_pt_fid"] = np.array([p.feature_id for p in m.points], np.int64).reshape(-1, 2)
        save[tag + "_pt_xyz"] = np.array([p.point for p in m.points], np.float64).reshape(-1, 3)
        save[tag + "_obs_cam"] = np.array([cid[o.camera_id] for o in m.observations], np.int32)
        save[tag + "_obs_pt"] = np.array([pid[o.point_id] for o in m.observations], np.int32)
        save[tag + "_obs_xy"] = np.array([o.image_coordinates for o in m.observations],
                                         np.float64).reshape(-1, 2)
        ipair = slam.current_image_pair
        save[tag + "_ba_log"] = np.array(BA_LOG, np.float64).reshape(-1, 6)
        save[tag + "_pnp_log"] = np.array(PNP_LOG, np.float64).reshape(-1, 9)
        if PNP_ARGS:
            save[tag + "_pnp_obj"], save[tag + "_pnp_img"], save[tag + "_pnp_inliers"] = PNP_ARGS
        save[tag + "_n_matches"] = len(ipair.filtered_matches)
        save[tag + "_n_3d"] = len(ipair.matches_with_3d_information)
        gt = np.eye(4)
        gt[:3, :3], gt[:3, 3] = seq.R_w2c[f], seq.t_w2c[f]
        print(f, "cams", save[tag + "_cam_frame"].tolist(), "points", len(m.points), "obs",
              len(m.observations), "matches", save[tag + "_n_matches"], "3d", save[tag + "_n_3d"])
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "slam_run.npz")
    np.savez_compressed(out, **save)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
