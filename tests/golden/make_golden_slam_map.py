"""Generates tests/golden/slam_map.npz by running the REAL reference VisualSlam + Map
(/root/reference/src/visual_slam.py, map.py, tracked_*.py, observation.py; GUI/g2o stubs of
oracle/load_reference.py) through its per-frame map bookkeeping on a seeded synthetic sequence:

  initialize_map (:43-87)  ->  estimate_current_camera_position (:190-268) per further frame =
  update_feature_mapper, mappointdict, MatchWithMap -> solvePnPRansac arguments, add camera,
  add_information_to_map (:152-181), calculate_reprojection_error, optimize_map (g2o: absent, made
  a no-op), remove_observations_with_reprojection_errors_above_threshold(1).

Replaced in the reference's namespace: cv2.solvePnPRansac (returns the scene's true pose, records
its arguments) and Map.optimize_map (no-op).  The image pair is a stand-in that carries synthetic
matches_with_3d_information (the hot path's own output is covered by the other fixtures).
Build container only (the GPU box has no /root/reference)."""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import load_reference as LR   # noqa: E402

F, N, NW = 6, 160, 260       # frames, features per frame, world points


class _Frame:
    def __init__(self, i):
        self.id, self.image = i, None


class _Pair:
    """what VisualSlam reads of an ImagePair (src/visual_slam.py:47-66, 165-172, 184, 203)"""

    def __init__(self, frame1, frame2, rel_matches, world_matches, R, t):
        self.frame1, self.frame2, self.R, self.t = frame1, frame2, R, t
        self.matches_with_3d_information = rel_matches
        self.filtered_matches = []
        self._world = world_matches

    def determine_essential_matrix(self, matches):
        return []

    def reconstruct_3d_points(self, matches, P1, P2):
        self.matches_with_3d_information = self._world


def main():
    import cv2
    mods = LR.load(extra=("visual_slam", "map", "tracked_camera", "tracked_point", "observation"))
    vs = mods["visual_slam"]
    Match3D = mods["initials"].Match3D
    rng = np.random.default_rng(20261104 + 7)
    slam = vs.VisualSlam("/nonexistent")
    slam.set_camera_matrix()
    K = slam.camera_matrix
    slam.map.optimize_map = lambda *a, **k: None

    # scene: world points in front of a slowly moving camera; a few far away (|X| > 50)
    W = np.c_[rng.uniform(-12, 12, NW), rng.uniform(-7, 7, NW), rng.uniform(8, 40, NW)]
    W[::17, 2] += 40.0
    poses = [np.eye(4)]
    for f in range(1, F):
        T = np.eye(4)
        T[:3, :3] = cv2.Rodrigues(rng.normal(0, 0.02, 3))[0]
        T[:3, 3] = rng.normal(0, 0.3, 3)
        poses.append(T @ poses[-1])
    vis = rng.random((F, NW)) < 0.75
    slot = np.array([rng.permutation(N) for _ in range(F)])          # world point j -> feature idx
    kp = np.zeros((F, N, 2), np.float32)
    for f in range(F):
        x = (K @ (poses[f] @ np.c_[W, np.ones(NW)].T)[:3]).T
        xy = x[:, :2] / x[:, 2:3] + rng.normal(0, 0.3, (NW, 2))
        gross = rng.random(NW) < 0.1
        xy[gross] += rng.normal(0, 4, (int(gross.sum()), 2))
        kp[f] = rng.uniform(0, 1000, (N, 2))
        for j in range(min(NW, N)):
            kp[f, slot[f, j]] = xy[j]
    n_track = min(NW, N)

    def matches_of(b, world):
        js = [j for j in rng.permutation(n_track) if vis[b, j] and vis[b + 1, j]]
        js = [j for j in js if rng.random() < 0.9]
        out = []
        for j in js:
            i1, i2 = int(slot[b, j]), int(slot[b + 1, j])
            out.append((j, i1, i2))
        return out

    captured = {}

    class _cv2:
        def __getattr__(self, k):
            return getattr(cv2, k)

        @staticmethod
        def solvePnPRansac(obj, img, Kc, dist):
            captured["obj"], captured["img"] = np.array(obj, np.float64), np.array(img, np.float64)
            T = captured["pose"]
            return True, cv2.Rodrigues(T[:3, :3])[0], T[:3, 3:4].copy(), None

    vs.cv2 = _cv2()
    snaps = {}
    orig_remove = slam.map.remove_observations_with_reprojection_errors_above_threshold
    orig_total = slam.map.calculate_reprojection_error

    def snap_map(tag):
        m = slam.map
        ids = {c.camera_id: i for i, c in enumerate(m.cameras)}
        pid = {p.point_id: i for i, p in enumerate(m.points)}
        snaps[tag + "_pt_fid"] = np.array([p.feature_id for p in m.points], np.int64).reshape(-1, 2)
        snaps[tag + "_pt_xyz"] = np.array([p.point for p in m.points], np.float64).reshape(-1, 3)
        snaps[tag + "_obs_cam"] = np.array([ids[o.camera_id] for o in m.observations], np.int32)
        snaps[tag + "_obs_pt"] = np.array([pid[o.point_id] for o in m.observations], np.int32)
        snaps[tag + "_obs_xy"] = np.array([o.image_coordinates for o in m.observations],
                                          np.float64).reshape(-1, 2)
        snaps[tag + "_cam_pose"] = np.array([c.pose() for c in m.cameras]).reshape(-1, 16)
        snaps[tag + "_cam_frame"] = np.array([c.frame_id for c in m.cameras], np.int32)

    frames = [_Frame(i) for i in range(F)]
    save = dict(F=F, N=N, K=K, kp=kp)
    for b in range(F - 1):
        tr = matches_of(b, W)
        rel, world = [], []
        for j, i1, i2 in tr:
            k1 = tuple(float(v) for v in kp[b, i1])
            k2 = tuple(float(v) for v in kp[b + 1, i2])
            pw = W[j] + rng.normal(0, 0.01, 3)
            pr = (poses[b + 1] @ np.r_[W[j], 1.0])[:3] + rng.normal(0, 0.01, 3)
            rel.append(Match3D((b, i1), (b + 1, i2), k1, k2, None, None, 0.0, None, pr))
            world.append(Match3D((b, i1), (b + 1, i2), k1, k2, None, None, 0.0, None, pw))
        save[f"q_{b}"] = np.array([m.featureid1[1] for m in rel], np.int32)
        save[f"t_{b}"] = np.array([m.featureid2[1] for m in rel], np.int32)
        save[f"Xrel_{b}"] = np.array([m.point for m in rel])
        save[f"Xworld_{b}"] = np.array([m.point for m in world])
        Trel = poses[b + 1] @ np.linalg.inv(poses[b])
        pair = _Pair(frames[b], frames[b + 1], rel, world, Trel[:3, :3].copy(), Trel[:3, 3:4].copy())
        slam.list_of_frames = frames[:b + 2]
        slam.current_image_pair = pair
        captured.clear()
        captured["pose"] = poses[b + 1]
        totals = []

        def total_wrap(threshold=10):
            r = orig_total(threshold)
            totals.append(float(np.asarray(r).ravel()[0]))
            return r

        def remove_wrap(threshold=100):
            snap_map(f"s{b}_before_prune")
            save[f"s{b}_prune_thr"] = float(threshold)
            return orig_remove(threshold)

        slam.map.calculate_reprojection_error = total_wrap
        slam.map.remove_observations_with_reprojection_errors_above_threshold = remove_wrap
        with contextlib.redirect_stdout(io.StringIO()):
            if b == 0:
                # the map of initialize_map uses the first pair's own (relative = world) points
                pair.matches_with_3d_information = world
                slam.initialize_map(pair)
                snap_map("s0_init")
            slam.estimate_current_camera_position(frames[b + 1])
        snap_map(f"s{b}_end")
        save[f"s{b}_pnp_called"] = int("obj" in captured)
        save[f"s{b}_pnp_obj"] = captured.get("obj", np.zeros((0, 3))).reshape(-1, 3)
        save[f"s{b}_pnp_img"] = captured.get("img", np.zeros((0, 2))).reshape(-1, 2)
        save[f"s{b}_totals"] = np.array(totals)
    # map culling by the reference's own Map.limit_number_of_camera_in_map / remove_camera_from_map
    # (src/map.py:188-232, 301-318): twice, so camera 0 and then the next-oldest go
    for k, limit in enumerate((len(slam.map.cameras) - 1, len(slam.map.cameras) - 2)):
        with contextlib.redirect_stdout(io.StringIO()):
            slam.map.limit_number_of_camera_in_map(limit)
        snap_map(f"cull{k}")
        save[f"cull{k}_limit"] = limit
    save.update(snaps)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "slam_map.npz")
    np.savez_compressed(out, **save)
    for b in range(F - 1):
        print(b, "matches", len(save[f"q_{b}"]), "pnp", len(save[f"s{b}_pnp_obj"]),
              "points", len(snaps[f"s{b}_end_pt_xyz"]), "obs", len(snaps[f"s{b}_end_obs_pt"]),
              "before prune", len(snaps.get(f"s{b}_before_prune_obs_pt", [])),
              "totals", save[f"s{b}_totals"])
    for k in range(2):
        print("cull", k, "cams", len(snaps[f"cull{k}_cam_frame"]), "points", len(snaps[f"cull{k}_pt_xyz"]),
              "obs", len(snaps[f"cull{k}_obs_pt"]))
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
