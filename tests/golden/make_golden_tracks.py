"""Generates tests/golden/tracks.npz by running the REAL reference VisualSlam
(/root/reference/src/visual_slam.py, imported with the GUI/g2o stubs of oracle/load_reference.py):
update_feature_mapper, track_feature_back_in_time and the MatchWithMap loop of
estimate_current_camera_position (:190-235) on a seeded synthetic match sequence.  The arguments
the reference hands to cv2.solvePnPRansac are captured by replacing that one function in the
reference module's namespace.  Build container only (the GPU box has no /root/reference)."""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import load_reference as LR   # noqa: E402

F, N = 7, 120            # frames, features per frame


class _Captured(BaseException):
    """leaves estimate_current_camera_position past its `except Exception` (:258)"""


class _Pair:
    def __init__(self, matches):
        self.matches_with_3d_information = matches


def synth_matches(rng, b, Match3D):
    """inlier matches of pair (b, b+1): a random partial injection, random order"""
    k = int(rng.integers(N - 30, N - 5))
    q = rng.permutation(N)[:k]
    t = rng.permutation(N)[:k]
    out = []
    for i in range(k):
        # float32-representable pixel coordinates (cv2.KeyPoint.pt is float32)
        kp1 = tuple(float(v) for v in rng.uniform(0, 1000, 2).astype(np.float32))
        kp2 = tuple(float(v) for v in rng.uniform(0, 1000, 2).astype(np.float32))
        out.append(Match3D((b, int(q[i])), (b + 1, int(t[i])), kp1, kp2, None, None, 0.0, None,
                           rng.normal(0, 1, 3)))
    return out


def main():
    mods = LR.load(extra=("visual_slam", "map", "tracked_camera", "tracked_point", "observation"))
    vs, tracked_point = mods["visual_slam"], mods["tracked_point"]
    Match3D = mods["initials"].Match3D
    rng = np.random.default_rng(20261104 + 6)
    slam = vs.VisualSlam("/nonexistent")
    slam.set_camera_matrix()
    captured = {}

    class _cv2:
        def __getattr__(self, k):
            return getattr(mods["initials"].cv2, k)

        @staticmethod
        def solvePnPRansac(obj, img, K, dist):
            captured["obj"], captured["img"] = np.array(obj, np.float64), np.array(img, np.float64)
            raise _Captured()

    vs.cv2 = _cv2()
    all_matches = [synth_matches(rng, b, Match3D) for b in range(F - 1)]
    # map: points keyed by featureid1 of pair 0 (initialize_map :49-52, :69-75), later additions
    # keyed by featureid1 of pairs 2 and 3 (add_new_match_to_map :101-107), and a few points
    # whose feature never occurs in any match
    map_ids, map_xyz = [], []

    def add_points(ids):
        for fid in ids:
            map_ids.append(fid)
            map_xyz.append(rng.uniform(-20, 20, 3))

    add_points([m.featureid1 for m in all_matches[0]][::2])
    steps = []
    for b in range(F - 1):
        if b in (2, 3):
            add_points([m.featureid1 for m in all_matches[b]][::5])
            add_points([(b, N + 3), (b + 1, N + 4)])
        slam.current_image_pair = _Pair(all_matches[b])
        slam.list_of_frames = [None] * (b + 2)
        slam.map.points = [tracked_point.TrackedPoint(xyz, None, None, fid)
                           for fid, xyz in zip(map_ids, map_xyz)]
        captured.clear()
        with contextlib.redirect_stdout(io.StringIO()):
            try:
                slam.estimate_current_camera_position(None)
            except _Captured:
                pass
        roots = np.array([slam.track_feature_back_in_time((b + 1, i)) for i in range(N)], np.int64)
        steps.append(dict(n_map=len(map_ids), roots=roots,
                          obj=captured.get("obj", np.zeros((0, 3))).reshape(-1, 3),
                          img=captured.get("img", np.zeros((0, 2))).reshape(-1, 2),
                          called=int("obj" in captured)))
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tracks.npz")
    save = dict(F=F, N=N, map_ids=np.array(map_ids, np.int64), map_xyz=np.array(map_xyz))
    for b, ms in enumerate(all_matches):
        save[f"f1_{b}"] = np.array([m.featureid1 for m in ms], np.int64)
        save[f"f2_{b}"] = np.array([m.featureid2 for m in ms], np.int64)
        save[f"kp2_{b}"] = np.array([m.keypoint2 for m in ms], np.float64)
        for k, v in steps[b].items():
            save[f"{k}_{b}"] = np.asarray(v)
    np.savez_compressed(out, **save)
    print("wrote", out, [(s["called"], len(s["obj"])) for s in steps])


if __name__ == "__main__":
    main()
