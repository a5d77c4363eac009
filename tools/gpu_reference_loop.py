"""Runs the reference's OWN, unmodified `visual_slam.py` frame loop on the B200 path.

VERDICT r1 #9: `VisualSlam.process_frame` -> `match_current_and_previous_frame`
(reference src/visual_slam.py:288-336) over `compat.install(patch_map=True)`: the reference file
imports OUR `initials` / `image_pair` / `frame` / `frame_generator` modules (seam 2), so every
cv2 call of the hot path (matcher, findEssentialMat, recoverPose, triangulatePoints,
solvePnPRansac, Rodrigues) and the reference `Map`'s optimize_map / reprojection loops run in
libvo_b200; the control flow, the dict bookkeeping and the Map / Tracked* classes are the
reference's.  Frames come from the seeded synthetic SIFT sequence of tests/golden/slam_run.npz
(made by the reference's loop on cv2 + the BA oracle), which the result is compared with.

The reference sources are NOT part of this repo.  On the build container they are read from
/root/reference/src; for the GPU box copy them into the git-ignored scratch directory
    mkdir -p baseline/_ref/reference_src && cp /root/reference/src/*.py baseline/_ref/reference_src/
(baseline/_ref is the sanctioned, git-ignored place of the reference install; it travels with
gpurun).  usage: python tools/gpu_reference_loop.py [n_frames]
"""
import contextlib
import importlib
import io
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def find_reference():
    for p in (os.environ.get("VO_REFERENCE_SRC"), "/root/reference/src",
              os.path.join(ROOT, "baseline", "_ref", "reference_src")):
        if p and os.path.isfile(os.path.join(p, "visual_slam.py")):
            return p
    return None


def main():
    ref = find_reference()
    if ref is None:
        print("reference sources not on this box (see the docstring): nothing run")
        return 0
    import torch
    from visual_odometry_b200 import compat
    from visual_odometry_b200.frame import SyntheticDetector
    from visual_odometry_b200.synth import K_REFERENCE, make_sequence
    sys.dont_write_bytecode = True
    compat.install(reference_src=ref, patch_map=True)
    vs = importlib.import_module("visual_slam")          # the reference's file, unmodified
    assert os.path.dirname(os.path.abspath(vs.__file__)) == os.path.abspath(ref), vs.__file__
    import frame_generator
    import image_pair
    assert "visual_odometry_b200" in image_pair.ImagePair.__module__

    g = np.load(os.path.join(ROOT, "tests", "golden", "slam_run.npz"))
    F, N, seed, max_cams = int(g["F"]), int(g["N"]), int(g["seed"]), int(g["max_cams"])
    F = min(F, int(sys.argv[1])) if len(sys.argv) > 1 else F
    seq = make_sequence(F, N, seed, descriptor="sift", K=K_REFERENCE)
    slam = vs.VisualSlam("/nonexistent")
    slam.set_camera_matrix()
    slam.frame_generator = frame_generator.FrameGenerator(SyntheticDetector(seq))
    ref_map = importlib.import_module("map")
    orig_limit = ref_map.Map.limit_number_of_camera_in_map
    ref_map.Map.limit_number_of_camera_in_map = lambda self, n: orig_limit(self, max_cams)
    print("reference file:", vs.__file__)
    print("matcher:", type(slam.bf).__module__, "| cv2 in visual_slam:", type(vs.cv2).__name__,
          "| Map.optimize_map:", ref_map.Map.optimize_map.__module__)
    img = np.zeros((int(2 * K_REFERENCE[1, 2]), int(2 * K_REFERENCE[0, 2]), 3), np.uint8)
    ok = True
    for f in range(F):
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            slam.process_frame(img)                       # src/visual_slam.py:333-336
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        if f == 0:
            continue
        m = slam.map
        tag = f"f{f}"
        cams = [c.frame_id for c in m.cameras]
        line = {"frame": f, "ms": round(ms, 1), "cameras": cams, "points": len(m.points),
                "observations": len(m.observations),
                "matches": len(slam.current_image_pair.filtered_matches),
                "golden_cameras": g[tag + "_cam_frame"].tolist(), "golden_points": int(len(g[tag + "_pt_xyz"])),
                "golden_observations": int(len(g[tag + "_obs_xy"])), "golden_matches": int(g[tag + "_n_matches"])}
        ok &= line["matches"] == line["golden_matches"]
        if f == 1:
            # initialize_map + the first adjustment: the map is still a deterministic function of
            # the two-view result, so the camera poses must agree with the reference-on-cv2 run
            pose = np.array([c.pose() for c in m.cameras]).reshape(-1, 16)
            line["max_abs_pose_diff_vs_golden"] = float(np.abs(pose - g[tag + "_cam_pose"]).max())
            ok &= cams == line["golden_cameras"] and line["points"] == line["golden_points"]
            ok &= line["max_abs_pose_diff_vs_golden"] < 1e-3
        print(line, flush=True)
    gt_last = np.eye(4)
    gt_last[:3, :3], gt_last[:3, 3] = seq.R_w2c[F - 1], seq.t_w2c[F - 1]
    print("frames processed:", F, "| per-pair match counts equal to the reference-on-cv2 run, "
          "frame-1 map equal:", ok)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
