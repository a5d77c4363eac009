"""Runs the reference's OWN, unmodified `visual_slam.py` — `VisualSlam.run()` over a directory of
image files, its live configuration (SIFT + BFMatcher(NORM_L2, crossCheck), src/visual_slam.py:17-19)
— with every stage on the B200: SIFT extraction (csrc/sift.cu), L2 matching (csrc/l2.cu),
findEssentialMat / recoverPose / triangulatePoints, solvePnPRansac and the Map's adjustment
(compat.install(patch_map=True)).  Control flow, bookkeeping and the Map / Tracked* classes are the
reference's.  Frames: a rendered sequence with known camera motion (three textured planes), written
as .jpg files into a scratch directory; the estimated inter-frame rotations are compared with the
truth and with cv2's own two-view chain on the same files.

Reference sources: see tools/gpu_reference_loop.py (read from /root/reference/src, or from the
git-ignored baseline/_ref/reference_src on the GPU box).  usage: python tools/gpu_reference_loop_images.py [n_frames]
"""
import contextlib
import importlib
import io
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def rot_angle(Ra, Rb):
    return float(np.degrees(np.arccos(np.clip((np.trace(Ra.T @ Rb) - 1) / 2, -1, 1))))


def main():
    from gpu_reference_loop import find_reference
    ref = find_reference()
    if ref is None:
        print("reference sources not on this box (see the docstring): nothing run")
        return 0
    import cv2 as real_cv2
    import torch
    from test_gpu_orb import _render_two_plane_sequence
    from visual_odometry_b200 import compat
    from visual_odometry_b200.sift import SiftDetector
    n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    images, K, cams = _render_two_plane_sequence(n_frames)
    tmp = tempfile.mkdtemp(prefix="vo_frames_")
    for k, img in enumerate(images):
        real_cv2.imwrite(os.path.join(tmp, "frame_%03d.jpg" % k), img, [real_cv2.IMWRITE_JPEG_QUALITY, 100])
    sys.dont_write_bytecode = True
    os.environ.setdefault("VO_B200_HEADLESS_QUIT", "3")
    compat.install(reference_src=ref, patch_map=True)
    vs = importlib.import_module("visual_slam")          # the reference's file, unmodified
    assert os.path.dirname(os.path.abspath(vs.__file__)) == os.path.abspath(ref), vs.__file__
    slam = vs.VisualSlam(tmp)
    slam.set_camera_matrix()
    slam.camera_matrix = K.copy()                        # the rendered camera, not the DJI one
    slam.scale_factor = 1
    assert isinstance(slam.detector, SiftDetector), type(slam.detector)
    print("reference file:", vs.__file__, "| detector:", type(slam.detector).__name__, "| matcher:",
          type(slam.bf).__name__)
    rel, stats = [], []
    orig = slam.match_current_and_previous_frame

    def spy():                                           # records what the reference computed
        orig()
        p = getattr(slam, "current_image_pair", None)
        if p is not None and len(slam.list_of_frames) >= 2:
            rel.append(np.array(p.R, float))
            stats.append((len(slam.list_of_frames[-1].keypoints), len(p.filtered_matches), len(slam.map.points),
                          len(slam.map.cameras)))

    slam.match_current_and_previous_frame = spy
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        slam.run()                                       # src/visual_slam.py:338-360
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ok = len(rel) == n_frames - 1
    sift, bf = real_cv2.SIFT_create(), real_cv2.BFMatcher(real_cv2.NORM_L2, crossCheck=True)
    files = sorted(os.listdir(tmp))
    for k in range(len(rel)):
        (R1, _), (R2, _) = cams[k], cams[k + 1]
        R_true = R2 @ R1.T
        a = real_cv2.imread(os.path.join(tmp, files[k])); b = real_cv2.imread(os.path.join(tmp, files[k + 1]))
        k1, d1 = sift.detectAndCompute(a, None); k2, d2 = sift.detectAndCompute(b, None)
        m = bf.match(d1, d2)
        p1 = np.float64([k1[x.queryIdx].pt for x in m]); p2 = np.float64([k2[x.trainIdx].pt for x in m])
        E, mask = real_cv2.findEssentialMat(p1, p2, K, real_cv2.FM_RANSAC, 0.99, 1)
        _, Rc, _, _ = real_cv2.recoverPose(E, p1[mask.ravel() == 1], p2[mask.ravel() == 1], K)
        line = {"pair": k, "key_points": stats[k][0], "cv2_key_points": len(k2), "matches": stats[k][1],
                "cv2_matches": len(m), "map_points": stats[k][2], "cameras": stats[k][3],
                "rotation_error_deg_vs_truth": round(rot_angle(rel[k], R_true), 3),
                "cv2_chain_rotation_error_deg_vs_truth": round(rot_angle(Rc, R_true), 3),
                "rotation_diff_deg_vs_cv2_chain": round(rot_angle(rel[k], Rc), 4)}
        ok &= line["key_points"] == line["cv2_key_points"] and abs(line["matches"] - line["cv2_matches"]) <= 2
        ok &= line["rotation_error_deg_vs_truth"] < 2.0
        print(line, flush=True)
    print("frames:", n_frames, "| wall s:", round(wall, 2), "| same key-point counts and match counts as "
          "cv2 on the same files, rotations within 2 degrees of the truth:", ok)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
