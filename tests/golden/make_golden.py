"""Generates tests/golden/*.npz by running the REAL reference ImagePair (imported from
/root/reference/src with GUI/BA stubs, oracle/load_reference.py) on seeded synthetic frames.
Run in the build container only:  python tests/golden/make_golden.py
Each file stores the inputs (descriptors, keypoints, K) and every hot-path output of
ImagePair.match_features / determine_essential_matrix / estimate_camera_movement /
reconstruct_3d_points (reference src/image_pair.py:233-354), plus the cv2 version used.
"""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402
from oracle import load_reference  # noqa: E402
from visual_odometry_b200.synth import make_sequence, K_DEFAULT, K_REFERENCE  # noqa: E402
from visual_odometry_b200.frame import SyntheticDetector  # noqa: E402

CASES = [  # name, descriptor, n_kp, seed, inlier_ratio, K
    ("orb_300", "orb", 300, 101, 0.7, K_DEFAULT),
    ("orb_1000_c1", "orb", 1000, 20261104, 0.7, K_DEFAULT),     # BASELINE.json configs[0]
    ("orb_700_refK", "orb", 700, 103, 0.5, K_REFERENCE),
    ("sift_400", "sift", 400, 104, 0.7, K_DEFAULT),
]


def run_case(ref, name, descriptor, n_kp, seed, inl, K):
    seq = make_sequence(2, n_kp, seed, inlier_ratio=inl, descriptor=descriptor, K=K)
    fg = ref["frame_generator"].FrameGenerator(SyntheticDetector(seq))
    f1 = fg.make_frame(np.zeros((4, 4, 3), np.uint8))
    f2 = fg.make_frame(np.zeros((4, 4, 3), np.uint8))
    norm = cv2.NORM_HAMMING if descriptor == "orb" else cv2.NORM_L2
    ip = ref["image_pair"].ImagePair(f1, f2, cv2.BFMatcher(norm, crossCheck=True), K)
    with contextlib.redirect_stdout(io.StringIO()):
        ip.match_features()
        inliers = ip.determine_essential_matrix(ip.filtered_matches)
        ip.estimate_camera_movement(inliers)
        ip.reconstruct_3d_points(inliers)
    q = np.array([m.featureid1[1] for m in ip.filtered_matches], np.int32)
    t = np.array([m.featureid2[1] for m in ip.filtered_matches], np.int32)
    d = np.array([m.distance for m in ip.filtered_matches], np.float32)
    inl_q = np.array([m.featureid1[1] for m in inliers], np.int32)
    desc = seq.desc.astype(np.uint8)  # SIFT data is integer valued in [0,255]: lossless
    assert np.array_equal(desc.astype(seq.desc.dtype), seq.desc)
    np.savez_compressed(
        os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".npz"),
        descriptor=descriptor, desc=desc, kp=seq.kp, K=K, q=q, t=t, dist=d, inlier_q=inl_q,
        E=ip.essential_matrix, R=ip.R, tvec=ip.t, relative_pose=ip.relative_pose,
        X=ip.points3d_reconstr, projection_matrix=ip.projection_matrix,
        null_projection_matrix=ip.null_projection_matrix, cv2_version=cv2.__version__)
    print(name, "matches", len(q), "inliers", len(inl_q))


if __name__ == "__main__":
    ref = load_reference.load()
    for case in CASES:
        run_case(ref, *case)
