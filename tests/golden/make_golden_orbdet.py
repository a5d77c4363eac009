"""Golden vectors for the ORB front end (N4): outputs of the cv2 binary installed in the build
container (cv2.ORB_create(...).detectAndCompute) on visual_odometry_b200.synth.textured_image
scenes, sorted by (octave, y, x) because cv2's order inside an octave is what std::nth_element
leaves.  Run from the repo root:  python tests/golden/make_golden_orbdet.py
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from visual_odometry_b200 import synth  # noqa: E402

CASES = {
    # name: (height, width, seed, nfeatures, nlevels, edgeThreshold, fastThreshold)
    "orbdet_480x640_500": (480, 640, 11, 500, 8, 31, 20),
    "orbdet_1080x1920_2000": (1080, 1920, 12, 2000, 8, 31, 20),
    "orbdet_240x320_all_1level": (240, 320, 13, 100000, 1, 31, 20),
    "orbdet_150x200_300_thr10": (150, 200, 14, 300, 4, 31, 10),
}


def canonical(kps, desc):
    kp = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps],
                  np.float32).reshape(-1, 6)
    desc = np.zeros((0, 32), np.uint8) if desc is None else desc
    # level coordinates are integers: pt / scale; sort on (octave, y, x) of the scaled points
    o = np.lexsort((kp[:, 0], kp[:, 1], kp[:, 5]))
    return kp[o], desc[o]


def main():
    for name, (h, w, seed, nf, nl, edge, thr) in CASES.items():
        img = synth.textured_image(h, w, seed)
        orb = cv2.ORB_create(nfeatures=nf, nlevels=nl, edgeThreshold=edge, fastThreshold=thr)
        kps, desc = orb.detectAndCompute(img, None)
        kp, desc = canonical(kps, desc)
        fast = cv2.FastFeatureDetector_create(thr, True).detect(img)
        fast = np.array([(k.pt[0], k.pt[1], k.response) for k in fast], np.float32).reshape(-1, 3)
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"),
                            params=np.array([h, w, seed, nf, nl, edge, thr]), kp=kp, desc=desc,
                            fast=fast, image_crc=np.array([int(img.astype(np.uint64).sum())]),
                            cv2_version=np.array(cv2.__version__))
        print(name, len(kp), "key points,", len(fast), "FAST corners on level 0")


if __name__ == "__main__":
    main()
