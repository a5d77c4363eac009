"""Golden vectors for the SIFT front end (N4): outputs of the cv2 binary installed in the build
container (cv2.SIFT_create().detectAndCompute) on visual_odometry_b200.synth.textured_image scenes,
in cv2's own output order.  Run from the repo root:  python tests/golden/make_golden_siftdet.py
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from visual_odometry_b200 import synth  # noqa: E402

CASES = {
    # name: (height, width, seed)
    "siftdet_120x160": (120, 160, 3),
    "siftdet_243x325": (243, 325, 9),        # odd sizes: scalar tails of cv2's blur loops
    "siftdet_480x640": (480, 640, 11),
}


def main():
    for name, (h, w, seed) in CASES.items():
        img = synth.textured_image(h, w, seed)
        kps, desc = cv2.SIFT_create().detectAndCompute(img, None)
        kp = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response) for k in kps], np.float32).reshape(-1, 5)
        octave = np.array([k.octave for k in kps], np.int32)
        assert np.array_equal(desc, np.rint(desc)) and desc.min() >= 0 and desc.max() <= 255
        out = os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".npz")
        np.savez_compressed(out, hws=np.array([h, w, seed]), kp=kp, octave=octave,
                            desc=desc.astype(np.uint8), cv2_version=cv2.__version__)
        print(name, len(kps), os.path.getsize(out))


if __name__ == "__main__":
    main()
