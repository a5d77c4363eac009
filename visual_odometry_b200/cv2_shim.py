"""`cv2` namespace handed to the reference's drivers through `from initials import *`.

src/visual_slam.py calls cv2.solvePnPRansac (:231) and cv2.Rodrigues (:243) directly, so the only
way to move them to the GPU with that file unchanged is to rebind the name `cv2`.
The matcher and the detectors the reference constructs through that name come from this
package too (B200Matcher, OrbDetector, SiftDetector).  Everything not overridden forwards to the
real OpenCV module (image I/O, resize, drawing).
"""
import numpy as np

import cv2 as _real


class _Cv2Namespace:
    def __init__(self):
        self._gui = "GUI:                           NONE" not in _real.getBuildInformation()

    def __getattr__(self, name):
        return getattr(_real, name)

    # ---- GPU overrides (K7) ----
    def solvePnPRansac(self, objectPoints, imagePoints, cameraMatrix, distCoeffs, *a, **k):
        from . import pnp
        return pnp.solve_pnp_ransac(objectPoints, imagePoints, cameraMatrix, distCoeffs, *a, **k)

    def Rodrigues(self, src, *a, **k):
        from . import pnp
        return pnp.rodrigues(src)

    # ---- GPU overrides (K1 / K6 matcher, N4 detector): the objects the reference builds in
    # VisualSlam.__init__ (src/visual_slam.py:16-19) and ImageAndKeypoints (src/image_and_keypoints.py:8-13)
    def BFMatcher(self, normType=None, crossCheck=False):
        from . import matcher
        normType = _real.NORM_L2 if normType is None else normType
        if normType not in (_real.NORM_HAMMING, _real.NORM_L2):
            raise NotImplementedError("B200 path: BFMatcher supports NORM_HAMMING and NORM_L2 only")
        return matcher.B200Matcher(matcher.NORM_HAMMING if normType == _real.NORM_HAMMING
                                   else matcher.NORM_L2, crossCheck=crossCheck)

    def DescriptorMatcher_create(self, matcherType):
        """src/feature_detection.py:20 builds its kNN matcher this way (FLANN, approximate); the
        B200 path answers with the EXACT brute-force matcher north_star specifies for the ratio
        test: FLANNBASED and BRUTEFORCE / _L2 -> NORM_L2, BRUTEFORCE_HAMMING -> NORM_HAMMING."""
        from . import matcher
        names = {"FlannBased": matcher.NORM_L2, "BruteForce": matcher.NORM_L2,
                 "BruteForce-Hamming": matcher.NORM_HAMMING}
        ids = {_real.DESCRIPTOR_MATCHER_FLANNBASED: matcher.NORM_L2,
               _real.DESCRIPTOR_MATCHER_BRUTEFORCE: matcher.NORM_L2,
               _real.DESCRIPTOR_MATCHER_BRUTEFORCE_HAMMING: matcher.NORM_HAMMING}
        norm = names.get(matcherType) if isinstance(matcherType, str) else ids.get(matcherType)
        if norm is None:
            raise NotImplementedError("B200 path: DescriptorMatcher_create supports FLANNBASED (as "
                                      "exact brute force), BRUTEFORCE and BRUTEFORCE_HAMMING")
        return matcher.B200Matcher(norm, crossCheck=False)

    def ORB_create(self, nfeatures=500, scaleFactor=1.2, nlevels=8, edgeThreshold=31, firstLevel=0,
                   WTA_K=2, scoreType=0, patchSize=31, fastThreshold=20):
        if firstLevel != 0 or WTA_K != 2 or scoreType != _real.ORB_HARRIS_SCORE or patchSize != 31:
            raise NotImplementedError("B200 path: ORB with firstLevel 0, WTA_K 2, HARRIS_SCORE, "
                                      "patchSize 31 only")
        from . import orb
        return orb.OrbDetector(nfeatures, scaleFactor, nlevels, edgeThreshold, fastThreshold)

    def SIFT_create(self, nfeatures=0, nOctaveLayers=3, contrastThreshold=0.04, edgeThreshold=10,
                    sigma=1.6, *a, **k):
        """the live detector of src/visual_slam.py:17 / src/image_and_keypoints.py:12"""
        if a or k:
            raise NotImplementedError("B200 path: SIFT_create(nfeatures, nOctaveLayers, contrastThreshold, "
                                      "edgeThreshold, sigma) only")
        from . import sift
        return sift.SiftDetector(nfeatures, nOctaveLayers, contrastThreshold, edgeThreshold, sigma)

    # ---- headless GUI ----
    def imread(self, *a, **k):
        self._idle_polls = 0           # a frame loop is alive (see waitKey)
        return _real.imread(*a, **k)

    def imshow(self, *a, **k):
        self._idle_polls = 0
        return _real.imshow(*a, **k) if self._gui else None

    def waitKey(self, delay=0, *a, **k):
        """Headless: there is no window and no key will ever arrive.  Sleep like cv2 would (capped
        at 10 ms so frame loops are not slowed) and, after `VO_B200_HEADLESS_QUIT` (default 20)
        consecutive polls with no imshow in between, answer ord('q'): the reference's
        `VisualSlam.run` ends with `while True: k = cv2.waitKey(100); if k == ord('q'): break`
        (src/visual_slam.py:344-349), which would otherwise spin forever."""
        if self._gui:
            return _real.waitKey(delay, *a, **k)
        import os
        import time
        time.sleep(min(max(int(delay), 0), 10) * 1e-3)
        self._idle_polls = getattr(self, "_idle_polls", 0) + 1
        limit = int(os.environ.get("VO_B200_HEADLESS_QUIT", "20"))
        if limit > 0 and self._idle_polls >= limit:
            self._idle_polls = 0
            return ord("q")
        return -1

    def destroyAllWindows(self, *a, **k):
        return _real.destroyAllWindows(*a, **k) if self._gui else None


cv2 = _Cv2Namespace()
