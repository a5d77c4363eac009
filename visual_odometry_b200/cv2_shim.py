"""`cv2` namespace handed to the reference's drivers through `from initials import *`.

src/visual_slam.py calls cv2.solvePnPRansac (:231) and cv2.Rodrigues (:243) directly, so the only
way to move them to the GPU with that file unchanged is to rebind the name `cv2`.
Everything not overridden forwards to the real OpenCV module (image I/O, resize, drawing).
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

    # ---- headless GUI ----
    def imshow(self, *a, **k):
        return _real.imshow(*a, **k) if self._gui else None

    def waitKey(self, *a, **k):
        return _real.waitKey(*a, **k) if self._gui else -1

    def destroyAllWindows(self, *a, **k):
        return _real.destroyAllWindows(*a, **k) if self._gui else None


cv2 = _Cv2Namespace()
