"""bare-name shim: `from image_pair import ImagePair` (reference src/visual_slam.py:8)"""
from visual_odometry_b200.image_pair import (ImagePair, isRotationMatrix,  # noqa: F401
                                             rotationMatrixToEulerAngles)
