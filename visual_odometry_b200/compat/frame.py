"""bare-name shim: `from frame import Frame` (reference src/visual_slam.py:3)"""
from visual_odometry_b200.frame import Frame  # noqa: F401
