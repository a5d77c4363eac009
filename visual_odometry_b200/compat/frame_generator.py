"""bare-name shim: `from frame_generator import FrameGenerator` (reference src/visual_slam.py:4)"""
from visual_odometry_b200.frame import FrameGenerator  # noqa: F401
