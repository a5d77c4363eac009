"""bare-name shim for the legacy two-image demo (reference src/image_and_keypoints.py)"""
from visual_odometry_b200.legacy import ImageAndKeypoints  # noqa: F401
