"""visual_odometry_b200 — B200-native two-view visual-odometry hot path.

Drop-in for the hot path of Samirez/Visual_odometry (ImagePair / FrameGenerator /
TriangulatePointsFromTwoImages call surface) on hand-written sm_100a CUDA kernels behind the
C-ABI in include/vo_b200.h.  No CPU fallback: importing the compute entry points without the
built library raises.
"""
__version__ = "0.1.0"
