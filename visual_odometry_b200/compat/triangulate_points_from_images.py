"""bare-name shim: reference src/main_triangulate.py:2"""
from visual_odometry_b200.legacy import TriangulatePointsFromTwoImages  # noqa: F401
