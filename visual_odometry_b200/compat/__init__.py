"""Bare-name modules (`initials`, `image_pair`, `frame`, `frame_generator`, `image_and_keypoints`,
`triangulate_points_from_images`) that shadow the reference's src/ modules of the same name.

Put this directory on sys.path AHEAD of the reference's src/ (see `install()`), and the
reference's own `main.py`, `main_triangulate.py`, `visual_slam.py`, `map.py`, `tracked_*.py`,
`observation.py` run unchanged on the B200 hot path (seam 2).
"""
import os
import sys


def install(reference_src=None, patch_map=False):
    """Prepend the compat directory (and optionally the reference's src/) to sys.path.
    patch_map=True also imports the reference's own `map` module and rebinds `Map.optimize_map`
    (g2o, not installable here) and the reprojection-error loops to the GPU versions
    (visual_odometry_b200.map_ops.patch_map)."""
    here = os.path.dirname(os.path.abspath(__file__))
    if reference_src and reference_src not in sys.path:
        sys.path.insert(0, reference_src)
    if here in sys.path:
        sys.path.remove(here)
    sys.path.insert(0, here)
    if patch_map:
        import importlib
        from visual_odometry_b200.map_ops import patch_map as _patch
        _patch(importlib.import_module("map").Map)
    return here
