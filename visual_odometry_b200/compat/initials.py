"""Drop-in for the reference's star-import hub (src/initials.py:1-33).

Exports the same names.  Differences, all outside the hot path:
  * `cv2` is a thin namespace: `solvePnPRansac` and `Rodrigues` (called directly by
    src/visual_slam.py:231,243) go to libvo_b200, `imshow`/`waitKey`/`destroyAllWindows` are
    no-ops when OpenCV was built without a GUI, everything else forwards to the real cv2;
  * GUI / BA / plotting dependencies that are not installed (OpenGL, g2o, pygame viewer,
    matplotlib, plotly, codetiming) are replaced by inert stand-ins — graphics are unavailable on
    this box; bundle adjustment runs through `map_ops.patch_map(Map)` -> vo_bundle_adjust instead of
    g2o (which is not installable here).
"""
import collections
import glob
import importlib
import math
import sys
import time
import types

import numpy as np

from visual_odometry_b200.initials_types import Feature, Match, Match3D, MatchWithMap
from visual_odometry_b200.cv2_shim import cv2

np.set_printoptions(precision=4, suppress=True)


class _Inert(types.ModuleType):
    """module stand-in: any attribute is a callable/attribute sink that does nothing"""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        sink = _Inert(f"{self.__name__}.{name}")
        setattr(self, name, sink)
        return sink

    def __call__(self, *a, **k):
        return _Inert(self.__name__ + "()")

    def __iter__(self):
        return iter(())

    def __bool__(self):
        return False


def _optional(name):
    """import `name` if present, else register and return an inert stand-in"""
    try:
        return importlib.import_module(name)
    except Exception:
        parts = name.split(".")
        for k in range(1, len(parts) + 1):
            sys.modules.setdefault(".".join(parts[:k]), _Inert(".".join(parts[:k])))
        return sys.modules[name]


gl = _optional("OpenGL.GL")
g2o = _optional("g2o")
plt = _optional("matplotlib.pyplot")
go = _optional("plotly.graph_objects")

try:
    from codetiming import Timer
except Exception:
    class Timer:
        """codetiming.Timer look-alike: decorator/context manager printing `text.format(s)`."""

        def __init__(self, text="Elapsed time: {:0.4f} seconds", **kw):
            self.text = text

        def __call__(self, fn):
            def wrapped(*a, **k):
                t0 = time.perf_counter()
                try:
                    return fn(*a, **k)
                finally:
                    print(self.text.format(time.perf_counter() - t0))
            wrapped.__name__ = getattr(fn, "__name__", "timed")
            return wrapped

        def __enter__(self):
            self._t0 = time.perf_counter()
            return self

        def __exit__(self, *exc):
            print(self.text.format(time.perf_counter() - self._t0))

try:
    import ThreeDimViewer  # the reference's PyOpenGL/pygame viewer, if importable
except Exception:
    ThreeDimViewer = _Inert("ThreeDimViewer")

    class _HeadlessViewer:
        """stand-in for ThreeDimViewer.ThreeDimViewer: collects what the caller hands it"""

        def __init__(self):
            self.vertices, self.colors, self.cameras = [], [], []

        def main(self):
            return None

    ThreeDimViewer.ThreeDimViewer = _HeadlessViewer
    sys.modules.setdefault("ThreeDimViewer", ThreeDimViewer)
