"""oracle/load_reference.py — import the REAL reference classes from /root/reference/src.

TEST INFRASTRUCTURE ONLY, and only usable in the build container (the GPU box has no
/root/reference).  src/initials.py:4-12 imports OpenGL, g2o, ThreeDimViewer(pygame), matplotlib,
plotly and codetiming unconditionally; none is installed, so empty stand-ins are seeded into
sys.modules first.  Nothing is written to the read-only reference tree.
"""
import os
import sys
import types

REFERENCE_SRC = "/root/reference/src"


def available():
    return os.path.isdir(REFERENCE_SRC)


def load(extra=()):
    """Returns the reference modules dict: image_pair, frame, frame_generator, initials (+ the
    module names in `extra`, e.g. ("map", "tracked_camera", "tracked_point", "observation"))."""
    if not available():
        raise RuntimeError("reference sources not present on this machine")
    sys.dont_write_bytecode = True

    def stub(name, **attrs):
        if name in sys.modules:
            return sys.modules[name]
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _Timer:
        def __init__(self, *a, **k):
            pass

        def __call__(self, fn):
            return fn

    stub("OpenGL"); stub("OpenGL.GL"); stub("g2o"); stub("ThreeDimViewer")
    stub("pygame"); stub("icecream", ic=print)
    for name in ("matplotlib", "matplotlib.pyplot", "plotly", "plotly.graph_objects"):
        try:
            __import__(name)
        except Exception:
            stub(name)
    try:
        __import__("codetiming")
    except Exception:
        stub("codetiming", Timer=_Timer)
    saved = list(sys.path)
    names = ("initials", "image_pair", "frame", "frame_generator") + tuple(extra)
    shadow = [k for k in names if k in sys.modules]
    kept = {k: sys.modules.pop(k) for k in shadow}
    sys.path.insert(0, REFERENCE_SRC)
    try:
        import initials, frame, frame_generator, image_pair  # noqa: E401
        mods = dict(initials=initials, frame=frame, frame_generator=frame_generator,
                    image_pair=image_pair)
        for name in extra:
            mods[name] = __import__(name)
    finally:
        sys.path[:] = saved
        for k in names:
            sys.modules.pop(k, None)
        sys.modules.update(kept)
    return mods
