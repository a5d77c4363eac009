"""CPU: the C-ABI library loads and exports every symbol include/vo_b200.h declares
(no compute calls — there is no GPU here)."""
import ctypes
import os
import re

from visual_odometry_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "vo_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vo_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build with __graft_entry__.build()"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), name


def test_loader_binds_and_reports_version():
    lib = _lib.load()
    assert lib.vo_version() >= 100
    assert isinstance(lib.vo_last_error_string(), bytes)
    # pure host-side size queries work without a device
    assert lib.vo_match_workspace_bytes(0, 3, 100) >= 3 * 100 * 33 * 4
    assert lib.vo_essential_workspace_bytes(2, 100, 1000) > 0
    assert lib.vo_pose_workspace_bytes(2, 100) >= 200


def test_invalid_arguments_fail_loudly_without_a_device():
    lib = _lib.load()
    rc = lib.vo_match_hamming_batch(None, None, None, 10, 2, None, 1, 0, 0.0, None, None, None, None,
                                    None, None, None, None, 0, None)
    assert rc == _lib.VO_ERR_INVALID_ARG
    assert b"null" in lib.vo_last_error_string()
    rc = lib.vo_ransac_samples(3, 10, ctypes.c_void_p(8), None)
    assert rc == _lib.VO_ERR_TOO_FEW_POINTS


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "visual_odometry_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
