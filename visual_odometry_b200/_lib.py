"""ctypes binding of csrc/libvo_b200.so (the C-ABI declared in include/vo_b200.h).

There is deliberately no CPU fallback: if the library is missing or a call fails, this module
raises.  PyTorch is used only for device memory and streams; signatures carry raw pointers.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libvo_b200.so")

VO_OK = 0
VO_ERR_INVALID_ARG = -1
VO_ERR_TOO_FEW_POINTS = -2
VO_ERR_CUDA = -3
VO_ERR_WORKSPACE = -4
MATCH_MUTUAL = 0
MATCH_KNN2_RATIO = 1


class VoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libvo_b200 error {code}: {msg}")
        self.code = code


_p = C.c_void_p
_i = C.c_int
_f = C.c_float
_d = C.c_double
_sz = C.c_size_t

# name -> (restype, argtypes); every symbol of include/vo_b200.h
SIGNATURES = {
    "vo_version": (_i, []),
    "vo_last_error_string": (C.c_char_p, []),
    "vo_match_workspace_bytes": (_sz, [_i, _i, _i]),
    "vo_match_hamming_batch": (_i, [_p, _p, _p, _i, _i, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p,
                                    _p, _p, _sz, _p]),
    "vo_debug_hamming": (_i, [_i]),
    "vo_debug_hamming_f4_acc": (_i, [_p, _p, _i, _p, _p, _p, _sz, _p]),
    "vo_match_hamming": (_i, [_p, _i, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_match_hamming_static_max_keypoints": (_i, []),
    "vo_match_hamming_static": (_i, [_p, _i, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p, _p]),
    "vo_match_l2_workspace_bytes": (_sz, [_i, _i, _i]),
    "vo_match_l2_batch": (_i, [_p, _p, _p, _i, _i, _i, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p, _p,
                               _p, _sz, _p]),
    "vo_essential_workspace_bytes": (_sz, [_i, _i, _i]),
    "vo_essential_ransac_batch": (_i, [_p, _p, _i, _i, _p, _d, _d, _i, _i, _p, _p, _p, _p, _p,
                                       _p, _sz, _p]),
    "vo_ransac_round_schedule": (_i, [_i, _i, _i, _p, _i]),
    "vo_fivept_solve": (_i, [_p, _i, _p, _i, _p, _p, _p]),
    "vo_sampson_score": (_i, [_p, _i, _p, _i, _f, _p, _p, _p]),
    "vo_ransac_samples": (_i, [_i, _i, _p, _p]),
    "vo_pose_workspace_bytes": (_sz, [_i, _i]),
    "vo_recover_pose_batch": (_i, [_p, _p, _p, _p, _i, _i, _p, _d, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_triangulate": (_i, [_p, _p, _p, _p, _i, _p, _i, _p]),
    "vo_triangulate_pairs": (_i, [_p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p]),
    "vo_chain_poses": (_i, [_p, _p, _i, _p, _p, _p]),
    "vo_pack_poses": (_i, [_p, _p, _i, _p, _p]),
    "vo_chain_poses_packed": (_i, [_p, _i, _p, _p, _p]),
    "vo_profile_enable": (_i, [_i]),
    "vo_profile_read": (_i, [_i, _p, _p]),
    "vo_measure_popc_peak": (_i, [_p, _p]),
    "vo_sift_octaves": (_i, [_i, _i, _p, _p]),
    "vo_sift_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "vo_sift_detect_batch": (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_host_alloc": (_i, [_p, _sz, _i]),
    "vo_host_free": (_i, [_p]),
    "vo_pnp_workspace_bytes": (_sz, [_i, _i]),
    "vo_pnp_ransac": (_i, [_p, _p, _i, _p, _i, _d, _d, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_pnp_batch_workspace_bytes": (_sz, [C.c_long, _i, _i]),
    "vo_pnp_ransac_batch": (_i, [_p, _p, _p, C.c_long, _i, _p, _i, _d, _d, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_tracks_reset": (_i, [_p, _p, _p, C.c_long, _p]),
    "vo_tracks_update": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p]),
    "vo_tracks_set_points": (_i, [_p, _p, _i, _p, _p]),
    "vo_tracks_match_map": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p,
                                 _p]),
    "vo_reprojection_workspace_bytes": (_sz, [_i]),
    "vo_reprojection_errors": (_i, [_p, _i, _p, _i, _p, _p, _p, _i, _p, _d, _p, _p, _p, _p, _p,
                                    _sz, _p]),
    "vo_map_extend": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _i, _i, _i, _d, _p, _p, _p,
                           _i, _p, _p, _p, _p, _i, _i, _p, _p]),
    "vo_map_remove_camera_workspace_bytes": (_sz, [_i, _i]),
    "vo_map_remove_camera": (_i, [_i, _p, _p, _i, _p, _p, _i, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p,
                                  _p, _p, _sz, _p]),
    "vo_bundle_adjust_workspace_bytes": (_sz, [_i, _i, _i]),
    "vo_bundle_adjust": (_i, [_p, _p, _i, _p, _i, _p, _p, _p, _i, _d, _d, _d, _d, _i, _p, _p, _p, _p,
                              _p, _p, _sz, _p]),
    "vo_orb_workspace_bytes": (_sz, [_i, _i, _i, _i, _f, _i, _i]),
    "vo_orb_level_info": (_i, [_i, _i, _i, _f, _i, _p, _p, _p, _p]),
    "vo_orb_detect_batch": (_i, [_p, _i, _i, _i, _i, _f, _i, _i, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "vo_orb_fast_score_map": (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "vo_map_prune_workspace_bytes": (_sz, [_i]),
    "vo_map_prune_observations": (_i, [_p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
}

_lib = None


def load():
    """Load the shared library (once) and bind every declared symbol; raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code):
    if code != VO_OK:
        raise VoError(code, load().vo_last_error_string().decode())


def ptr(t):
    """data_ptr of a torch tensor (or None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def cur_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
