"""N3: local bundle adjustment on the GPU — the reference's Map.optimize_map
(src/map.py:104-186: g2o Levenberg + Schur complement + EdgeProjectXYZ2UV + Huber, 40 iterations).

`bundle_adjust(...)` works on flat arrays (what DeviceMap holds); `optimize_map(self)` is the
drop-in for the reference's own Map class (flatten -> one C-ABI call -> write R, t and points back,
src/map.py:175-186).  All arithmetic runs in libvo_b200 (csrc/ba.cu); there is no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

F64, I32, U8 = torch.float64, torch.int32, torch.uint8


def bundle_adjust(poses, fixed, points, obs_cam, obs_pt, obs_xy, K, iterations=40, huber_delta=1.0,
                  n_cams=None, n_points=None, n_obs=None, ws=None):
    """In-place LM bundle adjustment of device tensors `poses` [C,16] f64 and `points` [P,3] f64.
    fixed uint8 [C]; obs_cam / obs_pt int32 [M] (row indices); obs_xy f64 [M,2]; K 3x3 (focal =
    K[0,0], as the reference passes it to g2o.CameraParameters).  Returns a dict with the LM log."""
    lib = _lib.load()
    n_cams = poses.shape[0] if n_cams is None else int(n_cams)
    n_points = points.shape[0] if n_points is None else int(n_points)
    n_obs = obs_cam.shape[0] if n_obs is None else int(n_obs)
    for t, dt in ((poses, F64), (points, F64), (obs_xy, F64), (obs_cam, I32), (obs_pt, I32), (fixed, U8)):
        if t.dtype != dt or not t.is_cuda or not t.is_contiguous():
            raise ValueError("bundle_adjust: contiguous CUDA tensors of the documented dtypes required")
    Kh = np.asarray(K.cpu() if isinstance(K, torch.Tensor) else K, np.float64).reshape(3, 3)
    wsb = lib.vo_bundle_adjust_workspace_bytes(n_cams, n_points, n_obs)
    if ws is None or ws.numel() < wsb:
        ws = torch.empty(wsb, dtype=U8, device=poses.device)
    log_chi = np.zeros(max(iterations, 1), np.float64)
    log_lam = np.zeros(max(iterations, 1), np.float64)
    log_tr = np.zeros(max(iterations, 1), np.int32)
    n_it, chi0 = C.c_int(0), C.c_double(0.0)
    _lib.check(lib.vo_bundle_adjust(
        _lib.ptr(poses), _lib.ptr(fixed), n_cams, _lib.ptr(points), n_points, _lib.ptr(obs_cam),
        _lib.ptr(obs_pt), _lib.ptr(obs_xy), n_obs, float(Kh[0, 0]), float(Kh[0, 2]), float(Kh[1, 2]),
        float(huber_delta), int(iterations), log_chi.ctypes.data_as(C.c_void_p),
        log_lam.ctypes.data_as(C.c_void_p), log_tr.ctypes.data_as(C.c_void_p), C.byref(n_it),
        C.byref(chi0), _lib.ptr(ws), ws.numel(), _lib.cur_stream()))
    k = n_it.value
    return dict(iterations=k, chi2_initial=chi0.value, chi2=log_chi[:k].copy(),
                lam=log_lam[:k].copy(), trials=log_tr[:k].copy())


def optimize_map(self, postfix=""):
    """Drop-in for Map.optimize_map (src/map.py:104-186)."""
    from .map_ops import flatten_map
    points, poses, obs_cam, obs_pt, obs_xy, K = flatten_map(self)
    if len(obs_cam) == 0:
        return None
    dev = torch.device("cuda")
    poses_d = torch.from_numpy(poses).to(dev)
    points_d = torch.from_numpy(points).to(dev)
    fixed = torch.tensor([bool(c.fixed) for c in self.cameras], dtype=U8, device=dev)
    out = bundle_adjust(poses_d, fixed, points_d, torch.from_numpy(obs_cam).to(dev),
                        torch.from_numpy(obs_pt).to(dev), torch.from_numpy(obs_xy).to(dev), K)
    P, X = poses_d.cpu().numpy().reshape(-1, 4, 4), points_d.cpu().numpy()
    for i, cam in enumerate(self.cameras):            # src/map.py:175-179
        cam.t = P[i, :3, 3].copy()
        cam.R = P[i, :3, :3].copy()
    for i, pt in enumerate(self.points):              # src/map.py:181-186
        pt.point = X[i].copy()
    return out
