"""N2: the reference Map's reprojection-error loops as one kernel launch.

Reference: src/map.py:32-97 — `calculate_reprojection_error(threshold=10)`,
`remove_observations_with_reprojection_errors_above_threshold(threshold=100)`,
`calculate_reprojection_error_for_point(point, camera_pose, image_coord)`; called at
src/visual_slam.py:181,266 and src/map.py:239-241.

`flatten_map(map)` turns the reference's object lists (cameras / points / observations with ids)
into the structure-of-arrays the kernel reads; `patch_map(MapClass)` rebinds the two methods of
the reference's own `Map` class so `visual_slam.py` keeps calling them unchanged.  All arithmetic
runs in libvo_b200 (csrc/mapops.cu); there is no CPU fallback.
"""
import numpy as np
import torch

from . import _lib

F64, I32, U8 = torch.float64, torch.int32, torch.uint8


def reprojection_errors(points, poses, obs_cam, obs_pt, obs_xy, K, threshold, device="cuda"):
    """Arrays in (numpy or torch) -> dict(sqerr [n_obs] f64, keep [n_obs] bool, total float,
    n_keep int), all computed on the GPU.  obs_cam / obs_pt are ROW indices."""
    lib = _lib.load()
    dev = torch.device(device)
    t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a) if not isinstance(a, torch.Tensor)
                                      else a, dtype=dt).to(dev).contiguous()
    points, poses = t(points, F64).reshape(-1, 3), t(poses, F64).reshape(-1, 16)
    obs_cam, obs_pt = t(obs_cam, I32).reshape(-1), t(obs_pt, I32).reshape(-1)
    obs_xy, K = t(obs_xy, F64).reshape(-1, 2), t(K, F64).reshape(9)
    n_obs = obs_cam.numel()
    if obs_pt.numel() != n_obs or obs_xy.shape[0] != n_obs:
        raise ValueError("observation arrays differ in length")
    sqerr = torch.empty(n_obs, dtype=F64, device=dev)
    keep = torch.empty(n_obs, dtype=U8, device=dev)
    total = torch.zeros(1, dtype=F64, device=dev)
    n_keep = torch.zeros(1, dtype=I32, device=dev)
    wsb = lib.vo_reprojection_workspace_bytes(n_obs)
    ws = torch.empty(wsb, dtype=U8, device=dev)
    _lib.check(lib.vo_reprojection_errors(
        _lib.ptr(points), points.shape[0], _lib.ptr(poses), poses.shape[0], _lib.ptr(obs_cam),
        _lib.ptr(obs_pt), _lib.ptr(obs_xy), n_obs, _lib.ptr(K), float(threshold),
        _lib.ptr(sqerr), _lib.ptr(keep), _lib.ptr(total), _lib.ptr(n_keep), _lib.ptr(ws), wsb,
        _lib.cur_stream()))
    return dict(sqerr=sqerr, keep=keep.bool(), total=float(total.item()),
                n_keep=int(n_keep.item()))


def flatten_map(m):
    """Reference Map object -> (points [P,3], poses [C,16], obs_cam, obs_pt, obs_xy, K) with ids
    replaced by row indices (the dict lookups of map.py:47-52 / 73-78 done once, vectorised)."""
    cam_row = {c.camera_id: i for i, c in enumerate(m.cameras)}
    pt_row = {p.point_id: i for i, p in enumerate(m.points)}
    points = np.array([np.asarray(p.point, np.float64).ravel()[:3] for p in m.points],
                      np.float64).reshape(-1, 3)
    poses = np.array([c.pose() for c in m.cameras], np.float64).reshape(-1, 16)
    obs_cam = np.fromiter((cam_row[o.camera_id] for o in m.observations), np.int32,
                          len(m.observations))
    obs_pt = np.fromiter((pt_row[o.point_id] for o in m.observations), np.int32,
                         len(m.observations))
    obs_xy = np.array([np.asarray(o.image_coordinates, np.float64).ravel()[:2]
                       for o in m.observations], np.float64).reshape(-1, 2)
    return points, poses, obs_cam, obs_pt, obs_xy, np.asarray(m.camera_matrix, np.float64)


def calculate_reprojection_error(self, threshold=10):
    """Drop-in for Map.calculate_reprojection_error (map.py:72-97): returns the total squared
    error (a 1-element array like the reference, whose dx/dy are 1-element arrays) and prints
    the observations above `threshold` as the reference does."""
    if len(self.observations) == 0:
        return 0        # like the reference's loop over no observations (camera_matrix may be None:
        #                 VisualSlam.run prints the statistics of the still empty map after frame 0)
    out = reprojection_errors(*flatten_map(self), threshold)
    sq = out["sqerr"].cpu().numpy()
    for i in np.nonzero(sq > threshold)[0]:
        print("high reprojection error: %f" % sq[i])
        print(self.observations[i])
    return np.array([out["total"]])


def remove_observations_with_reprojection_errors_above_threshold(self, threshold=100):
    """Drop-in for map.py:46-70: keeps observations with sqerr < threshold, in order."""
    if len(self.observations) == 0:
        return
    out = reprojection_errors(*flatten_map(self), threshold)
    keep = out["keep"].cpu().numpy()
    self.observations = [o for o, k in zip(self.observations, keep) if k]


def calculate_reprojection_error_for_point(self, point, camera_pose, image_coord):
    """Drop-in for map.py:32-44 (single observation; `point` homogeneous [4])."""
    p = np.asarray(point, np.float64).ravel()
    out = reprojection_errors(p[:3][None] / (p[3] if p.size > 3 else 1.0),
                              np.asarray(camera_pose, np.float64).reshape(1, 16), [0], [0],
                              np.asarray(image_coord, np.float64).ravel()[:2][None],
                              self.camera_matrix, np.inf)
    return float(out["sqerr"][0])


def patch_map(MapClass, bundle_adjustment=True):
    """Rebind the reprojection methods (and optimize_map) of the reference's Map class to the GPU
    versions."""
    if bundle_adjustment:
        from .ba import optimize_map
        MapClass.optimize_map = optimize_map
    MapClass.calculate_reprojection_error = calculate_reprojection_error
    MapClass.remove_observations_with_reprojection_errors_above_threshold = \
        remove_observations_with_reprojection_errors_above_threshold
    return MapClass
