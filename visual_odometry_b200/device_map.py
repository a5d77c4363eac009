"""N1 + N2: the reference's map bookkeeping on a flat, device-resident map.

Mirrors, as launches on structure-of-arrays buffers that never leave HBM:
  VisualSlam.initialize_map               (src/visual_slam.py:43-87)   -> DeviceMap.initialize
  VisualSlam.update_feature_mapper        (:183-188)                   -> DeviceMap.update_tracks
  MatchWithMap assembly for solvePnPRansac (:190-235)                  -> DeviceMap.matches_with_map
  map.add_camera(TrackedCamera(R, tvec))  (:244-250)                   -> DeviceMap.add_camera
  VisualSlam.add_information_to_map       (:152-181, :100-150)         -> DeviceMap.extend
  Map.calculate_reprojection_error        (src/map.py:72-97)           -> DeviceMap.reprojection_error
  Map.remove_observations_with_reprojection_errors_above_threshold (:46-70) -> DeviceMap.prune
Feature id (frame_id, idx) == frame_id * n_cap + idx; point / camera ids are array rows.
Kernels: csrc/tracks.cu, csrc/mapops.cu.  No CPU fallback.
"""
import numpy as np
import torch

from . import _lib
from .tracks import FeatureTracks

F64, I32, U8 = torch.float64, torch.int32, torch.uint8


class DeviceMap:
    def __init__(self, K, n_frames, n_cap, cap_points, cap_obs, cap_cams=64, device="cuda"):
        self.lib = _lib.load()
        self.dev = torch.device(device)
        self.n_cap = int(n_cap)
        self.tracks = FeatureTracks(n_frames, n_cap, device=device)
        self.K = torch.as_tensor(np.asarray(K, np.float64).reshape(9)).to(self.dev)
        self.cap_points, self.cap_obs, self.cap_cams = int(cap_points), int(cap_obs), int(cap_cams)
        e = lambda *s, dt: torch.empty(*s, dtype=dt, device=self.dev)
        # point and observation arrays are double-buffered: compactions write the other set
        self._pts = [dict(xyz=e(self.cap_points, 3, dt=F64), fid=e(self.cap_points, dt=I32))
                     for _ in range(2)]
        self._pcur = 0
        self._obs = [dict(xy=e(self.cap_obs, 2, dt=F64), cam=e(self.cap_obs, dt=I32),
                          pt=e(self.cap_obs, dt=I32)) for _ in range(2)]
        self._cur = 0
        self.poses = torch.zeros(self.cap_cams, 16, dtype=F64, device=self.dev)
        self.cam_frame = []                                   # frame id of every camera row
        self.fixed = torch.zeros(self.cap_cams, dtype=U8, device=self.dev)
        self.n_points = torch.zeros(1, dtype=I32, device=self.dev)
        self.n_obs = torch.zeros(1, dtype=I32, device=self.dev)
        self.status = torch.zeros(1, dtype=I32, device=self.dev)
        self.sqerr = e(self.cap_obs, dt=F64)
        self.keep = e(self.cap_obs, dt=U8)
        self._total = torch.zeros(1, dtype=F64, device=self.dev)
        self._n_keep = torch.zeros(1, dtype=I32, device=self.dev)
        self._ws = e(max(self.lib.vo_reprojection_workspace_bytes(self.cap_obs),
                         self.lib.vo_map_prune_workspace_bytes(self.cap_obs)), dt=U8)

    # ------------------------------------------------------------------ views
    @property
    def points(self):
        return self._pts[self._pcur]["xyz"]

    @property
    def point_fid(self):
        return self._pts[self._pcur]["fid"]

    @property
    def obs_xy(self):
        return self._obs[self._cur]["xy"]

    @property
    def obs_cam(self):
        return self._obs[self._cur]["cam"]

    @property
    def obs_pt(self):
        return self._obs[self._cur]["pt"]

    def counts(self):
        """(n_points, n_obs) — one small device->host read; raises if a call overflowed."""
        c = torch.stack([self.n_points[0], self.n_obs[0], self.status[0]]).cpu().tolist()
        if c[2]:
            raise RuntimeError("DeviceMap: cap_points / cap_obs too small for the last extend()")
        return c[0], c[1]

    # ------------------------------------------------------------------ cameras
    def add_camera(self, R, t, frame_id, fixed=False):
        """map.add_camera(TrackedCamera(R, t, frame_id, ..., fixed)): returns the camera row."""
        row = len(self.cam_frame)
        if row >= self.cap_cams:
            raise RuntimeError("DeviceMap: cap_cams exceeded")
        T = np.eye(4)
        T[:3, :3] = np.asarray(R, np.float64).reshape(3, 3)
        T[:3, 3] = np.asarray(t, np.float64).ravel()[:3]
        self.poses[row].copy_(torch.from_numpy(T.reshape(16)))
        self.cam_frame.append(int(frame_id))
        self.fixed[row] = int(bool(fixed))
        return row

    def freeze_nonlast_cameras(self):
        """VisualSlam.freeze_nonlast_cameras (src/visual_slam.py:270-275): only the last camera
        (and the one before it when there are more than two) stays free."""
        n = len(self.cam_frame)
        f = torch.ones(n, dtype=U8)
        f[n - 1] = 0
        if n > 2:
            f[n - 2] = 0
        self.fixed[:n].copy_(f)

    def unfreeze_cameras(self, number_of_fixed_cameras=5):
        """VisualSlam.unfreeze_cameras (src/visual_slam.py:281-286): camera rows above
        `number_of_fixed_cameras` become free, the others fixed."""
        n = len(self.cam_frame)
        self.fixed[:n].copy_((torch.arange(n) <= number_of_fixed_cameras).to(U8))

    def optimize(self, iterations=40, huber_delta=1.0):
        """Map.optimize_map (src/map.py:104-186) on the device arrays, in place."""
        from .ba import bundle_adjust
        n_pts, n_obs = self.counts()
        if n_obs == 0:
            return None
        return bundle_adjust(self.poses, self.fixed, self.points, self.obs_cam, self.obs_pt,
                             self.obs_xy, self.K.reshape(3, 3), iterations, huber_delta,
                             n_cams=len(self.cam_frame), n_points=n_pts, n_obs=n_obs)

    def remove_camera(self, row=0):
        """Map.remove_camera_from_map (src/map.py:188-232) for camera `row`."""
        n_pts, n_obs = self.counts()
        n_cams = len(self.cam_frame)
        wsb = self.lib.vo_map_remove_camera_workspace_bytes(n_pts, n_obs)
        ws = torch.empty(wsb, dtype=U8, device=self.dev)
        ps, pd = self._pts[self._pcur], self._pts[1 - self._pcur]
        os_, od = self._obs[self._cur], self._obs[1 - self._cur]
        _lib.check(self.lib.vo_map_remove_camera(
            int(row), _lib.ptr(self.poses), _lib.ptr(self.fixed), n_cams, _lib.ptr(ps["xyz"]),
            _lib.ptr(ps["fid"]), n_pts, _lib.ptr(pd["xyz"]), _lib.ptr(pd["fid"]),
            _lib.ptr(self.n_points), _lib.ptr(os_["cam"]), _lib.ptr(os_["pt"]), _lib.ptr(os_["xy"]),
            n_obs, _lib.ptr(od["cam"]), _lib.ptr(od["pt"]), _lib.ptr(od["xy"]), _lib.ptr(self.n_obs),
            _lib.ptr(self.tracks.root_point), _lib.ptr(ws), wsb, _lib.cur_stream()))
        self._pcur, self._cur = 1 - self._pcur, 1 - self._cur
        del self.cam_frame[int(row)]

    def limit_number_of_cameras(self, max_cameras=18):
        """Map.limit_number_of_camera_in_map (src/map.py:301-318): drop the oldest camera when
        there are more than `max_cameras`."""
        if len(self.cam_frame) > max_cameras:
            self.remove_camera(0)

    def camera_row(self, frame_id):
        return self.cam_frame.index(int(frame_id))

    # ------------------------------------------------------------------ tracks
    def update_tracks(self, q, t, mask, n_match, frame1, frame2):
        self.tracks.update(q, t, mask, n_match, frame1, frame2)

    def matches_with_map(self, q, t, mask, n_match, frame1, frame2, kp):
        return self.tracks.matches_with_map(q, t, mask, n_match, frame1, frame2, kp, self.points)

    # ------------------------------------------------------------------ growth
    def extend(self, q, t, mask, n_match, frame1, frame2, kp, X, max_norm=50.0, publish=True):
        """add_information_to_map's loop for one pair.  X [k, 3 or 4] f64: triangulated points of
        the masked matches in match order (row r = r-th masked match)."""
        X = X if X.is_contiguous() else X.contiguous()
        _lib.check(self.lib.vo_map_extend(
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(mask), _lib.ptr(n_match), int(frame1), int(frame2),
            self.n_cap, _lib.ptr(self.tracks.root), _lib.ptr(self.tracks.root_point), _lib.ptr(kp),
            _lib.ptr(X), int(X.shape[1]), self.camera_row(frame1), self.camera_row(frame2),
            float(max_norm), _lib.ptr(self.points), _lib.ptr(self.point_fid),
            _lib.ptr(self.n_points), self.cap_points, _lib.ptr(self.obs_xy), _lib.ptr(self.obs_cam),
            _lib.ptr(self.obs_pt), _lib.ptr(self.n_obs), self.cap_obs, int(bool(publish)),
            _lib.ptr(self.status), _lib.cur_stream()))

    def initialize(self, q, t, mask, n_match, frame1, frame2, kp, X, R, t_vec):
        """initialize_map: cameras [I|0] and [R|t], one point + two observations per match (no
        distance filter), feature_mapper links of the pair."""
        self.n_points.zero_(); self.n_obs.zero_(); self.status.zero_()
        self.cam_frame = []
        self.tracks.reset()
        self.update_tracks(q, t, mask, n_match, frame1, frame2)
        self.add_camera(np.eye(3), np.zeros(3), frame1, fixed=True)
        self.add_camera(R, t_vec, frame2, fixed=False)
        self.extend(q, t, mask, n_match, frame1, frame2, kp, X, max_norm=float("inf"))

    # ------------------------------------------------------------------ reprojection
    def _errors(self, threshold, n_obs):
        n_pts = self.cap_points      # row indices are validated against the allocated arrays
        _lib.check(self.lib.vo_reprojection_errors(
            _lib.ptr(self.points), n_pts, _lib.ptr(self.poses), len(self.cam_frame),
            _lib.ptr(self.obs_cam), _lib.ptr(self.obs_pt), _lib.ptr(self.obs_xy), n_obs,
            _lib.ptr(self.K), float(threshold), _lib.ptr(self.sqerr), _lib.ptr(self.keep),
            _lib.ptr(self._total), _lib.ptr(self._n_keep), _lib.ptr(self._ws), self._ws.numel(),
            _lib.cur_stream()))

    def reprojection_error(self, threshold=10.0):
        """Map.calculate_reprojection_error: total squared error over all observations."""
        _, n_obs = self.counts()
        self._errors(threshold, n_obs)
        return float(self._total.item())

    def prune(self, threshold=100.0):
        """Map.remove_observations_with_reprojection_errors_above_threshold; returns n kept."""
        _, n_obs = self.counts()
        self._errors(threshold, n_obs)
        src, dst = self._obs[self._cur], self._obs[1 - self._cur]
        _lib.check(self.lib.vo_map_prune_observations(
            _lib.ptr(self.keep), n_obs, _lib.ptr(src["cam"]), _lib.ptr(src["pt"]),
            _lib.ptr(src["xy"]), _lib.ptr(dst["cam"]), _lib.ptr(dst["pt"]), _lib.ptr(dst["xy"]),
            _lib.ptr(self.n_obs), _lib.ptr(self._ws), self._ws.numel(), _lib.cur_stream()))
        self._cur = 1 - self._cur
        return int(self.n_obs.item())
