"""N1: feature tracks and map association kept on the device.

Mirrors the bookkeeping of the reference's VisualSlam (src/visual_slam.py):
  update_feature_mapper (:183-188)          -> FeatureTracks.update(pair results)
  track_feature_back_in_time (:94-99)       -> FeatureTracks.root_of(feature_ids)
  mappointdict rebuild (:154-156, :193-196) -> FeatureTracks.set_map_points(feature_ids, rows)
  MatchWithMap assembly (:202-229)          -> FeatureTracks.matches_with_map(...)
Feature id (frame_id, idx) == flat index frame_id * n_cap + idx.  Kernels: csrc/tracks.cu.
"""
import numpy as np
import torch

from . import _lib

I32, F64 = torch.int32, torch.float64


class FeatureTracks:
    def __init__(self, n_frames, n_cap, device="cuda"):
        self.lib = _lib.load()
        self.dev = torch.device(device)
        self.n_frames, self.n_cap = int(n_frames), int(n_cap)
        n = self.n_frames * self.n_cap
        self.parent = torch.empty(n, dtype=I32, device=self.dev)
        self.root = torch.empty(n, dtype=I32, device=self.dev)
        self.root_point = torch.empty(n, dtype=I32, device=self.dev)
        self.reset()

    def feature_id(self, frame, idx):
        return int(frame) * self.n_cap + int(idx)

    def reset(self):
        _lib.check(self.lib.vo_tracks_reset(_lib.ptr(self.parent), _lib.ptr(self.root),
                                            _lib.ptr(self.root_point), self.parent.numel(),
                                            _lib.cur_stream()))

    def update(self, q, t, mask, n_match, frame1, frame2):
        """feature_mapper[(frame2, t[i])] = (frame1, q[i]) for the masked matches of one pair.
        q, t int32 device tensors, mask uint8 or None, n_match a 1-element int32 device tensor."""
        _lib.check(self.lib.vo_tracks_update(
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(mask), _lib.ptr(n_match), int(frame1), int(frame2),
            self.n_cap, _lib.ptr(self.parent), _lib.ptr(self.root), _lib.cur_stream()))

    def set_map_points(self, feature_ids, point_rows=None):
        """mappointdict[point.feature_id] = point for every map point (row = index into the
        map-point array; default: position in `feature_ids`)."""
        fid = torch.as_tensor(feature_ids, dtype=I32).to(self.dev).contiguous()
        rows = None if point_rows is None else \
            torch.as_tensor(point_rows, dtype=I32).to(self.dev).contiguous()
        _lib.check(self.lib.vo_tracks_set_points(_lib.ptr(fid), _lib.ptr(rows), fid.numel(),
                                                 _lib.ptr(self.root_point), _lib.cur_stream()))

    def root_of(self, feature_ids):
        idx = torch.as_tensor(feature_ids, dtype=torch.int64, device=self.dev)
        return self.root[idx]

    def matches_with_map(self, q, t, mask, n_match, frame1, frame2, kp, map_points):
        """-> (image_coords [k,2] f64, map_coords [k,3] f64, feature_id2 [k], point_row [k]) on the
        device, in match order: the arguments of solvePnPRansac (visual_slam.py:222-235)."""
        n = self.n_cap
        img = torch.empty(n, 2, dtype=F64, device=self.dev)
        xyz = torch.empty(n, 3, dtype=F64, device=self.dev)
        feat2 = torch.empty(n, dtype=I32, device=self.dev)
        prow = torch.empty(n, dtype=I32, device=self.dev)
        n_out = torch.zeros(1, dtype=I32, device=self.dev)
        _lib.check(self.lib.vo_tracks_match_map(
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(mask), _lib.ptr(n_match), int(frame1), int(frame2),
            n, _lib.ptr(self.root), _lib.ptr(self.root_point), _lib.ptr(kp),
            _lib.ptr(map_points), _lib.ptr(img), _lib.ptr(xyz), _lib.ptr(feat2), _lib.ptr(prow),
            _lib.ptr(n_out), _lib.cur_stream()))
        k = int(n_out.item())
        return img[:k], xyz[:k], feat2[:k], prow[:k]
