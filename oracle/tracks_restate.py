"""TEST INFRASTRUCTURE (oracle) — the reference VisualSlam's track bookkeeping with its own data
structures (Python dicts keyed by (frame_id, idx) tuples), restated from src/visual_slam.py:
  update_feature_mapper :183-188, track_feature_back_in_time :94-99,
  mappointdict :154-156 / :193-196, MatchWithMap assembly :202-229.
Pinned: tests/golden/tracks.npz holds the roots and the solvePnPRansac argument arrays produced by
the reference's own VisualSlam on a seeded match sequence (tests/golden/make_golden_tracks.py runs
/root/reference/src/visual_slam.py unmodified); tests/test_oracle_tracks.py checks this file
against it.  Only tests/ may import this."""


class TracksOracle:
    def __init__(self):
        self.feature_mapper = {}
        self.feature_history = {}
        self.mappointdict = {}

    def update_feature_mapper(self, matches):
        """matches: iterable of (featureid1, featureid2, keypoint2) of matches_with_3d_information"""
        for f1, f2, kp2 in matches:
            self.feature_mapper[f2] = f1
            self.feature_history[f1] = (f1, f2, kp2)

    def track_feature_back_in_time(self, featureid):
        while featureid in self.feature_mapper:
            featureid = self.feature_mapper[featureid]
        return featureid

    def set_map_points(self, points):
        """points: iterable of (feature_id, xyz) in map order"""
        self.mappointdict = {}
        for row, (fid, xyz) in enumerate(points):
            self.mappointdict[fid] = (row, xyz)

    def matches_with_map(self, matches):
        out = []
        for f1, f2, kp2 in matches:
            feature_id = f2
            while feature_id in self.feature_mapper:
                feature_id = self.feature_mapper[feature_id]
            if feature_id in self.mappointdict:
                image_feature = self.feature_history[f1]
                row, xyz = self.mappointdict[feature_id]
                out.append((image_feature[1], feature_id, image_feature[2], xyz, row))
        return out
