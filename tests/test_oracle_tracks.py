"""CPU: oracle/tracks_restate.py against tests/golden/tracks.npz, which was produced by the
reference's own VisualSlam (tests/golden/make_golden_tracks.py): roots of every feature after each
update and the exact object / image coordinate arrays the reference hands to solvePnPRansac."""
import os

import numpy as np

from oracle.tracks_restate import TracksOracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tracks.npz")


def replay(step_fn):
    g = np.load(GOLD)
    F, N = int(g["F"]), int(g["N"])
    for b in range(F - 1):
        ms = [(tuple(f1), tuple(f2), tuple(kp)) for f1, f2, kp in
              zip(g[f"f1_{b}"].tolist(), g[f"f2_{b}"].tolist(), g[f"kp2_{b}"].tolist())]
        n_map = int(g[f"n_map_{b}"])
        step_fn(b, N, ms, g["map_ids"][:n_map], g["map_xyz"][:n_map], g[f"roots_{b}"],
                g[f"obj_{b}"], g[f"img_{b}"], bool(g[f"called_{b}"]))
    return F


def test_oracle_equals_reference_visual_slam():
    orc = TracksOracle()
    seen = []

    def step(b, N, ms, map_ids, map_xyz, roots, obj, img, called):
        # estimate_current_camera_position: update_feature_mapper, mappointdict, (return if < 3
        # frames), MatchWithMap loop
        orc.update_feature_mapper(ms)
        orc.set_map_points([(tuple(f), x) for f, x in zip(map_ids.tolist(), map_xyz)])
        got_roots = [orc.track_feature_back_in_time((b + 1, i)) for i in range(N)]
        assert got_roots == [tuple(r) for r in roots.tolist()]
        assert called == (b + 2 >= 3)
        if called:
            mm = orc.matches_with_map(ms)
            assert np.array_equal(np.array([m[3] for m in mm]).reshape(-1, 3), obj)
            assert np.array_equal(np.array([m[2] for m in mm]).reshape(-1, 2), img)
            seen.append(len(mm))

    F = replay(step)
    assert len(seen) == F - 2 and min(seen) > 10
