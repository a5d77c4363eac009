"""Per-pair wall-clock latency of the ImagePair DROP-IN (the call sequence of
src/visual_slam.py:294-298) vs the reference's cv2 chain on the host, one pair at a time —
the way the reference's VisualSlam actually calls the path.  Python-side Match-tuple building is
included on our side (it is part of the surface) and excluded on the cv2 side (pair_cv2 returns
arrays), so the ratio is a lower bound."""
import sys, os, io, time, json, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, cv2, torch
from oracle import reference_path as RP
from visual_odometry_b200 import image_pair as IP
from visual_odometry_b200.frame import FrameGenerator, SyntheticDetector
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING, NORM_L2
from visual_odometry_b200.synth import make_sequence

for desc, N, norm, cvnorm in [("orb", 2000, NORM_HAMMING, cv2.NORM_HAMMING), ("sift", 2000, NORM_L2, cv2.NORM_L2),
                              ("sift", 5000, NORM_L2, cv2.NORM_L2)]:
    seq = make_sequence(6, N, 11, descriptor=desc)
    fg = FrameGenerator(SyntheticDetector(seq))
    frames = [fg.make_frame(np.zeros((4, 4, 3), np.uint8)) for _ in range(6)]
    matcher = B200Matcher(norm, crossCheck=True)
    stages = {"match_features": 0.0, "determine_essential_matrix": 0.0, "estimate_camera_movement": 0.0,
              "reconstruct_3d_points": 0.0}
    def run(b, acc=None):
        ip = IP.ImagePair(frames[b], frames[b + 1], matcher, seq.K)
        with contextlib.redirect_stdout(io.StringIO()):
            t0 = time.perf_counter(); ip.match_features(); t1 = time.perf_counter()
            ess = ip.determine_essential_matrix(ip.filtered_matches); t2 = time.perf_counter()
            ip.estimate_camera_movement(ess); t3 = time.perf_counter()
            ip.reconstruct_3d_points(ess); t4 = time.perf_counter()
        if acc is not None:
            for k, v in zip(acc, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)): acc[k] += v
    for b in range(5): run(b)
    reps = 0
    t0 = time.perf_counter()
    for _ in range(3):
        for b in range(5): run(b, stages); reps += 1
    ours = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for b in range(5): RP.pair_cv2(seq.desc[b], seq.kp[b], seq.desc[b + 1], seq.kp[b + 1], seq.K, cvnorm)
    ref = (time.perf_counter() - t0) / 5
    print(json.dumps({"descriptor": desc, "keypoints": N, "dropin_ms_per_pair": ours * 1e3,
                      "stages_ms": {k: v / reps * 1e3 for k, v in stages.items()},
                      "cv2_chain_ms_per_pair(all host threads)": ref * 1e3, "speedup": ref / ours}), flush=True)
