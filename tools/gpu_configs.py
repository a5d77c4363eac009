"""GPU timing of the BASELINE.json configurations that are not the bench line (C1, C3, C4, C5):
device-resident pairs/s (CUDA events, 3 warm-up + 5 timed passes) for documentation in DESIGN.md.
Parity for these sizes lives in tests/ (test_gpu_match_l2.py::test_c3_full_size_pipeline,
test_gpu_pose.py::test_full_size_c4_batch_vs_reference_path, test_gpu_config_c5.py)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200 import _lib, geometry as G, pnp as PNP
from visual_odometry_b200.pipeline import PairPipeline
from visual_odometry_b200.synth import make_sequence


def timed(fn, warm=3, reps=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def pipeline_case(name, frames, n_kp, tile, **kw):
    """`frames` distinct synthetic frames, tiled to `tile` pairs (descriptor content repeats; the
    work per pair is identical)."""
    seq = make_sequence(frames, n_kp, 20261104 + 7, descriptor=kw.get("descriptor", "orb"))
    F = frames
    B = tile
    pipe = PairPipeline(n_kp, B, F, K=seq.K, **kw)
    pairs = np.stack([np.arange(B) % (F - 1), np.arange(B) % (F - 1) + 1], 1).astype(np.int32)
    pipe.stage_host(seq.desc, seq.kp, seq.nkp, pairs)
    pipe.run_staged(F, B)
    ms = timed(lambda: pipe.run_device(B))
    print(json.dumps({"config": name, "pairs": B, "keypoints": n_kp, "ms_per_pass": ms,
                      "pairs_per_s": B / (ms * 1e-3), "mean_matches": float(pipe.n_match[:B].float().mean()),
                      "mean_inliers": float(pipe.n_inl[:B].float().mean()),
                      "mean_iters": float(pipe.n_iters[:B].float().mean())}), flush=True)


pipeline_case("C1 single pair ORB 1000", 2, 1000, 1)
pipeline_case("C3 SIFT 5000 knn+ratio, 4096 fixed hypotheses, 1 pair", 2, 5000, 1, descriptor="sift",
              max_iters=4096, adaptive=False, match_mode=_lib.MATCH_KNN2_RATIO, ratio=0.75)
pipeline_case("C3 same, 64 pairs", 9, 5000, 64, descriptor="sift",
              max_iters=4096, adaptive=False, match_mode=_lib.MATCH_KNN2_RATIO, ratio=0.75)
pipeline_case("C4 ORB 4000 x 1024 pairs (1/4 of the 4096; per-GPU share at 4 GPUs)", 33, 4000, 1024)

# C5: 100 k points triangulated + PnP of a new frame against the map
rng = np.random.default_rng(5)
n = 100_000
K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
X = np.c_[rng.uniform(-12, 12, n), rng.uniform(-7, 7, n), rng.uniform(8, 40, n)]
import cv2
R = cv2.Rodrigues(np.array([0.02, -0.03, 0.01]))[0]; t = np.array([0.6, 0.05, 0.1])
proj = lambda R, t: (lambda x: x[:, :2] / x[:, 2:])((K @ (R @ X.T + t[:, None])).T)
tt = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float64)).cuda()
P1, P2 = tt(K @ np.c_[np.eye(3), np.zeros(3)]), tt(K @ np.c_[R, t])
x1, x2 = tt(proj(np.eye(3), np.zeros(3))), tt(proj(R, t))
ms = timed(lambda: G.triangulate(P1, P2, x1, x2, dehomog=True))
print(json.dumps({"config": "C5 triangulate 100k points", "ms": ms, "points_per_s": n / (ms * 1e-3),
                  "GB_per_s_algorithmic(64B/pt)": 64 * n / (ms * 1e-3) / 1e9}), flush=True)
for n_obs in (500, 5000):
    sel = rng.permutation(n)[:n_obs]
    img = proj(R, t)[sel] + rng.normal(0, 0.5, (n_obs, 2))
    for _ in range(3): PNP.solve_pnp_ransac(X[sel], img, K, np.zeros(4))
    t0 = time.perf_counter()
    for _ in range(20): PNP.solve_pnp_ransac(X[sel], img, K, np.zeros(4))
    ours = (time.perf_counter() - t0) / 20
    t0 = time.perf_counter()
    for _ in range(20): cv2.solvePnPRansac(X[sel], img, K, np.zeros(4))
    ref = (time.perf_counter() - t0) / 20
    print(json.dumps({"config": f"C5 PnP-RANSAC {n_obs} observations (host call, wall clock)",
                      "ours_ms": ours * 1e3, "cv2_ms": ref * 1e3}), flush=True)
