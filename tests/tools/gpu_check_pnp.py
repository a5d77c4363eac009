import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, cv2, torch
from visual_odometry_b200 import pnp
from oracle import cv_pnp as P
K = np.array([[1400, 0, 960], [0, 1400, 540], [0, 0, 1.]])
def scene(seed, n=500, out=0.2):
    rng = np.random.default_rng(seed)
    X = np.c_[rng.uniform(-12, 12, n), rng.uniform(-7, 7, n), rng.uniform(8, 40, n)]
    rv = rng.normal(0, 0.1, 3); t = rng.normal(0, 1, 3)
    x = P.project(X, rv, t, K) + rng.normal(0, 0.5, (n, 2))
    no = int(n * out); x[:no] = np.c_[rng.uniform(0, 1920, no), rng.uniform(0, 1080, no)]
    return X, x
same = 0; worst_r = worst_t = 0; N = 24; ok_all = True
for seed in range(N):
    X, x = scene(seed, n=200 + 150 * seed, out=0.1 + 0.02 * (seed % 12))
    okc, r, tv, inl = cv2.solvePnPRansac(X, x, K, np.zeros(4))
    t0 = time.time(); okg, r2, tv2, inl2 = pnp.solve_pnp_ransac(X, x, K, np.zeros(4)); tg = time.time() - t0
    s = inl is not None and inl2 is not None and np.array_equal(inl.ravel(), inl2.ravel())
    same += s
    dr = np.abs(r - r2).max(); dt = np.abs(tv - tv2).max() / max(1.0, np.abs(tv).max())
    if s: worst_r = max(worst_r, dr); worst_t = max(worst_t, dt)
    sym = 0 if inl is None or inl2 is None else len(set(inl.ravel()) ^ set(inl2.ravel()))
    print(seed, len(X), okc, okg, len(inl), len(inl2), "same" if s else "DIFF(%d)" % sym, "dr %.2e dt %.2e" % (dr, dt), "gpu ms %.2f" % (tg * 1e3))
    ok_all &= okg and dr < 1e-3 and dt < 1e-3
print("identical inlier sets", same, "of", N, "; worst (same sets) dr %.2e dt %.2e" % (worst_r, worst_t))
ok_all &= same >= N * 0.6 and worst_r < 1e-6
print("ALL OK" if ok_all else "FAILED"); sys.exit(0 if ok_all else 1)
