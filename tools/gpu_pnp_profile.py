import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, cv2, torch
from visual_odometry_b200 import pnp as PNP
rng = np.random.default_rng(5)
K = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
R = cv2.Rodrigues(np.array([0.02, -0.03, 0.01]))[0]; t = np.array([0.6, 0.05, 0.1])
for n in (500, 5000):
    X = np.c_[rng.uniform(-12, 12, n), rng.uniform(-7, 7, n), rng.uniform(8, 40, n)]
    x = (K @ (R @ X.T + t[:, None])).T; x = x[:, :2] / x[:, 2:] + rng.normal(0, 0.5, (n, 2))
    out = rng.random(n) < 0.2; x[out] = np.c_[rng.uniform(0, 1920, out.sum()), rng.uniform(0, 1080, out.sum())]
    for _ in range(3): PNP.solve_pnp_ransac(X, x, K, np.zeros(4))
    t0 = time.perf_counter()
    for _ in range(20): PNP.solve_pnp_ransac(X, x, K, np.zeros(4))
    print(n, "ours ms", (time.perf_counter() - t0) / 20 * 1e3)
    t0 = time.perf_counter()
    for _ in range(20): cv2.solvePnPRansac(X, x, K, np.zeros(4))
    print(n, "cv2 ms", (time.perf_counter() - t0) / 20 * 1e3)
