"""Synthetic multi-frame scene for the map-bookkeeping tests: world points seen by a slowly moving
camera, per-frame feature slots, per-pair inlier matches with triangulated points (world frame)."""
import numpy as np


def make_scene(F, N, NW, seed, K):
    import cv2
    rng = np.random.default_rng(seed)
    W = np.c_[rng.uniform(-12, 12, NW), rng.uniform(-7, 7, NW), rng.uniform(8, 40, NW)]
    W[::17, 2] += 40.0                                   # some points beyond |X| = 50
    poses = [np.eye(4)]
    for _ in range(1, F):
        T = np.eye(4)
        T[:3, :3] = cv2.Rodrigues(rng.normal(0, 0.02, 3))[0]
        T[:3, 3] = rng.normal(0, 0.3, 3)
        poses.append(T @ poses[-1])
    n_track = min(NW, N)
    vis = rng.random((F, NW)) < 0.8
    slot = np.array([rng.permutation(N) for _ in range(F)])
    kp = rng.uniform(0, 1000, (F, N, 2)).astype(np.float32)
    for f in range(F):
        x = (K @ (poses[f] @ np.c_[W, np.ones(NW)].T)[:3]).T
        xy = x[:, :2] / x[:, 2:3] + rng.normal(0, 0.3, (NW, 2))
        gross = rng.random(NW) < 0.1
        xy[gross] += rng.normal(0, 4, (int(gross.sum()), 2))
        kp[f, slot[f, :n_track]] = xy[:n_track]
    pairs = []
    for b in range(F - 1):
        js = [j for j in rng.permutation(n_track) if vis[b, j] and vis[b + 1, j]]
        q = slot[b, js].astype(np.int32)
        t = slot[b + 1, js].astype(np.int32)
        # E-RANSAC mask over the raw matches: ~85 % inliers; triangulated points only for those
        mask = (rng.random(len(js)) < 0.85).astype(np.uint8)
        Xw = W[js][mask.astype(bool)] + rng.normal(0, 0.01, (int(mask.sum()), 3))
        pairs.append(dict(q=q, t=t, mask=mask, X=Xw))
    return dict(W=W, poses=poses, kp=kp, pairs=pairs)
