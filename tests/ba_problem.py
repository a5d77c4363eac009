"""Synthetic bundle-adjustment problems for the N3 tests: a few cameras on a random walk, world
points in front of them, noisy observations with gross outliers, perturbed starting values."""
import numpy as np


def make_ba(C, P, seed, noise=0.5, pert=0.02, frac=0.7, n_fixed=1, outliers=0.05):
    import cv2
    rng = np.random.default_rng(seed)
    f, cx, cy = 802.832, 565.427, 240.124          # the reference's live intrinsics (K * 0.3)
    W = np.c_[rng.uniform(-12, 12, P), rng.uniform(-7, 7, P), rng.uniform(8, 40, P)]
    R, t = np.zeros((C, 3, 3)), np.zeros((C, 3))
    T = np.eye(4)
    for c in range(C):
        R[c], t[c] = T[:3, :3], T[:3, 3]
        S = np.eye(4)
        S[:3, :3] = cv2.Rodrigues(rng.normal(0, 0.02, 3))[0]
        S[:3, 3] = rng.normal(0, 0.4, 3)
        T = S @ T
    see = rng.random((C, P)) < frac
    oc, op = np.nonzero(see)
    Xc = np.einsum("mij,mj->mi", R[oc], W[op]) + t[oc]
    xy = Xc[:, :2] / Xc[:, 2:3] * f + [cx, cy] + rng.normal(0, noise, (len(oc), 2))
    bad = rng.random(len(oc)) < outliers
    xy[bad] += rng.normal(0, 10, (int(bad.sum()), 2))
    order = rng.permutation(len(oc))               # observation order is not sorted by anything
    oc, op, xy = oc[order], op[order], xy[order]
    R0 = np.array([cv2.Rodrigues(rng.normal(0, pert * 0.3, 3))[0] @ R[c] for c in range(C)])
    t0 = t + rng.normal(0, pert, t.shape)
    fixed = np.zeros(C, bool)
    fixed[:n_fixed] = True
    R0[fixed], t0[fixed] = R[fixed], t[fixed]
    W0 = W + rng.normal(0, pert * 5, W.shape)
    return dict(R=R0, t=t0, fixed=fixed, points=W0, obs_cam=oc.astype(np.int32),
                obs_pt=op.astype(np.int32), obs_xy=xy, f=f, cx=cx, cy=cy)
