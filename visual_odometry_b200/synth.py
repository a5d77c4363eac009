"""Synthetic drone sequences for parity tests and bench.py (the reference ships no
data: inputs are git-ignored, /root/reference/.gitignore:1-6).

Random 3-D points in a slab in front of a 1920x1080 camera that random-walks like a drone;
keypoints are projections + N(0, noise) px rounded through float32 (cv2.KeyPoint.pt is float32);
a fraction `inlier_ratio` of keypoints carries over to the next frame, the rest are replaced by
new points.  ORB descriptors: iid Bernoulli(0.5) 256-bit strings, each observation flips every
bit w.p. 10/256 (pair Hamming ~20); ~1 % exact duplicates exercise tie-breaks.  SIFT descriptors:
integer-valued float32 [128], ||d|| ~ 512, per-observation noise round(N(0,4.2)).
Everything is numpy on the host: this is input generation, not part of the timed path.
"""
from dataclasses import dataclass

import numpy as np

K_DEFAULT = np.array([[1400.0, 0.0, 960.0], [0.0, 1400.0, 540.0], [0.0, 0.0, 1.0]])
# the reference's live intrinsics: DJI 4K matrix scaled by 0.3 (src/visual_slam.py:30-37; the
# committed test.g2o:1 pins f=802.832 cx=565.427 cy=240.124)
K_REFERENCE = np.array([[2676.1051390718389, 0.0, 3840 / 2 - 35.243952918157035],
                        [0.0, 2676.1051390718389, 2160 / 2 - 279.58562078697361],
                        [0.0, 0.0, 1.0]]) * 0.3
K_REFERENCE[2, 2] = 1.0
WIDTH, HEIGHT = 1920, 1080


def rodrigues(rv):
    """Axis-angle -> rotation matrix (same map as cv2.Rodrigues, visual_slam.py:243)."""
    th = float(np.linalg.norm(rv))
    if th < 1e-300:
        return np.eye(3)
    k = np.asarray(rv, np.float64) / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


@dataclass
class Sequence:
    desc: np.ndarray      # [F,N,32] uint8 (ORB) or [F,N,128] float32 (SIFT)
    kp: np.ndarray        # [F,N,2] float32 pixel coordinates
    nkp: np.ndarray       # [F] int32
    point_id: np.ndarray  # [F,N] int64 id of the 3-D point behind each keypoint
    points3d: np.ndarray  # [n_points,3] world coordinates (index = point id)
    R_w2c: np.ndarray     # [F,3,3] world -> camera
    t_w2c: np.ndarray     # [F,3]
    K: np.ndarray

    def relative_pose(self, a, b):
        """(R, t) with X_b = R X_a + t — the convention cv2.recoverPose returns."""
        R = self.R_w2c[b] @ self.R_w2c[a].T
        t = self.t_w2c[b] - R @ self.t_w2c[a]
        return R, t

    def consecutive_pairs(self):
        f = len(self.nkp)
        return np.stack([np.arange(f - 1), np.arange(1, f)], 1).astype(np.int32)


def _new_points_cam(rng, n):
    return np.stack([rng.uniform(-12, 12, n), rng.uniform(-7, 7, n), rng.uniform(8, 40, n)], 1)


def _orb_base(rng, n):
    return rng.integers(0, 256, (n, 32), dtype=np.uint8)


def _sift_base(rng, n):
    d = np.abs(rng.standard_normal((n, 128)))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = np.minimum(d, 0.2)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.clip(np.rint(d * 512.0), 0, 255).astype(np.float32)


def make_sequence(n_frames, n_kp, seed, inlier_ratio=0.7, noise_px=0.5, descriptor="orb",
                  K=K_DEFAULT, dup_fraction=0.01):
    rng = np.random.default_rng(seed)
    orb = descriptor == "orb"
    K = np.asarray(K, np.float64)
    R = np.eye(3)
    t = np.zeros(3)
    next_id = 0
    all_pts, all_base, all_noflip = [], [], []

    def spawn(n, R, t):
        nonlocal next_id
        Xc = _new_points_cam(rng, n)
        # keep only points projecting inside the image (resample the others closer to the axis)
        for _ in range(8):
            uv = (K @ Xc.T).T
            uv = uv[:, :2] / uv[:, 2:]
            bad = (uv[:, 0] < 4) | (uv[:, 0] > WIDTH - 4) | (uv[:, 1] < 4) | (uv[:, 1] > HEIGHT - 4)
            if not bad.any():
                break
            Xc[bad] = _new_points_cam(rng, int(bad.sum())) * np.array([0.5, 0.5, 1.0])
        Xw = (R.T @ (Xc - t).T).T
        base = _orb_base(rng, n) if orb else _sift_base(rng, n)
        noflip = np.zeros(n, bool)
        nd = int(round(n * dup_fraction / 2)) if n >= 4 else 0
        if nd > 0:  # exact duplicate descriptors on distinct points, never perturbed
            a = rng.choice(n, 2 * nd, replace=False)
            base[a[nd:]] = base[a[:nd]]
            noflip[a] = True
        ids = np.arange(next_id, next_id + n)
        next_id += n
        all_pts.append(Xw)
        all_base.append(base)
        all_noflip.append(noflip)
        return ids, Xw, base, noflip

    ids, Xw, base, noflip = spawn(n_kp, R, t)
    D = 32 if orb else 128
    desc = np.empty((n_frames, n_kp, D), np.uint8 if orb else np.float32)
    kp = np.empty((n_frames, n_kp, 2), np.float32)
    pid = np.empty((n_frames, n_kp), np.int64)
    Rs = np.empty((n_frames, 3, 3))
    ts = np.empty((n_frames, 3))
    for f in range(n_frames):
        if f > 0:
            dR = rodrigues(rng.normal(0, 0.03, 3))
            d = np.array([rng.normal(0, 0.35), rng.normal(0, 0.15), -abs(rng.normal(0.8, 0.3))])
            d *= rng.uniform(0.3, 1.0) / np.linalg.norm(d)
            R = dR @ R
            t = dR @ t + d
            Xc = (R @ Xw.T).T + t
            uv = (K @ Xc.T).T
            with np.errstate(divide="ignore", invalid="ignore"):
                uv = uv[:, :2] / uv[:, 2:]
            gone = (Xc[:, 2] < 1.0) | (uv[:, 0] < 0) | (uv[:, 0] > WIDTH - 1) | \
                   (uv[:, 1] < 0) | (uv[:, 1] > HEIGHT - 1)
            n_rep = max(int(round(n_kp * (1 - inlier_ratio))), int(gone.sum()))
            rep = np.nonzero(gone)[0]
            if n_rep > len(rep):
                rest = np.nonzero(~gone)[0]
                rep = np.concatenate([rep, rng.choice(rest, n_rep - len(rep), replace=False)])
            nid, nX, nb, nf = spawn(len(rep), R, t)
            ids[rep], Xw[rep], base[rep], noflip[rep] = nid, nX, nb, nf
        Xc = (R @ Xw.T).T + t
        uv = (K @ Xc.T).T
        uv = uv[:, :2] / uv[:, 2:] + rng.normal(0, noise_px, (n_kp, 2))
        if orb:
            flips = rng.random((n_kp, 256)) < (10.0 / 256.0)
            flips[noflip] = False
            obs = base ^ np.packbits(flips, axis=1)
        else:
            nz = np.rint(rng.normal(0, 4.2, (n_kp, 128)))
            nz[noflip] = 0
            obs = np.clip(base + nz, 0, 255).astype(np.float32)
        perm = rng.permutation(n_kp)
        desc[f] = obs[perm]
        kp[f] = uv[perm].astype(np.float32)
        pid[f] = ids[perm]
        Rs[f], ts[f] = R, t
    return Sequence(desc=desc, kp=kp, nkp=np.full(n_frames, n_kp, np.int32), point_id=pid,
                    points3d=np.concatenate(all_pts), R_w2c=Rs, t_w2c=ts, K=K)


def make_pair(n_kp, seed, **kw):
    """One frame pair (config C1): a 2-frame sequence."""
    return make_sequence(2, n_kp, seed, **kw)


def textured_image(height, width, seed, n_shapes=None):
    """Synthetic gray image with corners for the ORB front end (N4): random filled rectangles and
    discs on a mid-gray ground, 3x3 box smoothing, Gaussian pixel noise.  numpy only, so the GPU
    box regenerates exactly the image the golden fixtures were made from."""
    rng = np.random.default_rng(seed)
    img = np.full((height, width), 110.0, np.float32)
    n_shapes = n_shapes if n_shapes is not None else height * width // 1200
    yy, xx = np.mgrid[0:64, 0:64]
    for _ in range(n_shapes):
        x, y = int(rng.integers(0, width)), int(rng.integers(0, height))
        w, h = int(rng.integers(6, 60)), int(rng.integers(6, 60))
        c = float(rng.integers(20, 236))
        x1, y1 = min(x + w, width), min(y + h, height)
        if rng.random() < 0.5:
            img[y:y1, x:x1] = c
        else:
            m = ((xx[:y1 - y, :x1 - x] - w / 2) ** 2 / (w / 2) ** 2 + (yy[:y1 - y, :x1 - x] - h / 2) ** 2 / (h / 2) ** 2) <= 1
            img[y:y1, x:x1][m] = c
    p = np.pad(img, 1, mode="edge")
    img = sum(p[dy:dy + height, dx:dx + width] for dy in range(3) for dx in range(3)) / 9.0
    img = img + rng.normal(0, 2.5, img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)
