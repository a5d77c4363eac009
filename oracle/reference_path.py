"""oracle/reference_path.py — the reference's per-pair call sequence restated on cv2 arrays.

TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline); never imported by the product package.

The reference's hot path is four cv2 calls made by ImagePair (src/image_pair.py:234 match,
:280 findEssentialMat, :304 recoverPose, :332 triangulatePoints) plus list bookkeeping.  This
module performs the same calls with the same arguments on packed arrays (no Match namedtuples —
SURVEY.md §8d excludes tuple materialisation on both sides), so it can run on the GPU box where
/root/reference does not exist.  tests/test_oracle.py checks it against the golden vectors that
tests/golden/make_golden.py produced by running the REAL reference classes.
"""
import numpy as np

import cv2


def match_cv2(desc1, desc2, norm, cross_check=True):
    """src/image_pair.py:234 with the matcher of src/visual_slam.py:18-19."""
    ms = cv2.BFMatcher(norm, crossCheck=cross_check).match(desc1, desc2)
    q = np.fromiter((m.queryIdx for m in ms), np.int32, len(ms))
    t = np.fromiter((m.trainIdx for m in ms), np.int32, len(ms))
    d = np.fromiter((m.distance for m in ms), np.float32, len(ms))
    return q, t, d


def pair_cv2(desc1, kp1, desc2, kp2, K, norm=cv2.NORM_HAMMING, triangulate=True):
    """match_features -> determine_essential_matrix -> estimate_camera_movement ->
    reconstruct_3d_points for one frame pair; returns a dict of arrays."""
    q, t, d = match_cv2(desc1, desc2, norm)
    p1 = np.asarray(kp1, np.float32)[q].astype(np.float64)   # get_image_points (:294-299)
    p2 = np.asarray(kp2, np.float32)[t].astype(np.float64)
    out = dict(q=q, t=t, dist=d, p1=p1, p2=p2)
    E, mask = cv2.findEssentialMat(p1, p2, K, cv2.FM_RANSAC, 0.99, 1)
    out["E"], out["mask"] = E, mask
    if E is None or E.shape != (3, 3):
        return out
    inl = mask.ravel() == 1
    p1i, p2i = p1[inl], p2[inl]
    n_good, R, tt, pmask = cv2.recoverPose(E, p1i, p2i, K)
    out.update(R=R, tvec=tt, n_good=n_good, pose_mask=pmask)
    if triangulate:
        P_est = K @ np.hstack((R.T, -R.T @ tt))        # projection_matrix (:320-323)
        P_null = K @ np.eye(3, 4)                      # null_projection_matrix
        X = cv2.triangulatePoints(P_est, P_null, p1i.T, p2i.T)
        X /= X[3, :]
        out["X"] = X
    return out


def chain_poses(Rs, ts):
    """north_star trajectory chaining: P_0 = I, P_k = [R_k|t_k] P_{k-1}."""
    P = np.eye(4)
    out = [P.copy()]
    for R, t in zip(Rs, ts):
        rel = np.eye(4)
        rel[:3, :3] = R
        rel[:3, 3] = np.asarray(t).ravel()
        P = rel @ P
        out.append(P.copy())
    return np.stack(out)
