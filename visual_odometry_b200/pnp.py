"""cv2.solvePnPRansac / cv2.Rodrigues replacements for the map-tracking step
(reference src/visual_slam.py:231-235,243), bound to csrc/pnp.cu through the C-ABI."""
import numpy as np
import torch

from . import _lib

try:
    import cv2 as _cv2
    CvError = _cv2.error
except Exception:  # pragma: no cover
    class CvError(Exception):
        pass


def solve_pnp_ransac(objectPoints, imagePoints, cameraMatrix, distCoeffs=None, rvec=None, tvec=None,
                     useExtrinsicGuess=False, iterationsCount=100, reprojectionError=8.0,
                     confidence=0.99, inliers=None, flags=0):
    """Same return as cv2.solvePnPRansac: (retval, rvec [3,1] f64, tvec [3,1] f64,
    inliers int32 [k,1]).  Defaults = the ones the reference relies on (100 iterations, 8 px,
    0.99, SOLVEPNP_ITERATIVE)."""
    if distCoeffs is not None and np.any(np.asarray(distCoeffs, np.float64) != 0):
        raise NotImplementedError("libvo_b200 PnP: lens distortion is not modelled "
                                  "(the reference passes zeros, src/visual_slam.py:234)")
    if useExtrinsicGuess or flags not in (0,):
        raise NotImplementedError("libvo_b200 PnP: only the default SOLVEPNP_ITERATIVE path")
    X = np.ascontiguousarray(np.asarray(objectPoints, np.float64).reshape(-1, 3))
    x = np.ascontiguousarray(np.asarray(imagePoints, np.float64).reshape(-1, 2))
    n = len(X)
    if n < 4 or len(x) != n:
        raise CvError("solvePnPRansac: npoints >= 4 && npoints == max(ipoints.checkVector(2))")
    if n < 5:
        raise NotImplementedError("libvo_b200 PnP: the n == 4 (P3P) special case is not implemented")
    lib = _lib.load()
    dev = torch.device("cuda")
    Xd, xd = torch.from_numpy(X).to(dev), torch.from_numpy(x).to(dev)
    Kd = torch.from_numpy(np.ascontiguousarray(np.asarray(cameraMatrix, np.float64))).to(dev)
    rv = torch.zeros(3, dtype=torch.float64, device=dev)
    tv = torch.zeros(3, dtype=torch.float64, device=dev)
    inl = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.zeros(2, dtype=torch.int32, device=dev)
    wsb = lib.vo_pnp_workspace_bytes(n, int(iterationsCount))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    _lib.check(lib.vo_pnp_ransac(_lib.ptr(Xd), _lib.ptr(xd), n, _lib.ptr(Kd), int(iterationsCount),
                                 float(reprojectionError), float(confidence), _lib.ptr(rv),
                                 _lib.ptr(tv), _lib.ptr(inl), _lib.ptr(cnt[0:1]), _lib.ptr(cnt[1:2]),
                                 _lib.ptr(ws), wsb, _lib.cur_stream()))
    n_inl, ok = (int(v) for v in cnt.tolist())
    if not ok:
        return False, np.zeros((3, 1)), np.zeros((3, 1)), None
    return (True, rv.cpu().numpy().reshape(3, 1), tv.cpu().numpy().reshape(3, 1),
            inl[:n_inl].cpu().numpy().reshape(-1, 1))


def rodrigues(src):
    """cv2.Rodrigues: rotation vector [3]/[3,1] -> (R [3,3], jacobian None) or matrix -> vector.
    Nine flops on the host ('trivial device/host function')."""
    a = np.asarray(src, np.float64)
    if a.size == 3:
        r = a.ravel()
        th = float(np.linalg.norm(r))
        if th < 2.220446049250313e-16:
            return np.eye(3), None
        k = r / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * Kx, None
    R = a.reshape(3, 3)
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = np.sqrt((r @ r) * 0.25)
    c = np.clip((np.trace(R) - 1) * 0.5, -1.0, 1.0)
    th = np.arccos(c)
    if s < 1e-5:
        if c > 0:
            return np.zeros((3, 1)), None
        t = np.sqrt(np.maximum((np.diag(R) + 1) * 0.5, 0))
        t[1] *= -1 if R[0, 1] < 0 else 1
        t[2] *= -1 if R[0, 2] < 0 else 1
        return (t * (th / np.linalg.norm(t))).reshape(3, 1), None
    return (r * (0.5 / s) * th).reshape(3, 1), None
