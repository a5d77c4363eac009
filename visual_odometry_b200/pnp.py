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
    # one upload (X | x | K packed as float64) and one download (rvec, tvec | n_inl, ok | inliers)
    host_in = np.concatenate([X.ravel(), x.ravel(),
                              np.asarray(cameraMatrix, np.float64).ravel()])
    din = torch.from_numpy(host_in).to(dev)
    base = din.data_ptr()
    out = torch.zeros(48 + 8 + 4 * n, dtype=torch.uint8, device=dev)
    ob = out.data_ptr()
    vp = _lib.C.c_void_p
    wsb = lib.vo_pnp_workspace_bytes(n, int(iterationsCount))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    _lib.check(lib.vo_pnp_ransac(vp(base), vp(base + 24 * n), n, vp(base + 40 * n),
                                 int(iterationsCount), float(reprojectionError), float(confidence),
                                 vp(ob), vp(ob + 24), vp(ob + 56), vp(ob + 48), vp(ob + 52),
                                 _lib.ptr(ws), wsb, _lib.cur_stream()))
    host = out.cpu().numpy()
    rt = host[:48].view(np.float64)
    n_inl, ok = (int(v) for v in host[48:56].view(np.int32))
    if not ok:
        return False, np.zeros((3, 1)), np.zeros((3, 1)), None
    return (True, rt[:3].copy().reshape(3, 1), rt[3:6].copy().reshape(3, 1),
            host[56:56 + 4 * n_inl].view(np.int32).copy().reshape(-1, 1))


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


def solve_pnp_ransac_device(X, x, K_dev, iterationsCount=100, reprojectionError=8.0, confidence=0.99):
    """solvePnPRansac on DEVICE inputs (X [n,3], x [n,2] float64 CUDA tensors, K_dev [9] float64 on
    the device): no upload; one 56-byte read-back.  -> (retval, rvec [3,1], tvec [3,1], n_inliers)."""
    n = int(X.shape[0])
    if n < 4 or int(x.shape[0]) != n:
        raise CvError("solvePnPRansac: npoints >= 4 && npoints == max(ipoints.checkVector(2))")
    if n < 5:
        raise NotImplementedError("libvo_b200 PnP: the n == 4 (P3P) special case is not implemented")
    lib = _lib.load()
    X, x = X.contiguous(), x.contiguous()
    out = torch.zeros(48 + 8 + 4 * n, dtype=torch.uint8, device=X.device)
    ob = out.data_ptr()
    vp = _lib.C.c_void_p
    wsb = lib.vo_pnp_workspace_bytes(n, int(iterationsCount))
    ws = torch.empty(wsb, dtype=torch.uint8, device=X.device)
    _lib.check(lib.vo_pnp_ransac(_lib.ptr(X), _lib.ptr(x), n, _lib.ptr(K_dev), int(iterationsCount),
                                 float(reprojectionError), float(confidence), vp(ob), vp(ob + 24),
                                 vp(ob + 56), vp(ob + 48), vp(ob + 52), _lib.ptr(ws), wsb,
                                 _lib.cur_stream()))
    host = out[:56].cpu().numpy()
    rt = host[:48].view(np.float64)
    n_inl, ok = (int(v) for v in host[48:56].view(np.int32))
    if not ok:
        return False, np.zeros((3, 1)), np.zeros((3, 1)), 0
    return True, rt[:3].copy().reshape(3, 1), rt[3:6].copy().reshape(3, 1), n_inl
