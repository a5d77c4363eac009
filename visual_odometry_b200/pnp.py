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


def solve_pnp_ransac_batch_device(X, x, offsets, K_dev, iterationsCount=100, reprojectionError=8.0,
                                  confidence=0.99, out=None):
    """cv2.solvePnPRansac for MANY frames against one device-resident map in one launch group
    (BASELINE config C5; reference call src/visual_slam.py:231-235, once per frame there).
    X [n_total,3], x [n_total,2] float64 CUDA tensors, offsets int32 [F+1] CUDA (CSR: frame f owns
    rows offsets[f]:offsets[f+1]), K_dev [9] float64 CUDA.  Asynchronous, no host read-back.
    Returns dict(rvec [F,3], tvec [F,3], inliers [n_total] (frame-local indices from offsets[f]),
    n_inl [F], ok [F]) of CUDA tensors; pass the dict back as `out` to reuse the buffers."""
    lib = _lib.load()
    F, n_total = int(offsets.shape[0]) - 1, int(X.shape[0])
    dev = X.device
    if out is None:
        wsb = lib.vo_pnp_batch_workspace_bytes(max(n_total, 1), max(F, 1), int(iterationsCount))
        out = dict(rvec=torch.zeros(F, 3, dtype=torch.float64, device=dev),
                   tvec=torch.zeros(F, 3, dtype=torch.float64, device=dev),
                   inliers=torch.zeros(max(n_total, 1), dtype=torch.int32, device=dev),
                   n_inl=torch.zeros(F, dtype=torch.int32, device=dev),
                   ok=torch.zeros(F, dtype=torch.int32, device=dev),
                   ws=torch.empty(max(wsb, 256), dtype=torch.uint8, device=dev))
    _lib.check(lib.vo_pnp_ransac_batch(
        _lib.ptr(X), _lib.ptr(x), _lib.ptr(offsets), n_total, F, _lib.ptr(K_dev), int(iterationsCount),
        float(reprojectionError), float(confidence), _lib.ptr(out["rvec"]), _lib.ptr(out["tvec"]),
        _lib.ptr(out["inliers"]), _lib.ptr(out["n_inl"]), _lib.ptr(out["ok"]), _lib.ptr(out["ws"]),
        out["ws"].numel(), _lib.cur_stream()))
    return out


def solve_pnp_ransac_batch(object_points, image_points, cameraMatrix, iterationsCount=100,
                           reprojectionError=8.0, confidence=0.99):
    """Host lists in, list of cv2-shaped results out: [(retval, rvec [3,1], tvec [3,1],
    inliers int32 [k,1] or None), ...] — one upload, one launch group, one download."""
    Xs = [np.asarray(a, np.float64).reshape(-1, 3) for a in object_points]
    xs = [np.asarray(a, np.float64).reshape(-1, 2) for a in image_points]
    if any(len(a) != len(b) for a, b in zip(Xs, xs)):
        raise CvError("solvePnPRansac: npoints == max(ipoints.checkVector(2))")
    off = np.zeros(len(Xs) + 1, np.int32)
    off[1:] = np.cumsum([len(a) for a in Xs])
    dev = torch.device("cuda")
    cat = lambda parts, w: np.concatenate(parts, 0) if off[-1] else np.zeros((0, w))
    X = torch.from_numpy(np.ascontiguousarray(cat(Xs, 3))).to(dev)
    x = torch.from_numpy(np.ascontiguousarray(cat(xs, 2))).to(dev)
    K_dev = torch.from_numpy(np.asarray(cameraMatrix, np.float64).reshape(9).copy()).to(dev)
    o = solve_pnp_ransac_batch_device(X, x, torch.from_numpy(off).to(dev), K_dev, iterationsCount,
                                      reprojectionError, confidence)
    rv, tv = o["rvec"].cpu().numpy(), o["tvec"].cpu().numpy()
    inl, n_inl, ok = o["inliers"].cpu().numpy(), o["n_inl"].cpu().numpy(), o["ok"].cpu().numpy()
    res = []
    for f in range(len(Xs)):
        if not ok[f]:
            res.append((False, np.zeros((3, 1)), np.zeros((3, 1)), None))
        else:
            res.append((True, rv[f].copy().reshape(3, 1), tv[f].copy().reshape(3, 1),
                        inl[off[f]:off[f] + n_inl[f]].copy().reshape(-1, 1)))
    return res


def bench_batch(dev, map_points, K, rng, timed, n_frames=64, n_obs=5000):
    """bench.py `configs` leg for C5: `n_frames` new frames, each observing `n_obs` points of the
    100 k-point map (0.5 px noise, 20 % outliers), PnP-RANSAC for all of them in one launch group,
    device-resident; beside cv2.solvePnPRansac on the host for two of the frames."""
    import time
    import cv2
    n_map = len(map_points)
    Xs, xs = [], []
    for f in range(n_frames):
        rv = rng.normal(0, 0.03, 3)
        tv = rng.normal(0, 0.4, 3)
        sel = rng.permutation(n_map)[:n_obs]
        X = map_points[sel]
        img = cv2.projectPoints(X, rv, tv, K, None)[0].reshape(-1, 2) + rng.normal(0, 0.5, (n_obs, 2))
        bad = rng.random(n_obs) < 0.2
        img[bad] = rng.uniform([0, 0], [1920, 1080], (int(bad.sum()), 2))
        Xs.append(X)
        xs.append(img)
    off = torch.from_numpy(np.arange(n_frames + 1, dtype=np.int32) * n_obs).to(dev)
    Xd = torch.from_numpy(np.ascontiguousarray(np.concatenate(Xs))).to(dev)
    xd = torch.from_numpy(np.ascontiguousarray(np.concatenate(xs))).to(dev)
    Kd = torch.from_numpy(np.asarray(K, np.float64).reshape(9).copy()).to(dev)
    out = solve_pnp_ransac_batch_device(Xd, xd, off, Kd)
    ms = timed(lambda: solve_pnp_ransac_batch_device(Xd, xd, off, Kd, out=out), warm=2, reps=5)
    t0 = time.perf_counter()
    refs = [cv2.solvePnPRansac(Xs[f], xs[f], K, np.zeros(4)) for f in range(2)]
    cv_ms = (time.perf_counter() - t0) / 2 * 1e3
    rv = out["rvec"][:2].cpu().numpy()
    dr = max(float(np.abs(rv[f] - refs[f][1].ravel()).max()) for f in range(2))
    return {"what": f"configs[4]: PnP-RANSAC of {n_frames} new frames x {n_obs} observations against the "
                    "100 k-point map, one vo_pnp_ransac_batch launch group (device-resident)",
            "ms": ms, "frames_per_s": n_frames / (ms * 1e-3), "cv2_ms_per_frame_host": cv_ms,
            "cv2_frames_per_s_1_thread": 1e3 / cv_ms, "n_inliers_mean": float(out["n_inl"].float().mean()),
            "max_abs_rvec_diff_vs_cv2_first_2_frames": dr}
