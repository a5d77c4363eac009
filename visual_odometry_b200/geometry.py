"""Device-tensor wrappers over the geometry entry points of libvo_b200 (include/vo_b200.h).

Each function mirrors one cv2 call the reference makes (file:line in the docstrings) but works on
batches of frame pairs resident in HBM.  Inputs/outputs are torch CUDA tensors; nothing here
computes on the CPU.
"""
import torch

from . import _lib

F64 = torch.float64
I32 = torch.int32
U8 = torch.uint8

_ws_cache = {}


def _workspace(nbytes, device, tag):
    """Grow-only scratch buffer per (device, tag) so repeated calls do not re-allocate."""
    key = (str(device), tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=U8, device=device)
        _ws_cache[key] = buf
    return buf


def find_essential_batch(pts, m, K, prob=0.99, threshold=1.0, max_iters=1000, adaptive=True):
    """Batched cv2.findEssentialMat(p1, p2, K, FM_RANSAC, prob, threshold)
    (reference src/image_pair.py:280-286).

    pts [B,n_cap,4] f64 (x1,y1,x2,y2) pixels, m [B] i32 valid rows, K [3,3] f64.
    Returns dict(E [B,3,3], mask [B,n_cap] u8 {0,1}, n_inl [B], n_iters [B], status [B])."""
    lib = _lib.load()
    B, n_cap = pts.shape[0], pts.shape[1]
    dev = pts.device
    assert pts.dtype == F64 and pts.is_contiguous() and pts.shape[2] == 4
    assert m.dtype == I32 and K.dtype == F64
    K = K.contiguous()
    E = torch.empty(B, 3, 3, dtype=F64, device=dev)
    mask = torch.empty(B, n_cap, dtype=U8, device=dev)
    n_inl = torch.empty(B, dtype=I32, device=dev)
    n_iters = torch.empty(B, dtype=I32, device=dev)
    status = torch.empty(B, dtype=I32, device=dev)
    wsb = lib.vo_essential_workspace_bytes(B, n_cap, max_iters)
    ws = _workspace(wsb, dev, "essential")
    _lib.check(lib.vo_essential_ransac_batch(
        _lib.ptr(pts), _lib.ptr(m), n_cap, B, _lib.ptr(K), float(prob), float(threshold),
        int(max_iters), int(bool(adaptive)), _lib.ptr(E), _lib.ptr(mask), _lib.ptr(n_inl),
        _lib.ptr(n_iters), _lib.ptr(status), _lib.ptr(ws), ws.numel(), _lib.cur_stream()))
    return dict(E=E, mask=mask, n_inl=n_inl, n_iters=n_iters, status=status)


def fivept_solve(pts_n, samples):
    """Five-point minimal solver (OpenCV five-point.cpp runKernel) on normalised points.
    pts_n [m,4] f64, samples [S,5] i32 -> (E [S,10,3,3], n_models [S])."""
    lib = _lib.load()
    S = samples.shape[0]
    E = torch.zeros(S, 10, 3, 3, dtype=F64, device=pts_n.device)
    n = torch.zeros(S, dtype=I32, device=pts_n.device)
    _lib.check(lib.vo_fivept_solve(_lib.ptr(pts_n.contiguous()), pts_n.shape[0],
                                   _lib.ptr(samples.contiguous()), S, _lib.ptr(E), _lib.ptr(n),
                                   _lib.cur_stream()))
    return E, n


def sampson_score(E, pts_n, thr2, want_masks=False):
    """EMEstimatorCallback::computeError + findInliers for H hypotheses.
    E [H,3,3] f64, pts_n [m,4] f64 -> counts [H] (and masks [H,m] u8)."""
    lib = _lib.load()
    H, mm = E.shape[0], pts_n.shape[0]
    counts = torch.empty(H, dtype=I32, device=E.device)
    masks = torch.empty(H, mm, dtype=U8, device=E.device) if want_masks else None
    _lib.check(lib.vo_sampson_score(_lib.ptr(E.contiguous()), H, _lib.ptr(pts_n.contiguous()), mm,
                                    float(thr2), _lib.ptr(counts), _lib.ptr(masks),
                                    _lib.cur_stream()))
    return (counts, masks) if want_masks else counts


def ransac_samples(count, n_iters, device="cuda"):
    """cv::RNG(2^64-1) replay of getSubset: [n_iters,5] i32."""
    lib = _lib.load()
    out = torch.empty(n_iters, 5, dtype=I32, device=device)
    _lib.check(lib.vo_ransac_samples(int(count), int(n_iters), _lib.ptr(out), _lib.cur_stream()))
    return out


def recover_pose_batch(E, pts, m, mask, K, dist_thresh=50.0, want_mask=False):
    """Batched cv2.recoverPose(E, p1, p2, K) over the inliers of `mask`
    (reference src/image_pair.py:304-308).  Returns dict(R [B,3,3], t [B,3], n_good [B],
    pose_mask [B,n_cap] u8 {0,255} if requested)."""
    lib = _lib.load()
    B, n_cap = pts.shape[0], pts.shape[1]
    dev = pts.device
    R = torch.empty(B, 3, 3, dtype=F64, device=dev)
    t = torch.empty(B, 3, dtype=F64, device=dev)
    n_good = torch.empty(B, dtype=I32, device=dev)
    pm = torch.empty(B, n_cap, dtype=U8, device=dev) if want_mask else None
    wsb = lib.vo_pose_workspace_bytes(B, n_cap)
    ws = _workspace(wsb, dev, "pose")
    _lib.check(lib.vo_recover_pose_batch(
        _lib.ptr(E.contiguous()), _lib.ptr(pts), _lib.ptr(m), _lib.ptr(mask), n_cap, B,
        _lib.ptr(K.contiguous()), float(dist_thresh), _lib.ptr(R), _lib.ptr(t), _lib.ptr(pm),
        _lib.ptr(n_good), _lib.ptr(ws), ws.numel(), _lib.cur_stream()))
    return dict(R=R, t=t, n_good=n_good, pose_mask=pm)


def triangulate(P1, P2, x1, x2, dehomog=False):
    """cv2.triangulatePoints(P1, P2, x1, x2) (reference src/image_pair.py:332-336) for points
    given as [n,2] f64 tensors; returns X [n,4] (divide by w if dehomog, image_pair.py:339)."""
    lib = _lib.load()
    n = x1.shape[0]
    X = torch.empty(n, 4, dtype=F64, device=x1.device)
    _lib.check(lib.vo_triangulate(_lib.ptr(P1.contiguous()), _lib.ptr(P2.contiguous()),
                                  _lib.ptr(x1.contiguous()), _lib.ptr(x2.contiguous()), n,
                                  _lib.ptr(X), int(bool(dehomog)), _lib.cur_stream()))
    return X


def triangulate_pairs(R, t, pts, m, mask, K):
    """Per pair: ImagePair.reconstruct_3d_points default path (src/image_pair.py:316-339):
    P_first = K[R^T|-R^T t] with frame-1 points, P_second = K[I|0] with frame-2 points, inliers
    only, X /= w.  Returns (X [B,n_cap,4] compacted, n_x [B])."""
    lib = _lib.load()
    B, n_cap = pts.shape[0], pts.shape[1]
    X = torch.empty(B, n_cap, 4, dtype=F64, device=pts.device)
    n_x = torch.empty(B, dtype=I32, device=pts.device)
    _lib.check(lib.vo_triangulate_pairs(_lib.ptr(R.contiguous()), _lib.ptr(t.contiguous()),
                                        _lib.ptr(pts), _lib.ptr(m), _lib.ptr(mask), n_cap, B,
                                        _lib.ptr(K.contiguous()), _lib.ptr(X), _lib.ptr(n_x),
                                        _lib.cur_stream()))
    return X, n_x


def chain_poses(R, t, P0=None):
    """Trajectory chaining (north_star): P_0 = I (or P0), P_k = [R_k|t_k] * P_{k-1}.
    Returns poses [B+1,4,4]."""
    lib = _lib.load()
    B = R.shape[0]
    poses = torch.empty(B + 1, 4, 4, dtype=F64, device=R.device)
    _lib.check(lib.vo_chain_poses(_lib.ptr(R.contiguous()), _lib.ptr(t.contiguous()), B,
                                  _lib.ptr(P0.contiguous()) if P0 is not None else None,
                                  _lib.ptr(poses), _lib.cur_stream()))
    return poses
