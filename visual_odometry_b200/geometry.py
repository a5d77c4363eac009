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
