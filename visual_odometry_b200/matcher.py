"""Seam 1 of the drop-in boundary: a matcher object with cv2.BFMatcher's call surface.

`B200Matcher(norm, crossCheck=True)` can be handed to the reference's ImagePair as its `matcher`
argument (reference src/image_pair.py:227-236; built at src/visual_slam.py:19 and
src/image_and_keypoints.py:9,13).  `.match` = mutual nearest neighbour when crossCheck else
nearest neighbour; `.knnMatch(k=2)` = two best, stable by (distance, index); `.ratio_match`
= Lowe ratio test as in src/feature_detection.py:20-26.  All arithmetic runs in libvo_b200.
"""
import weakref

import numpy as np
import torch

from . import _lib

NORM_L2 = 4        # cv2.NORM_L2
NORM_HAMMING = 6   # cv2.NORM_HAMMING


class DMatch:
    """Field-compatible stand-in for cv2.DMatch."""
    __slots__ = ("queryIdx", "trainIdx", "imgIdx", "distance")

    def __init__(self, queryIdx, trainIdx, distance, imgIdx=0):
        self.queryIdx = queryIdx
        self.trainIdx = trainIdx
        self.imgIdx = imgIdx
        self.distance = distance

    def __repr__(self):
        return f"DMatch(q={self.queryIdx}, t={self.trainIdx}, d={self.distance})"


# Device copies of descriptor arrays, keyed by the identity of the host array (a Frame's
# descriptors are matched twice: as frame2 of one pair, then as frame1 of the next).  Only arrays
# the caller has made READ-ONLY (`a.flags.writeable = False`) are cached: a writable array may be
# refilled in place between two calls (a video loop reusing one buffer) and cv2.BFMatcher would
# see the new contents, so it is uploaded on every call.  Entries die with their array.
_DEV_CACHE = {}


def _dev(a, dtype):
    if isinstance(a, torch.Tensor):
        return a.to(device="cuda", dtype=dtype).contiguous()
    key = id(a)
    cacheable = isinstance(a, np.ndarray) and not a.flags.writeable
    hit = _DEV_CACHE.get(key) if cacheable else None
    if hit is not None and hit[0]() is a and hit[1].dtype == dtype:
        return hit[1]
    t = torch.from_numpy(np.ascontiguousarray(a, dtype={torch.uint8: np.uint8,
                                                        torch.float32: np.float32}[dtype])).cuda()
    if cacheable:
        try:
            _DEV_CACHE[key] = (weakref.ref(a, lambda _r, k=key: _DEV_CACHE.pop(k, None)), t)
        except TypeError:
            pass
    return t


class B200Matcher:
    def __init__(self, normType=NORM_HAMMING, crossCheck=False):
        if normType not in (NORM_HAMMING, NORM_L2):
            raise ValueError("B200Matcher supports NORM_HAMMING (ORB) and NORM_L2 (SIFT)")
        self.normType = normType
        self.crossCheck = bool(crossCheck)
        self.lib = _lib.load()

    # ---- raw device-side call: returns torch tensors (q, t, dist, t2, dist2) trimmed to n ----
    def match_arrays(self, desc1, desc2, mode=None, ratio=0.0):
        if mode is None:
            mode = _lib.MATCH_MUTUAL
        if self.normType == NORM_HAMMING:
            d1 = _dev(desc1, torch.uint8)
            d2 = _dev(desc2, torch.uint8)
            if d1.ndim != 2 or d2.ndim != 2 or (d1.shape[0] and d1.shape[1] != 32) or \
                    (d2.shape[0] and d2.shape[1] != 32):
                raise ValueError("Hamming descriptors must be uint8 [N,32]")
            return self._hamming(d1, d2, mode, ratio)
        from .l2 import match_l2_single
        return match_l2_single(self.lib, desc1, desc2, mode, ratio)

    def _hamming(self, d1, d2, mode, ratio):
        n1, n2 = d1.shape[0], d2.shape[0]
        cap = max(n1, n2, 1)
        dev = d1.device
        q = torch.empty(max(n1, 1), dtype=torch.int32, device=dev)
        t = torch.empty_like(q)
        t2 = torch.empty_like(q)
        dist = torch.empty(max(n1, 1), dtype=torch.float32, device=dev)
        dist2 = torch.empty_like(dist)
        n_out = torch.zeros(1, dtype=torch.int32, device=dev)
        wsb = self.lib.vo_match_workspace_bytes(0, 1, cap)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        _lib.check(self.lib.vo_match_hamming(
            _lib.ptr(d1), n1, _lib.ptr(d2), n2, mode, float(ratio), _lib.ptr(q), _lib.ptr(t),
            _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2), _lib.ptr(n_out), _lib.ptr(ws), wsb,
            _lib.cur_stream()))
        n = int(n_out.item())
        return q[:n], t[:n], dist[:n], t2[:n], dist2[:n]

    # ---- cv2.BFMatcher surface ----
    def match_host(self, queryDescriptors, trainDescriptors):
        """(queryIdx, trainIdx, distance) as three Python lists — one device->host transfer."""
        if self.crossCheck:
            q, t, d, _, _ = self.match_arrays(queryDescriptors, trainDescriptors,
                                              _lib.MATCH_MUTUAL)
        else:
            q, t, d, _, _ = self._all_nn(queryDescriptors, trainDescriptors)
        if q.numel() == 0:
            return [], [], []
        packed = torch.stack((q, t, d.view(torch.int32))).cpu().numpy()
        return packed[0].tolist(), packed[1].tolist(), packed[2].view(np.float32).tolist()

    def match(self, queryDescriptors, trainDescriptors, mask=None):
        q, t, d = self.match_host(queryDescriptors, trainDescriptors)
        return [DMatch(a, b, c) for a, b, c in zip(q, t, d)]

    def _all_nn(self, d1, d2):
        # nearest neighbour for every query = knn2 mode with the ratio test disabled
        return self.match_arrays(d1, d2, _lib.MATCH_KNN2_RATIO, -1.0)

    def knnMatch(self, queryDescriptors, trainDescriptors, k=2, mask=None, compactResult=False):
        if k not in (1, 2):
            raise ValueError("B200Matcher.knnMatch supports k in (1, 2)")
        q, t, d, t2, d2 = self._all_nn(queryDescriptors, trainDescriptors)
        q, t, d, t2, d2 = q.tolist(), t.tolist(), d.tolist(), t2.tolist(), d2.tolist()
        out = []
        for a, b, c, b2, c2 in zip(q, t, d, t2, d2):
            row = [DMatch(a, b, c)]
            if k == 2 and b2 >= 0:
                row.append(DMatch(a, b2, c2))
            out.append(row)
        return out

    def ratio_match(self, queryDescriptors, trainDescriptors, ratio=0.3):
        """knn(k=2) + `m.distance < ratio * n.distance` (src/feature_detection.py:22-26)."""
        q, t, d, _, _ = self.match_arrays(queryDescriptors, trainDescriptors,
                                          _lib.MATCH_KNN2_RATIO, ratio)
        return [DMatch(a, b, c) for a, b, c in zip(q.tolist(), t.tolist(), d.tolist())]
