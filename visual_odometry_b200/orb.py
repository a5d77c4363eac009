"""N4: ORB feature extraction on the GPU behind the detector interface the reference
uses (src/frame_generator.py:25 `self.detector.detectAndCompute(frame.image, None)`; the detector
is built at src/visual_slam.py:16 `cv2.ORB_create()` / src/image_and_keypoints.py:8).

`OrbDetector(nfeatures, scaleFactor, nlevels, edgeThreshold, fastThreshold)` mirrors
`cv2.ORB_create` (firstLevel 0, WTA_K 2, HARRIS_SCORE, patchSize 31 are fixed); pass it to
FrameGenerator in place of the cv2 detector.  `detect_batch` takes all frames of a sequence at once
and leaves key points and descriptors on the device in the layout the matcher reads
(desc uint8 [F, cap, 32], kp float32 [F, cap, 2]).  All arithmetic runs in libvo_b200
(csrc/orb.cu); there is no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .frame import KeyPointLite

U8, I32, F32 = torch.uint8, torch.int32, torch.float32


def bgr_to_gray(img):
    """cv2.cvtColor(BGR2GRAY) for uint8 on the device: (3735 B + 19235 G + 9798 R + 2^14) >> 15."""
    x = img.to(torch.int32)
    return ((x[..., 0] * 3735 + x[..., 1] * 19235 + x[..., 2] * 9798 + (1 << 14)) >> 15).to(U8)


class OrbDetector:
    def __init__(self, nfeatures=500, scaleFactor=1.2, nlevels=8, edgeThreshold=31,
                 fastThreshold=20, device="cuda", slack=256):
        self.nfeatures, self.scaleFactor, self.nlevels = int(nfeatures), float(scaleFactor), int(nlevels)
        self.edgeThreshold, self.fastThreshold = int(edgeThreshold), int(fastThreshold)
        self.device = torch.device(device)
        self.slack = int(slack)       # extra output rows for ties at the cuts
        self._ws = None

    def level_info(self, width, height):
        lib = _lib.load()
        n = self.nlevels
        w, h, nf = (C.c_int32 * n)(), (C.c_int32 * n)(), (C.c_int32 * n)()
        sc = (C.c_float * n)()
        _lib.check(lib.vo_orb_level_info(width, height, self.nfeatures, self.scaleFactor, n,
                                         C.cast(w, C.c_void_p), C.cast(h, C.c_void_p),
                                         C.cast(nf, C.c_void_p), C.cast(sc, C.c_void_p)))
        return dict(w=list(w), h=list(h), nfeatures=list(nf), scale=[np.float32(s) for s in sc])

    def detect_batch(self, images, cap=None):
        """images: uint8 [F,H,W] gray or [F,H,W,3] BGR (numpy or torch) ->
        dict(kp f32 [F,cap,6] = (x, y, size, angle, response, octave), desc u8 [F,cap,32],
        n i32 [F]) on the device.  Raises if more than `cap` key points survive somewhere."""
        lib = _lib.load()
        if not isinstance(images, torch.Tensor):
            images = torch.from_numpy(np.ascontiguousarray(images))
        images = images.to(self.device)
        if images.dtype != U8:
            raise TypeError("OrbDetector: uint8 images required")
        if images.dim() == 4:
            images = bgr_to_gray(images)
        if images.dim() != 3:
            raise ValueError("OrbDetector: images must be [F,H,W] or [F,H,W,3]")
        images = images.contiguous()
        F, H, W = images.shape
        cap = int(cap) if cap is not None else self.nfeatures + self.slack
        wsb = lib.vo_orb_workspace_bytes(F, W, H, self.nfeatures, self.scaleFactor, self.nlevels, cap)
        if wsb == 0:
            raise _lib.VoError(_lib.VO_ERR_INVALID_ARG, "OrbDetector: bad geometry")
        if self._ws is None or self._ws.numel() < wsb:
            self._ws = torch.empty(wsb, dtype=U8, device=self.device)
        kp = torch.zeros((F, cap, 6), dtype=F32, device=self.device)
        desc = torch.zeros((F, cap, 32), dtype=U8, device=self.device)
        n = torch.zeros(F, dtype=I32, device=self.device)
        status = torch.zeros(F, dtype=I32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.vo_orb_detect_batch(
                _lib.ptr(images), F, W, H, self.nfeatures, self.scaleFactor, self.nlevels,
                self.edgeThreshold, self.fastThreshold, cap, _lib.ptr(kp), _lib.ptr(desc),
                _lib.ptr(n), _lib.ptr(status), _lib.ptr(self._ws), wsb, _lib.cur_stream()))
        return dict(kp=kp, desc=desc, n=n, status=status, cap=cap)

    def detectAndCompute(self, image, mask=None):
        """cv2 detector interface: (tuple of key points, uint8 [n,32] descriptors)."""
        if mask is not None:
            raise NotImplementedError("OrbDetector: masks are not supported")
        img = np.asarray(image)
        out = self.detect_batch(img[None])
        if int(out["status"][0].item()) != 0:   # ties at a cut overflowed the slack: retry larger
            out = self.detect_batch(img[None], cap=4 * out["cap"])
            if int(out["status"][0].item()) != 0:
                raise _lib.VoError(_lib.VO_ERR_WORKSPACE, "OrbDetector: key-point capacity exceeded")
        n = int(out["n"][0].item())
        kp = out["kp"][0, :n].cpu().numpy()
        desc = out["desc"][0, :n].cpu().numpy()
        try:
            import cv2
            kps = tuple(cv2.KeyPoint(float(r[0]), float(r[1]), float(r[2]), float(r[3]), float(r[4]),
                                     int(r[5]), -1) for r in kp)
        except ImportError:
            kps = []
            for r in kp:
                k = KeyPointLite(r[0], r[1], float(r[2]))
                k.angle, k.response, k.octave = float(r[3]), float(r[4]), int(r[5])
                kps.append(k)
            kps = tuple(kps)
        return kps, desc


def fast_corners(images, threshold=20, edge=0, device="cuda"):
    """FAST-9/16 stage alone (cv2.FastFeatureDetector_create(threshold, True)) on uint8 [F,H,W]:
    -> (scores u8 [F,H,W] = response + 1 before suppression, list of int32 [n,3] = (x, y,
    response) per frame in (y, x) order)."""
    lib = _lib.load()
    dev = torch.device(device)
    images = torch.as_tensor(np.ascontiguousarray(images)).to(dev).contiguous()
    F, H, W = images.shape
    cap = (W // 2 + 1) * (H // 2 + 1)
    scores = torch.empty((F, H, W), dtype=U8, device=dev)
    corners = torch.zeros((F, cap), dtype=torch.int64, device=dev).to(torch.int32)
    n = torch.zeros(F, dtype=I32, device=dev)
    hist = torch.zeros((F, 256), dtype=I32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.vo_orb_fast_score_map(_lib.ptr(images), F, W, H, int(threshold), int(edge),
                                             _lib.ptr(scores), _lib.ptr(corners), _lib.ptr(n),
                                             _lib.ptr(hist), _lib.cur_stream()))
    out = []
    for f in range(F):
        c = corners[f, :int(n[f].item())].cpu().numpy().view(np.uint32)
        c = np.sort(c)
        out.append(np.stack([(c >> 8) & 0xfff, c >> 20, (c & 255).astype(np.int64) - 1], 1).astype(np.int32))
    return scores, out
