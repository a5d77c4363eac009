"""N4: SIFT feature extraction on the GPU behind the detector interface the reference
uses: `cv2.SIFT_create()` (src/visual_slam.py:17, src/image_and_keypoints.py:12) and
`self.detector.detectAndCompute(frame.image, None)` (src/frame_generator.py:25,
src/image_and_keypoints.py:46).

`SiftDetector(nfeatures=0)` mirrors `cv2.SIFT_create(nfeatures)` with its default nOctaveLayers 3,
contrastThreshold 0.04, edgeThreshold 10, sigma 1.6; pass it to FrameGenerator
in place of the cv2 detector.  `detect_batch` takes all frames of a sequence at once and leaves
key points and descriptors on the device in the layout the L2 matcher reads (desc uint8
[F, cap, 128] — cv2's float32 descriptors hold exactly these integers —, kp float32 [F, cap, 6]).
All arithmetic runs in libvo_b200 (csrc/sift.cu); there is no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .frame import KeyPointLite
from .orb import bgr_to_gray

U8, I32, F32 = torch.uint8, torch.int32, torch.float32


class SiftDetector:
    def __init__(self, nfeatures=0, nOctaveLayers=3, contrastThreshold=0.04, edgeThreshold=10,
                 sigma=1.6, device="cuda", cap=8192):
        if (int(nOctaveLayers), float(contrastThreshold), float(edgeThreshold),
                float(sigma)) != (3, 0.04, 10.0, 1.6) or int(nfeatures) < 0:
            raise NotImplementedError("SiftDetector: cv2.SIFT_create()'s default nOctaveLayers, "
                                      "thresholds and sigma only (the reference uses no others)")
        self.nfeatures = int(nfeatures)       # 0 = all; n > 0 = retainBest (response >= n-th largest)
        self.device = torch.device(device)
        self.cap = int(cap)
        self._ws = None

    @staticmethod
    def octaves(width, height):
        """[(w, h)] of the pyramid's octaves (octave 0 = the doubled image)"""
        lib = _lib.load()
        w, h = (C.c_int32 * 16)(), (C.c_int32 * 16)()
        n = lib.vo_sift_octaves(int(width), int(height), C.cast(w, C.c_void_p), C.cast(h, C.c_void_p))
        return [(w[i], h[i]) for i in range(n)]

    def detect_batch(self, images, cap=None, with_pyramid=False):
        """images: uint8 [F,H,W] gray or [F,H,W,3] BGR (numpy or torch) ->
        dict(kp f32 [F,cap,6] = (x, y, size, angle, response, octave word), desc u8 [F,cap,128],
        n i32 [F], status i32 [F] (1 = more than cap key points in that frame)) on the device."""
        lib = _lib.load()
        if not isinstance(images, torch.Tensor):
            images = torch.from_numpy(np.ascontiguousarray(images))
        images = images.to(self.device)
        if images.dtype != U8:
            raise TypeError("SiftDetector: uint8 images required")
        if images.dim() == 4:
            images = bgr_to_gray(images)
        if images.dim() != 3:
            raise ValueError("SiftDetector: images must be [F,H,W] or [F,H,W,3]")
        images = images.contiguous()
        F, H, W = images.shape
        cap = int(cap) if cap is not None else self.cap
        wsb = lib.vo_sift_workspace_bytes(F, W, H, cap)
        if wsb == 0:
            raise _lib.VoError(_lib.VO_ERR_INVALID_ARG, "SiftDetector: bad geometry")
        if self._ws is None or self._ws.numel() < wsb:
            self._ws = None
            self._ws = torch.empty(wsb, dtype=U8, device=self.device)
        kp = torch.zeros((F, cap, 6), dtype=F32, device=self.device)
        desc = torch.zeros((F, cap, 128), dtype=U8, device=self.device)
        n = torch.zeros(F, dtype=I32, device=self.device)
        status = torch.zeros(F, dtype=I32, device=self.device)
        pyr = None
        if with_pyramid:
            floats = sum(6 * w * h for w, h in self.octaves(W, H))
            pyr = torch.empty((F, floats), dtype=F32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.vo_sift_detect_batch(
                _lib.ptr(images), F, W, H, cap, self.nfeatures, _lib.ptr(kp), _lib.ptr(desc), _lib.ptr(n),
                _lib.ptr(status), _lib.ptr(pyr), _lib.ptr(self._ws), wsb, _lib.cur_stream()))
        out = dict(kp=kp, desc=desc, n=n, status=status, cap=cap)
        if with_pyramid:
            out["pyramid"] = pyr
        return out

    def detectAndCompute(self, image, mask=None):
        """cv2 detector interface: (tuple of key points, float32 [n,128] descriptors)."""
        if mask is not None:
            raise NotImplementedError("SiftDetector: masks are not supported")
        img = np.asarray(image)
        out = self.detect_batch(img[None])
        while int(out["status"][0].item()) != 0:      # more key points than rows: retry larger
            if out["cap"] >= (1 << 20):
                raise _lib.VoError(_lib.VO_ERR_WORKSPACE, "SiftDetector: key-point capacity exceeded")
            out = self.detect_batch(img[None], cap=4 * out["cap"])
        n = int(out["n"][0].item())
        kp = out["kp"][0, :n].cpu().numpy()
        desc = out["desc"][0, :n].cpu().numpy().astype(np.float32)
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
