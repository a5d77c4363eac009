"""N4 measurement: ORB extraction (vo_orb_detect_batch) on a batch of 1920x1080 synthetic frames,
CUDA-event timing with the frames resident in HBM, beside cv2.ORB_create(nfeatures) on the host
cores of the same box.  Algorithmic bytes per frame: every pyramid level is written
once and read twice (next level + FAST) = 3 x 6.8 MB - the base level is only read - about 18 MB.
usage: python tools/gpu_orb_bench.py [frames] [nfeatures] [reps]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from visual_odometry_b200 import synth  # noqa: E402
from visual_odometry_b200.orb import OrbDetector  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 32
NF = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
REPS = int(sys.argv[3]) if len(sys.argv) > 3 else 10
H, W = 1080, 1920

base = [synth.textured_image(H, W, 100 + k) for k in range(4)]
imgs = np.stack([np.roll(base[k % 4], 7 * (k // 4), axis=1) for k in range(F)])
dev = torch.device("cuda:0")
d_imgs = torch.from_numpy(imgs).to(dev)
det = OrbDetector(nfeatures=NF)
info = det.level_info(W, H)
level_px = [w * h for w, h in zip(info["w"], info["h"])]
alg_bytes = level_px[0] + 3 * sum(level_px[1:]) + NF * (6 * 4 + 32)
for _ in range(3):
    out = det.detect_batch(d_imgs)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
times = []
for _ in range(REPS):
    flush.fill_(1)                                   # L2 flush between timed runs
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = det.detect_batch(d_imgs)
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = float(np.median(times))
n = out["n"].cpu().numpy()

res = {"workload": "%d frames %dx%d, ORB nfeatures %d, 8 levels" % (F, W, H, NF),
       "gpu_ms_per_batch": ms, "gpu_ms_per_frame": ms / F, "gpu_frames_per_s": F / ms * 1e3,
       "keypoints_per_frame_mean": float(n.mean()),
       "algorithmic_bytes_per_frame": alg_bytes,
       "achieved_gbs": alg_bytes * F / ms / 1e6, "launches_per_batch": 7 + 2 * 8 - 1}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    res["hbm_peak_gbs"] = peaks["hbm_gbs"]
    res["frac_of_hbm_peak"] = res["achieved_gbs"] / peaks["hbm_gbs"]
except Exception:
    pass
try:
    import cv2
    orb = cv2.ORB_create(nfeatures=NF)
    for threads in (1, 0):
        cv2.setNumThreads(threads if threads else -1)
        t = []
        for k in range(min(F, 8)):
            t0 = time.perf_counter()
            orb.detectAndCompute(imgs[k], None)
            t.append(time.perf_counter() - t0)
        res["cv2_ms_per_frame_%s" % ("1thread" if threads == 1 else "allthreads")] = float(np.median(t)) * 1e3
    res["cv2_threads"] = cv2.getNumThreads()
    res["cv2_version"] = cv2.__version__
    res["speedup_vs_cv2_allthreads"] = res["cv2_ms_per_frame_allthreads"] / res["gpu_ms_per_frame"]
except ImportError:
    pass
print(json.dumps(res))
