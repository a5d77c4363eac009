"""GPU: SIFT extraction of ONE batch of many different frames against cv2 frame by frame: how many key
points / descriptor rows are identical (parity statistics beyond the test scenes, and a large batch).
usage: python tools/gpu_sift_stats.py [n_frames] [H] [W]"""
import json
import os
import sys

import cv2
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visual_odometry_b200 import synth  # noqa: E402
from visual_odometry_b200.sift import SiftDetector  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 128
H = int(sys.argv[2]) if len(sys.argv) > 2 else 480
W = int(sys.argv[3]) if len(sys.argv) > 3 else 640
rng = np.random.default_rng(7)
imgs = []
for i in range(F):
    kind = i % 4
    if kind == 0:
        img = synth.textured_image(H, W, 1000 + i)
    elif kind == 1:      # blurred texture: fewer, larger key points
        img = cv2.GaussianBlur(synth.textured_image(H, W, 2000 + i), (0, 0), 2.5)
    elif kind == 2:      # low contrast + noise
        img = np.clip(synth.textured_image(H, W, 3000 + i) * 0.3 + 90 + rng.normal(0, 3, (H, W)), 0, 255).astype(np.uint8)
    else:                # upscaled coarse pattern: strong edges, few blobs
        img = cv2.resize(rng.integers(0, 256, (H // 16, W // 16), dtype=np.uint8), (W, H), interpolation=cv2.INTER_CUBIC)
    imgs.append(np.ascontiguousarray(img))
imgs = np.stack(imgs)
det = SiftDetector(cap=16384)
out = det.detect_batch(imgs)
torch.cuda.synchronize()
assert int(out["status"].max().item()) == 0
n = out["n"].cpu().numpy()
tot = dict(frames=F, cv2_key_points=0, ours=0, paired=0, identical_records=0, identical_descriptor_rows=0,
           max_abs_xy=0.0, max_abs_angle=0.0, max_descriptor_level_diff=0, frames_with_equal_count=0)
sift = cv2.SIFT_create()
for f in range(F):
    kps, rd = sift.detectAndCompute(imgs[f], None)
    ref = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave) for k in kps], np.float64).reshape(-1, 6)
    rd = np.zeros((0, 128), np.float32) if rd is None else rd
    kp = out["kp"][f, :n[f]].cpu().numpy().astype(np.float64)
    de = out["desc"][f, :n[f]].cpu().numpy()
    tot["cv2_key_points"] += len(ref); tot["ours"] += len(kp)
    tot["frames_with_equal_count"] += int(len(ref) == len(kp))
    if len(ref) == len(kp) and len(ref):
        # same order: compare row by row where the octave word agrees
        same = kp[:, 5] == ref[:, 5]
        d = np.abs(kp[:, :5] - ref[:, :5])
        d[:, 3] = np.minimum(d[:, 3], 360 - d[:, 3])
        ok = same & (d[:, 0] < 0.05) & (d[:, 1] < 0.05) & (d[:, 3] < 0.5)
        tot["paired"] += int(ok.sum())
        tot["identical_records"] += int((ok & (d.max(1) == 0)).sum())
        dd = np.abs(de[ok].astype(np.int32) - rd[ok].astype(np.int32))
        tot["identical_descriptor_rows"] += int((dd.max(1) == 0).sum()) if len(dd) else 0
        if ok.any():
            tot["max_abs_xy"] = max(tot["max_abs_xy"], float(d[ok, :2].max()))
            tot["max_abs_angle"] = max(tot["max_abs_angle"], float(d[ok, 3].max()))
            tot["max_descriptor_level_diff"] = max(tot["max_descriptor_level_diff"], int(dd.max()) if dd.size else 0)
print(json.dumps(tot))
