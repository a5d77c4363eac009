"""Development aid: stage-by-stage comparison of the GPU ORB front end with the oracle."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cv_orb
from visual_odometry_b200 import synth
from visual_odometry_b200.orb import OrbDetector
PATTERN = np.load(os.path.join(ROOT, "tests", "golden", "orb_pattern_31.npy"))
img = synth.textured_image(240, 320, 13)
out = OrbDetector(nfeatures=100000, nlevels=1).detect_batch(img[None], cap=5000)
torch.cuda.synchronize()
n = int(out["n"][0]); kp = out["kp"][0, :n].cpu().numpy(); desc = out["desc"][0, :n].cpu().numpy()
xy, sc = cv_orb.fast9_detect(img, 20, True)
m = cv_orb.run_by_image_border(xy, img.shape, 31); xy = xy[m]
mine = set(map(tuple, kp[:, :2].astype(int).tolist())); ref = set(map(tuple, xy.tolist()))
print("gpu", len(mine), "oracle", len(ref), "common", len(mine & ref))
print("only gpu", sorted(mine - ref)[:10]); print("only oracle", sorted(ref - mine)[:10])
r = cv_orb.orb_detect_and_compute(img, PATTERN, nfeatures=100000, nlevels=1)
if len(r["pt"]) == n:
    for k, name in enumerate(["x", "y", "size"]):
        pass
    print("pt eq", np.array_equal(kp[:, :2], r["pt"]), "angle eq", (kp[:, 3] == r["angle"]).mean(),
          "resp eq", (kp[:, 4] == r["response"]).mean(), "desc eq", (desc == r["desc"]).all(1).mean())
    bad = np.nonzero(kp[:, 3] != r["angle"])[0][:5]
    print(kp[bad, 3], r["angle"][bad])
    bad = np.nonzero(kp[:, 4] != r["response"])[0][:5]
    print(kp[bad, 4], r["response"][bad])
from visual_odometry_b200.orb import fast_corners
scores, cs = fast_corners(img[None], 20, 0)
scores = scores[0].cpu().numpy().astype(int)
ref = cv_orb.fast9_score_map(img, 20); raw = np.where(ref > 0, ref + 1, 0)
print("score map mismatches", (scores != raw).sum(), "of", raw.size, "nonzero", (scores > 0).sum(), (raw > 0).sum())
ys, xs = np.nonzero(scores != raw)
print(list(zip(xs[:12], ys[:12], scores[ys[:12], xs[:12]], raw[ys[:12], xs[:12]])))
xy0, sc0 = cv_orb.fast9_detect(img, 20, True)
print("nms corners gpu", len(cs[0]), "oracle", len(xy0), "equal", len(cs[0]) == len(xy0) and np.array_equal(cs[0][:, :2], xy0) and np.array_equal(cs[0][:, 2], sc0))
