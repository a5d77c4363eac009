"""GPU: one batch of SIFT extraction (for an ncu launch list / timing): python tools/gpu_sift_bench.py H W F reps"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visual_odometry_b200 import synth  # noqa: E402
from visual_odometry_b200.sift import SiftDetector  # noqa: E402

h, w, F, reps = (int(v) for v in sys.argv[1:5])
imgs = torch.from_numpy(np.stack([synth.textured_image(h, w, 100 + i) for i in range(F)])).cuda()
det = SiftDetector()
det.detect_batch(imgs)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    out = det.detect_batch(imgs)
e1.record()
torch.cuda.synchronize()
print(f"{F} x {h}x{w}: {e0.elapsed_time(e1) / reps:.3f} ms per batch, key points/frame {out['n'].float().mean().item():.0f}")
