"""Scratch: end-to-end step time vs upload chunking (run_staged chunks / first fraction)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200.pipeline import PairPipeline
from visual_odometry_b200.synth import make_sequence
F, N = 1000, 2000
seq = make_sequence(F, N, 20261105)
pipe = PairPipeline(N, F - 1, F, K=seq.K)
pipe.stage_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for chunks, first in [(1, 1.0), (2, 0.15), (2, 0.08), (2, 0.04), (3, 0.05), (3, 0.1), (4, 0.05), (5, 0.04)]:
    for _ in range(2): pipe.run_staged(F, F - 1, chunks=chunks, first=first)
    tot = 0.0
    for _ in range(8):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); pipe.run_staged(F, F - 1, chunks=chunks, first=first); e1.record()
        torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
    print(json.dumps({"chunks": chunks, "first": first, "ms": tot / 8, "pairs_per_s": (F - 1) / (tot / 8 * 1e-3)}))
