import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING, NORM_L2
from visual_odometry_b200.synth import make_sequence
for desc, norm in (("orb", NORM_HAMMING), ("sift", NORM_L2)):
    seq = make_sequence(2, 2000, 11, descriptor=desc)
    m = B200Matcher(norm, crossCheck=True)
    for _ in range(5): m.match_arrays(seq.desc[0], seq.desc[1])
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50): m.match_arrays(seq.desc[0], seq.desc[1])
    torch.cuda.synchronize(); t1 = time.perf_counter()
    for _ in range(50): m.match_host(seq.desc[0], seq.desc[1])
    t2 = time.perf_counter()
    for _ in range(50): m.match(seq.desc[0], seq.desc[1])
    t3 = time.perf_counter()
    print(desc, "match_arrays ms", (t1 - t0) / 50 * 1e3, "match_host ms", (t2 - t1) / 50 * 1e3, "match ms", (t3 - t2) / 50 * 1e3)
