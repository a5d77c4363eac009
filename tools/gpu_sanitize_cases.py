"""Small cases of the kernels that hand-roll mbarrier / bulk-copy / TMEM synchronisation, for
`compute-sanitizer --tool memcheck|racecheck|synccheck python tools/gpu_sanitize_cases.py`
(VERDICT r1 #8).  Every case is checked against cv2, so a sanitizer run is also a parity run:
  K1  hamming_top2_kernel   (popc, bulk-copy staged train tiles)
  K1t htc::match_kernel     (tcgen05 kind::i8, TMEM accumulators, mbarrier pipelines)
  K6i l8::match_kernel      (uint8 SIFT, same skeleton + column-constant ring)
  K6  l2_top2_tc_kernel     (bf16, vo_debug_l2(128))
  N4  orb_* kernels         (one 640 x 480 frame)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402
from visual_odometry_b200 import _lib  # noqa: E402
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING, NORM_L2  # noqa: E402
from visual_odometry_b200.synth import make_sequence  # noqa: E402

lib = _lib.load()
ok = True


def tr(ms):
    return [(m.queryIdx, m.trainIdx, m.distance) for m in ms]


# K1 (single pair -> popc kernel)
rng = np.random.default_rng(0)
d1 = rng.integers(0, 256, (700, 32), dtype=np.uint8)
d2 = rng.integers(0, 256, (513, 32), dtype=np.uint8)
d1[:200] = d2[:200]
same = tr(cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d1, d2)) == \
    tr(B200Matcher(NORM_HAMMING, crossCheck=True).match(d1, d2))
print("K1 popc 700 x 513:", "OK" if same else "MISMATCH", flush=True)
ok &= same

# K1t (forced tensor path, 3 pairs x 600 kp, ragged)
seq = make_sequence(4, 600, 3)
nkp = np.array([600, 513, 600, 257], np.int32)
pairs = np.array([[0, 1], [1, 2], [2, 3]], np.int32)
dev = torch.device("cuda")
F, B, N = 4, 3, 600
desc, kp = torch.from_numpy(seq.desc).to(dev), torch.from_numpy(seq.kp).to(dev)
q = torch.empty(B, N, dtype=torch.int32, device=dev)
t = torch.empty_like(q)
dist = torch.empty(B, N, dtype=torch.float32, device=dev)
n_out = torch.empty(B, dtype=torch.int32, device=dev)
wsb = lib.vo_match_workspace_bytes(F, B, N)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
nkp_d, pairs_d = torch.from_numpy(nkp).to(dev), torch.from_numpy(pairs).to(dev)
for mode in (0, 1):
    lib.vo_debug_hamming(2)
    _lib.check(lib.vo_match_hamming_batch(
        _lib.ptr(desc), _lib.ptr(nkp_d), _lib.ptr(kp), N, F,
        _lib.ptr(pairs_d), B, mode, -1.0, _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist),
        None, None, None, _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    torch.cuda.synchronize()
    lib.vo_debug_hamming(0)
    for b, (fa, fb) in enumerate(pairs):
        ref = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=(mode == 0)).match(seq.desc[fa, :nkp[fa]], seq.desc[fb, :nkp[fb]])
        n = int(n_out[b])
        same = n == len(ref) and q[b, :n].tolist() == [m.queryIdx for m in ref] and \
            t[b, :n].tolist() == [m.trainIdx for m in ref] and dist[b, :n].tolist() == [m.distance for m in ref]
        ok &= same
        print("K1t tcgen05 pair", b, "mode", mode, "OK" if same else "MISMATCH", flush=True)

# K6i (uint8 SIFT) and K6 (bf16)
sq = make_sequence(2, 600, 5, descriptor="sift")
a, b_ = sq.desc[0][:600], sq.desc[1][:513]
lib.vo_debug_l2.argtypes = [_lib.C.c_int]
for flags, name in ((0, "K6i uint8"), (128, "K6 bf16")):
    lib.vo_debug_l2(flags)
    for cc in (True, False):
        same = tr(cv2.BFMatcher(cv2.NORM_L2, crossCheck=cc).match(a, b_)) == tr(B200Matcher(NORM_L2, crossCheck=cc).match(a, b_))
        ok &= same
        print(name, "crossCheck", cc, "OK" if same else "MISMATCH", flush=True)
    same = [tr(r) for r in cv2.BFMatcher(cv2.NORM_L2).knnMatch(a, b_, k=2)] == [tr(r) for r in B200Matcher(NORM_L2).knnMatch(a, b_, k=2)]
    ok &= same
    print(name, "knn2", "OK" if same else "MISMATCH", flush=True)
lib.vo_debug_l2(0)

# N4 ORB: one 640 x 480 frame
from visual_odometry_b200.orb import OrbDetector  # noqa: E402
img = (rng.random((480, 640)) * 255).astype(np.uint8)
img = cv2.GaussianBlur(img, (5, 5), 1.5)
kps, des = OrbDetector(300).detectAndCompute(img, None)
kr, dr = cv2.ORB_create(300).detectAndCompute(img, None)
same = len(kps) == len(kr) and sorted(map(bytes, des)) == sorted(map(bytes, dr))
ok &= same
print("N4 ORB 640 x 480:", "OK" if same else "MISMATCH", len(kps), len(kr), flush=True)
print("ALL OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
