"""Scratch GPU check: K1 vs cv2.BFMatcher on random + planted-tie descriptors, plus raw timing."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, cv2
from visual_odometry_b200 import _lib
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING

rng = np.random.default_rng(1)
ok = True
for n1, n2 in [(1000, 1000), (2000, 1777), (37, 4000), (1, 1), (513, 129), (4000, 4000)]:
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    k = min(n1, n2) // 2
    d1[:k] = d2[rng.permutation(n2)[:k]]
    flips = rng.random((k, 256)) < 20 / 256
    d1[:k] ^= np.packbits(flips, axis=1)
    if n1 > 10 and n2 > 10:   # exact duplicates -> ties
        d1[5] = d1[3]; d2[7] = d2[2]; d1[9] = d2[2]
    for cc in (True, False):
        ref = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=cc).match(d1, d2)
        got = B200Matcher(NORM_HAMMING, crossCheck=cc).match(d1, d2)
        a = [(m.queryIdx, m.trainIdx, m.distance) for m in ref]
        b = [(m.queryIdx, m.trainIdx, m.distance) for m in got]
        same = a == b
        ok &= same
        print(n1, n2, "crossCheck", cc, len(a), len(b), "OK" if same else "MISMATCH")
    refk = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2)
    gotk = B200Matcher(NORM_HAMMING).knnMatch(d1, d2, k=2)
    a = [[(m.queryIdx, m.trainIdx, m.distance) for m in r] for r in refk]
    b = [[(m.queryIdx, m.trainIdx, m.distance) for m in r] for r in gotk]
    same = a == b
    ok &= same
    print(n1, n2, "knn2", "OK" if same else "MISMATCH")

# batched timing
lib = _lib.load()
for N, B in [(2000, 999), (4000, 512), (1000, 1)]:
    F = B + 1
    desc = torch.randint(0, 256, (F, N, 32), dtype=torch.uint8, device="cuda")
    nkp = torch.full((F,), N, dtype=torch.int32, device="cuda")
    kp = torch.rand(F, N, 2, device="cuda") * 1000
    pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
    q = torch.empty(B, N, dtype=torch.int32, device="cuda"); t = torch.empty_like(q)
    dist = torch.empty(B, N, dtype=torch.float32, device="cuda")
    pts = torch.empty(B, N, 4, dtype=torch.float64, device="cuda")
    n_out = torch.empty(B, dtype=torch.int32, device="cuda")
    wsb = lib.vo_match_workspace_bytes(B, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    def run():
        _lib.check(lib.vo_match_hamming_batch(_lib.ptr(desc), _lib.ptr(nkp), _lib.ptr(kp), N,
            _lib.ptr(pairs), B, 0, 0.0, _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), None, None,
            _lib.ptr(pts), _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    popc = B * N * N * 8
    print(json.dumps({"N": N, "B": B, "ms": ms, "us_per_pair": ms * 1e3 / B,
                      "popc_per_s": popc / (ms * 1e-3), "n_out_mean": float(n_out.float().mean())}))
print("ALL OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
