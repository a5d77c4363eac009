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

# batched timing: popc kernel (path 1) vs tcgen05 int8 kernel (path 2; + timing variants of its epilogue)
import ctypes
lib = _lib.load()
for N, B in [(2000, 999), (4000, 512), (1000, 64)]:
    F = B + 1
    desc = torch.randint(0, 256, (F, N, 32), dtype=torch.uint8, device="cuda")
    nkp = torch.full((F,), N, dtype=torch.int32, device="cuda")
    kp = torch.rand(F, N, 2, device="cuda") * 1000
    pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
    q = torch.empty(B, N, dtype=torch.int32, device="cuda"); t = torch.empty_like(q)
    dist = torch.empty(B, N, dtype=torch.float32, device="cuda")
    pts = torch.empty(B, N, 4, dtype=torch.float64, device="cuda")
    n_out = torch.empty(B, dtype=torch.int32, device="cuda")
    wsb = lib.vo_match_workspace_bytes(F, B, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    res = {}
    for label, flags, mode in [("popc", 1, 0), ("tc", 2, 0), ("tc_knn2", 2, 1), ("tc_norank", 2 | (2 << 2), 0),
                               ("tc_noload", 2 | (6 << 2), 0)]:
        lib.vo_debug_hamming(flags)
        def run():
            _lib.check(lib.vo_match_hamming_batch(_lib.ptr(desc), _lib.ptr(nkp), _lib.ptr(kp), N, F,
                _lib.ptr(pairs), B, mode, 0.8, _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), None, None,
                _lib.ptr(pts), _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
        for _ in range(3): run()
        torch.cuda.synchronize()
        lib.vo_profile_enable(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        k_ms, k_n, x_ms, x_n = ctypes.c_double(0), ctypes.c_int(0), ctypes.c_double(0), ctypes.c_int(0)
        lib.vo_profile_read(0, ctypes.byref(k_ms), ctypes.byref(k_n))
        lib.vo_profile_read(4, ctypes.byref(x_ms), ctypes.byref(x_n))
        lib.vo_profile_enable(0)
        kern = k_ms.value / max(k_n.value, 1)
        ndir = 2 if mode == 0 else 1
        res[label] = (q.clone(), t.clone(), dist.clone(), n_out.clone())
        print(json.dumps({"N": N, "B": B, "path": label, "ms_total": ms, "ms_kernel": kern,
                          "ms_expand": x_ms.value / max(x_n.value, 1), "us_per_pair": ms * 1e3 / B,
                          "algorithmic_popc32_per_s": B * N * N * 8 / (kern * 1e-3),
                          "executed_int8_tops": 2.0 * ndir * B * N * N * 288 / (kern * 1e-3) / 1e12,
                          "n_out_mean": float(n_out.float().mean())}), flush=True)
    lib.vo_debug_hamming(0)
    a, b = res["popc"], res["tc"]
    same = bool(torch.equal(a[3], b[3]))
    for p in range(B):
        n = int(a[3][p])
        same &= bool(torch.equal(a[0][p, :n], b[0][p, :n]) and torch.equal(a[1][p, :n], b[1][p, :n])
                     and torch.equal(a[2][p, :n], b[2][p, :n]))
    ok &= same
    print(N, B, "tensor path == popc path:", "OK" if same else "MISMATCH", flush=True)
print("ALL OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
