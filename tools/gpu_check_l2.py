"""Scratch GPU check: K6 L2 tensor-core matching vs cv2.BFMatcher(NORM_L2)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, cv2
from visual_odometry_b200 import _lib
from visual_odometry_b200.matcher import B200Matcher, NORM_L2
from visual_odometry_b200.synth import make_sequence

ok = True
if len(sys.argv) > 1:      # diagnostics: e.g. 8 = MAGIC added by the epilogue instead of the MMA
    import ctypes
    ctypes.CDLL(_lib.LIB_PATH).vo_debug_l2(int(sys.argv[1]))
    print("vo_debug_l2 flags", sys.argv[1])
def tr(ms): return [(m.queryIdx, m.trainIdx, m.distance) for m in ms]
for n, seed in [(100, 1), (400, 2), (1000, 3), (5000, 4)]:
    seq = make_sequence(2, n, seed, descriptor="sift")
    d1, d2 = seq.desc[0], seq.desc[1][: n - n // 7]
    if n > 20:
        d1[5] = d1[3]; d2[7] = d2[2]; d1[9] = d2[2]
    for cc in (True, False):
        ref = cv2.BFMatcher(cv2.NORM_L2, crossCheck=cc).match(d1, d2)
        got = B200Matcher(NORM_L2, crossCheck=cc).match(d1, d2)
        same = tr(ref) == tr(got); ok &= same
        print(n, "crossCheck", cc, len(ref), len(got), "OK" if same else "MISMATCH")
        if not same:
            a, b = tr(ref), tr(got)
            bad = [(x, y) for x, y in zip(a, b) if x != y][:5]; print("  first diffs", bad)
    refk = cv2.BFMatcher(cv2.NORM_L2).knnMatch(d1, d2, k=2)
    gotk = B200Matcher(NORM_L2).knnMatch(d1, d2, k=2)
    same = [tr(r) for r in refk] == [tr(r) for r in gotk]; ok &= same
    print(n, "knn2", "OK" if same else "MISMATCH")
# non-integer descriptors (RootSIFT-like): indices equal, distances within 1e-5 relative
seq = make_sequence(2, 800, 9, descriptor="sift")
r1 = np.sqrt(seq.desc[0] / seq.desc[0].sum(1, keepdims=True)).astype(np.float32)
r2 = np.sqrt(seq.desc[1] / seq.desc[1].sum(1, keepdims=True)).astype(np.float32)
ref = cv2.BFMatcher(cv2.NORM_L2, crossCheck=True).match(r1, r2)
got = B200Matcher(NORM_L2, crossCheck=True).match(r1, r2)
same_idx = [(m.queryIdx, m.trainIdx) for m in ref] == [(m.queryIdx, m.trainIdx) for m in got]
maxrel = max(abs(a.distance - b.distance) / max(a.distance, 1e-9) for a, b in zip(ref, got)) if same_idx else -1
print("rootsift: same indices", same_idx, len(ref), len(got), "max rel dist diff", maxrel)
ok &= same_idx and maxrel < 1e-5
# timing C3: 5000 x 5000 single pair and a batch
lib = _lib.load()
import ctypes
for mode, mname in [(0, "mutual"), (1, "knn2+ratio")]:
  for N, B in [(5000, 1), (5000, 16), (2000, 200)]:
    F = B + 1
    base = make_sequence(2, N, 5, descriptor="sift")
    desc = torch.from_numpy(base.desc).cuda().repeat((F + 1) // 2, 1, 1)[:F].contiguous()
    nkp = torch.full((F,), N, dtype=torch.int32, device="cuda")
    kp = torch.rand(F, N, 2, device="cuda") * 1000
    pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
    q = torch.empty(B, N, dtype=torch.int32, device="cuda"); t = torch.empty_like(q)
    dist = torch.empty(B, N, dtype=torch.float32, device="cuda")
    pts = torch.empty(B, N, 4, dtype=torch.float64, device="cuda")
    n_out = torch.empty(B, dtype=torch.int32, device="cuda")
    wsb = lib.vo_match_l2_workspace_bytes(F, B, N); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    def run():
        _lib.check(lib.vo_match_l2_batch(_lib.ptr(desc), _lib.ptr(nkp), _lib.ptr(kp), N, 128, F, _lib.ptr(pairs), B, mode, 0.75,
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), None, None, _lib.ptr(pts), _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    for _ in range(3): run()
    torch.cuda.synchronize()
    lib.vo_profile_enable(1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): run()
    e1.record(); torch.cuda.synchronize()
    ms_k, nk = ctypes.c_double(0), ctypes.c_int(0); lib.vo_profile_read(2, ctypes.byref(ms_k), ctypes.byref(nk)); lib.vo_profile_enable(0)
    ms = e0.elapsed_time(e1) / 10
    kms = ms_k.value / max(nk.value, 1)
    ndir = 2 if mode == 0 else 1
    alg = 2.0 * B * N * N * 128
    print(json.dumps({"mode": mname, "N": N, "B": B, "ms_total": ms, "gemm_kernel_ms": kms,
                      "tflops_executed": ndir * alg / (kms * 1e-3) / 1e12,
                      "tflops_algorithmic": alg / (kms * 1e-3) / 1e12, "n_out": float(n_out.float().mean())}))
print("ALL OK" if ok else "FAILED"); sys.exit(0 if ok else 1)
