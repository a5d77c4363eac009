"""GPU check of K1f (fp4 tensor-core Hamming, csrc/hamming_f4.cuh).
  probe : raw accumulator of one 128 x 240 block against the numpy value, with a fit when it differs
  check : path 3 (fp4) against path 1 (popc) on ragged batches, both modes, bit for bit
  time  : popc / int8 / fp4 kernels on the bench sizes"""
import sys, os, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200 import _lib

lib = _lib.load()
stage = sys.argv[1] if len(sys.argv) > 1 else "all"
ok = True


def batch(desc, nkp, pairs, mode, flags, ratio=0.8):
    F, N = desc.shape[0], desc.shape[1]
    B = pairs.shape[0]
    kp = torch.zeros(F, N, 2, device="cuda")
    q = torch.full((B, N), -7, dtype=torch.int32, device="cuda"); t = torch.full_like(q, -7)
    t2 = torch.full_like(q, -7)
    dist = torch.zeros(B, N, dtype=torch.float32, device="cuda"); dist2 = torch.zeros_like(dist)
    n_out = torch.zeros(B, dtype=torch.int32, device="cuda")
    wsb = lib.vo_match_workspace_bytes(F, B, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    lib.vo_debug_hamming(flags)
    _lib.check(lib.vo_match_hamming_batch(_lib.ptr(desc), _lib.ptr(nkp), _lib.ptr(kp), N, F, _lib.ptr(pairs), B,
               mode, ratio, _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2), None,
               _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    torch.cuda.synchronize()
    lib.vo_debug_hamming(0)
    return q.cpu(), t.cpu(), dist.cpu(), t2.cpu(), dist2.cpu(), n_out.cpu()


if stage in ("probe", "all"):
    rng = np.random.default_rng(3)
    N = 1000
    d = rng.integers(0, 256, (2, N, 32), dtype=np.uint8)
    d[0, :64] = d[1, rng.permutation(208)[:64]]          # some near neighbours
    d[0, :64, :2] ^= rng.integers(0, 256, (64, 2), dtype=np.uint8)
    desc = torch.from_numpy(d).cuda()
    nkp = torch.tensor([N, N], dtype=torch.int32, device="cuda")
    pairs = torch.tensor([[0, 1]], dtype=torch.int32, device="cuda")
    acc = torch.full((128, 208), -12345.0, dtype=torch.float32, device="cuda")
    wsb = lib.vo_match_workspace_bytes(2, 1, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    _lib.check(lib.vo_debug_hamming_f4_acc(_lib.ptr(desc), _lib.ptr(nkp), N, _lib.ptr(pairs), _lib.ptr(acc),
                                           _lib.ptr(ws), wsb, _lib.cur_stream()))
    torch.cuda.synchronize()
    a = acc.cpu().numpy().astype(np.float64)
    bits = np.unpackbits(d, axis=2).astype(np.int32)
    H = (bits[0, :128, None, :] != bits[1, None, :208, :]).sum(-1)
    dot36 = 36.0 * (256 - 2 * H)
    idx = 0.25 * (207 - np.arange(208))[None, :]
    exp = dot36 + idx
    diff = np.abs(a - exp).max()
    print("probe: max |acc - expected| =", diff, flush=True)
    if diff != 0.0:
        ok = False
        np.set_printoptions(linewidth=200, precision=3, suppress=True)
        print(" acc[0,:8]     ", a[0, :8]); print(" exp[0,:8]     ", exp[0, :8])
        print(" acc[0,200:208]", a[0, 200:]); print(" exp[0,200:208]", exp[0, 200:])
        print(" acc[1,:4], acc[32,:4], acc[127,:4]", a[1, :4], a[32, :4], a[127, :4])
        print(" exp[1,:4], exp[32,:4], exp[127,:4]", exp[1, :4], exp[32, :4], exp[127, :4])
        X = np.stack([dot36.ravel(), np.broadcast_to(idx, a.shape).ravel(), np.ones(a.size)], 1)
        coef, res, *_ = np.linalg.lstsq(X, a.ravel(), rcond=None)
        print(" fit acc ~ a * 36 dot + b * idx + c: a, b, c =", coef, "residual rms",
              float(np.sqrt(np.mean((X @ coef - a.ravel()) ** 2))))
        print(" untouched (-12345) entries:", int((a == -12345.0).sum()), " zeros:", int((a == 0).sum()),
              " non-finite:", int((~np.isfinite(a)).sum()))
        # per K group: which of the 8 bit groups contribute (correlation with the per-group dots)
        for g in range(8):
            Hg = (bits[0, :128, None, 32 * g:32 * g + 32] != bits[1, None, :208, 32 * g:32 * g + 32]).sum(-1)
            dg = 36.0 * (32 - 2 * Hg)
            print("  corr(acc - idx, dot of K group %d) = %.3f" % (g, np.corrcoef((a - idx).ravel(), dg.ravel())[0, 1]))

if stage in ("check", "all"):
    rng = np.random.default_rng(5)
    cases = [(2000, 12, None), (1777, 9, "ragged"), (4000, 6, None), (600, 40, "ragged"), (520, 300, None)]
    for N, B, kind in cases:
        F = B + 1
        d = rng.integers(0, 256, (F, N, 32), dtype=np.uint8)
        for f in range(1, F):          # neighbours with a few flipped bits, some exact duplicates (ties)
            k = N // 2
            src = rng.permutation(N)[:k]
            d[f, :k] = d[f - 1, src]
            flips = rng.random((k, 256)) < 12 / 256
            d[f, :k] ^= np.packbits(flips, axis=1)
            d[f, 5] = d[f, 3]; d[f, N - 1] = d[f, 3]
        nk = np.full(F, N, np.int32)
        if kind == "ragged":
            special = [0, 1, 103, 104, 105, 207, 208, 209, 255, 256, 257, 415, 416, 417, N - 1]
            for i in range(F):
                nk[i] = min(N, special[i % len(special)]) if i % 2 else rng.integers(N // 2, N + 1)
        desc = torch.from_numpy(d).cuda(); nkp = torch.from_numpy(nk).cuda()
        pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
        for mode in (0, 1):
            ref = batch(desc, nkp, pairs, mode, 1)
            got = batch(desc, nkp, pairs, mode, 3)
            same = bool(torch.equal(ref[5], got[5]))
            bad = []
            for p in range(B):
                n = int(ref[5][p])
                e = all(torch.equal(ref[i][p, :n], got[i][p, :n]) for i in (range(5) if mode else range(3))) and int(got[5][p]) == n   # (the second neighbour is a knn-mode output)
                if not e:
                    bad.append(p)
            same &= not bad
            ok &= same
            print("check N", N, "B", B, kind, "mode", mode, "matches", int(ref[5].sum()), "OK" if same else "MISMATCH pairs %s" % bad[:8], flush=True)
            if bad:
                p = bad[0]
                n = min(int(ref[5][p]), int(got[5][p]))
                print("  pair", p, "nkp", nk[p], nk[p + 1], "n_out popc / fp4", int(ref[5][p]), int(got[5][p]))
                for i in range(n):
                    r = [int(ref[0][p, i]), int(ref[1][p, i]), float(ref[2][p, i]), int(ref[3][p, i]), float(ref[4][p, i])]
                    g = [int(got[0][p, i]), int(got[1][p, i]), float(got[2][p, i]), int(got[3][p, i]), float(got[4][p, i])]
                    if r != g:
                        print("  first difference at", i, "popc", r, "fp4", g)
                        break

if stage in ("time", "all"):
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
        for label, flags, mode in [("popc", 1, 0), ("int8", 2, 0), ("fp4", 3, 0), ("int8_knn2", 2, 1), ("fp4_knn2", 3, 1),
                                   ("fp4_noload", 3 | (1 << 3), 0), ("fp4_norank", 3 | (2 << 3), 0)]:
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
            res[label] = (q.clone(), t.clone(), dist.clone(), n_out.clone())
            print(json.dumps({"N": N, "B": B, "path": label, "ms_total": ms, "ms_kernel": kern,
                              "ms_expand": x_ms.value / max(x_n.value, 1),
                              "algorithmic_tops_2N1N2x256": 2.0 * B * N * N * 256 / (kern * 1e-3) / 1e12,
                              "n_out_mean": float(n_out.float().mean())}), flush=True)
        lib.vo_debug_hamming(0)
        for x, y in (("popc", "fp4"), ("int8_knn2", "fp4_knn2")):
            a, b = res[x], res[y]
            same = bool(torch.equal(a[3], b[3]))
            for p in range(B):
                n = int(a[3][p])
                same &= bool(torch.equal(a[0][p, :n], b[0][p, :n]) and torch.equal(a[1][p, :n], b[1][p, :n])
                             and torch.equal(a[2][p, :n], b[2][p, :n]))
            ok &= same
            print(N, B, y, "==", x, ":", "OK" if same else "MISMATCH", flush=True)
print("ALL OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
