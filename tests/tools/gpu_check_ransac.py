"""Scratch GPU check: cv::RNG replay, five-point solver and batched E-RANSAC vs cv2."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, cv2
from visual_odometry_b200 import geometry as G
from visual_odometry_b200.synth import make_sequence
from oracle import cv_restate, cv_geometry

ok = True
# 1. samples
for count in (5, 6, 17, 1500):
    a = G.ransac_samples(count, 200).cpu().numpy()
    b = cv_restate.ransac_samples(count, 200)
    same = np.array_equal(a, b); ok &= same
    print("samples", count, "OK" if same else "MISMATCH")

# 2. five-point vs cv2's own 5-point output
seq = make_sequence(2, 600, 11)
mt = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(seq.desc[0], seq.desc[1])
p1 = np.array([seq.kp[0][m.queryIdx] for m in mt], np.float64)
p2 = np.array([seq.kp[1][m.trainIdx] for m in mt], np.float64)
K = seq.K
ptsn = np.c_[cv_geometry.normalize_points(p1, K), cv_geometry.normalize_points(p2, K)]
rng = np.random.default_rng(0)
S = 400
samples = np.stack([rng.choice(len(p1), 5, replace=False) for _ in range(S)]).astype(np.int32)
E, n = G.fivept_solve(torch.from_numpy(ptsn).cuda(), torch.from_numpy(samples).cuda())
E = E.cpu().numpy(); n = n.cpu().numpy()
bad_cnt = 0; worst = 0.0; nbig = 0
for s in range(S):
    idx = samples[s]
    Ecv, _ = cv2.findEssentialMat(p1[idx], p2[idx], K, cv2.FM_RANSAC, 0.99, 1)
    ncv = 0 if Ecv is None else Ecv.shape[0] // 3
    if ncv != n[s]:
        bad_cnt += 1; continue
    for k in range(ncv):
        e = Ecv[3 * k:3 * k + 3]
        d = min(min(np.abs(e - E[s, j]).max(), np.abs(e + E[s, j]).max()) for j in range(n[s]))
        worst = max(worst, d); nbig += d > 1e-6
print("fivept: model-count mismatches", bad_cnt, "of", S, "worst |dE|", worst, "n>1e-6:", nbig)
ok &= bad_cnt <= 2

# 3. batched E-RANSAC vs cv2 on a sequence
Fr, N = 41, 2000
seq = make_sequence(Fr, N, 5)
B = Fr - 1
pts = np.zeros((B, N, 4)); m = np.zeros(B, np.int32)
cvE = []; cvM = []
tcv = 0
for b in range(B):
    mt = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(seq.desc[b], seq.desc[b + 1])
    q = np.array([x.queryIdx for x in mt]); t = np.array([x.trainIdx for x in mt])
    m[b] = len(q)
    pts[b, :len(q), :2] = seq.kp[b][q]; pts[b, :len(q), 2:] = seq.kp[b + 1][t]
    t0 = time.perf_counter()
    Ecv, mk = cv2.findEssentialMat(pts[b, :m[b], :2], pts[b, :m[b], 2:], K, cv2.FM_RANSAC, 0.99, 1)
    tcv += time.perf_counter() - t0
    cvE.append(Ecv); cvM.append(mk.ravel())
out = G.find_essential_batch(torch.from_numpy(pts).cuda(), torch.from_numpy(m).cuda(),
                             torch.from_numpy(K).cuda())
Eg = out["E"].cpu().numpy(); Mg = out["mask"].cpu().numpy(); ni = out["n_inl"].cpu().numpy()
nit = out["n_iters"].cpu().numpy(); st = out["status"].cpu().numpy()
nsame = 0
for b in range(B):
    dE = min(np.abs(cvE[b] - Eg[b]).max(), np.abs(cvE[b] + Eg[b]).max())
    dm = int((cvM[b] != Mg[b, :m[b]]).sum())
    nsame += (dm == 0)
    if dm or dE > 1e-8:
        print("pair", b, "dE %.2e" % dE, "maskdiff", dm, "inl cv", int(cvM[b].sum()), "gpu", ni[b], "niters", nit[b])
print("E-RANSAC: identical masks on", nsame, "of", B, "pairs; cv2 time/pair %.2f ms" % (tcv / B * 1e3), "status", st.min(), st.max())
ok &= nsame >= B - 2

# 4. timing at C2 scale: replicate pairs to 999
rep = (999 + B - 1) // B
ptsb = torch.from_numpy(np.tile(pts, (rep, 1, 1))[:999]).cuda().contiguous()
mb = torch.from_numpy(np.tile(m, rep)[:999]).cuda().contiguous()
Kd = torch.from_numpy(K).cuda()
for _ in range(2): G.find_essential_batch(ptsb, mb, Kd)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): o = G.find_essential_batch(ptsb, mb, Kd)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(json.dumps({"essential_999_pairs_ms": ms, "us_per_pair": ms * 1e3 / 999,
                  "mean_niters": float(o["n_iters"].float().mean())}))
print("ALL OK" if ok else "FAILED")
sys.exit(0 if ok else 1)
