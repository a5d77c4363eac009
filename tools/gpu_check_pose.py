"""Scratch GPU check: recoverPose / triangulate / chain vs cv2."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, cv2
from visual_odometry_b200 import geometry as G
from visual_odometry_b200.synth import make_sequence

ok = True
Fr, N = 31, 2000
seq = make_sequence(Fr, N, 9); K = seq.K; B = Fr - 1
pts = np.zeros((B, N, 4)); m = np.zeros(B, np.int32)
for b in range(B):
    mt = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(seq.desc[b], seq.desc[b + 1])
    q = np.array([x.queryIdx for x in mt]); t = np.array([x.trainIdx for x in mt]); m[b] = len(q)
    pts[b, :len(q), :2] = seq.kp[b][q]; pts[b, :len(q), 2:] = seq.kp[b + 1][t]
ptsd = torch.from_numpy(pts).cuda(); md = torch.from_numpy(m).cuda(); Kd = torch.from_numpy(K).cuda()
out = G.find_essential_batch(ptsd, md, Kd)
pose = G.recover_pose_batch(out["E"], ptsd, md, out["mask"], Kd, want_mask=True)
X, nx = G.triangulate_pairs(pose["R"], pose["t"], ptsd, md, out["mask"], Kd)
E = out["E"].cpu().numpy(); M = out["mask"].cpu().numpy()
Rg = pose["R"].cpu().numpy(); tg = pose["t"].cpu().numpy(); ng = pose["n_good"].cpu().numpy()
pm = pose["pose_mask"].cpu().numpy(); Xg = X.cpu().numpy(); nx = nx.cpu().numpy()
worstR = worstt = worstX = 0; badmask = 0
for b in range(B):
    inl = M[b, :m[b]] == 1
    p1 = pts[b, :m[b], :2][inl]; p2 = pts[b, :m[b], 2:][inl]
    n, R, t, mk = cv2.recoverPose(E[b], p1, p2, K)
    worstR = max(worstR, np.abs(R - Rg[b]).max()); worstt = max(worstt, np.abs(t.ravel() - tg[b]).max())
    dm = int((mk.ravel() != pm[b, :m[b]][inl]).sum()); badmask += dm
    if n != ng[b] or dm: print("pair", b, "n_good", n, ng[b], "maskdiff", dm)
    ok &= n == ng[b]
    # reference reconstruct_3d_points
    Pf = K @ np.hstack([R.T, -R.T @ t]); P0 = K @ np.eye(3, 4)
    Xc = cv2.triangulatePoints(Pf, P0, p1.T, p2.T); Xc /= Xc[3]
    assert nx[b] == inl.sum()
    rel = np.abs(Xc.T - Xg[b, :nx[b]]).max(1) / np.maximum(1e-12, np.abs(Xc.T).max(1))
    worstX = max(worstX, float(np.median(rel))); 
    if (rel > 1e-4).sum(): print("pair", b, "points rel>1e-4:", int((rel > 1e-4).sum()), "max", rel.max())
print("recoverPose worst |dR| %.2e |dt| %.2e pose-mask diffs %d ; triangulate median-rel worst %.2e" % (worstR, worstt, badmask, worstX))
ok &= worstR < 1e-9 and worstt < 1e-9
# generic triangulate
P1 = np.hstack([cv2.Rodrigues(np.array([0.01, 0.02, -0.01]))[0], [[0.4], [0.1], [-0.3]]]); P2 = np.eye(3, 4)
x1 = np.random.default_rng(0).normal(0, 0.3, (1000, 2)); x2 = x1 + 0.01
Xc = cv2.triangulatePoints(P1, P2, x1.T, x2.T).T
Xg = G.triangulate(torch.from_numpy(P1).cuda(), torch.from_numpy(P2).cuda(), torch.from_numpy(x1).cuda(), torch.from_numpy(x2).cuda()).cpu().numpy()
sg = np.sign((Xc * Xg).sum(1, keepdims=True))
d = np.abs(Xc - sg * Xg).max(); print("triangulate generic max diff %.2e" % d); ok &= d < 1e-8
# chain
poses = G.chain_poses(pose["R"], pose["t"]).cpu().numpy()
P = np.eye(4); worst = 0
for b in range(B):
    rel = np.eye(4); rel[:3, :3] = Rg[b]; rel[:3, 3] = tg[b]; P = rel @ P
    worst = max(worst, np.abs(P - poses[b + 1]).max())
print("chain worst %.2e" % worst); ok &= worst < 1e-12
# timing
rep = (999 + B - 1) // B
ptsb = torch.from_numpy(np.tile(pts, (rep, 1, 1))[:999]).cuda().contiguous(); mb = torch.from_numpy(np.tile(m, rep)[:999]).cuda()
o = G.find_essential_batch(ptsb, mb, Kd)
def timeit(fn, n=3):
    for _ in range(2): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): r = fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n, r
tp, po = timeit(lambda: G.recover_pose_batch(o["E"], ptsb, mb, o["mask"], Kd))
tt, _ = timeit(lambda: G.triangulate_pairs(po["R"], po["t"], ptsb, mb, o["mask"], Kd))
tc, _ = timeit(lambda: G.chain_poses(po["R"], po["t"]))
print(json.dumps({"recover_pose_999_ms": tp, "triangulate_999_ms": tt, "chain_999_ms": tc}))
print("ALL OK" if ok else "FAILED"); sys.exit(0 if ok else 1)
