"""B200: DeviceSlam on a synthetic sequence; per frame: map sizes, PnP, BA, alignment of the camera
centres to the ground truth (similarity transform).  usage: descriptor F N seed [K=ref]"""
import sys

import numpy as np

sys.path.insert(0, ".")
from visual_odometry_b200.slam import DeviceSlam          # noqa: E402
from visual_odometry_b200.synth import K_REFERENCE, K_DEFAULT, make_sequence      # noqa: E402

desc, F, N, seed = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
K = K_REFERENCE if len(sys.argv) > 5 and sys.argv[5] == "ref" else K_DEFAULT
seq = make_sequence(F, N, seed, descriptor=desc, K=K)
slam = DeviceSlam(K, N, F, descriptor=desc)
p = slam.pipe
pairs = np.stack([np.arange(F - 1), np.arange(1, F)], 1).astype(np.int32)
p.run_host(seq.desc, seq.kp, seq.nkp, pairs)
import time
import torch
for b in range(F - 1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = slam.step(b)
    torch.cuda.synchronize()
    info["ms"] = 1e3 * (time.perf_counter() - t0)
    m = slam.map
    n_cam = len(m.cam_frame)
    if n_cam < 3:
        bi = info["ba_init"]
        X = m.points[:info["n_points"]].cpu().numpy()
        print("init: matches", int(p.n_match[0]), "inliers", int(p.mask[0, :int(p.n_match[0])].sum()), "ba",
              (bi["iterations"], round(bi["chi2_initial"], 1), round(float(bi["chi2"][-1]), 1), bi["trials"].tolist()),
              "|t1|", np.linalg.norm(m.poses[1].cpu().numpy().reshape(4, 4)[:3, 3]).round(4),
              "median |X|", np.median(np.linalg.norm(X, axis=1)).round(3), "true |t|",
              np.linalg.norm(seq.relative_pose(0, 1)[1]).round(3), flush=True)
        continue
    poses = m.poses[:n_cam].cpu().numpy().reshape(n_cam, 4, 4)
    C = np.array([-P[:3, :3].T @ P[:3, 3] for P in poses])
    Cg = np.array([-seq.R_w2c[f].T @ seq.t_w2c[f] for f in m.cam_frame])
    A, B = C - C.mean(0), Cg - Cg.mean(0)
    U, S, Vt = np.linalg.svd(A.T @ B)
    D = np.diag([1, 1, np.sign(np.linalg.det(U @ Vt))])
    s = (S * np.diag(D)).sum() / (A ** 2).sum()
    err = np.linalg.norm(s * ((U @ D @ Vt).T @ A.T).T - B, axis=1)
    ba = info["ba"]
    print(b + 1, "cams", n_cam, "pts", info["n_points"], "obs", info["n_obs"], "pnp", info["pnp_points"],
          info.get("pnp_inliers"), info["pnp_ok"], "ba", None if ba is None else
          (ba["iterations"], round(ba["chi2_initial"], 1), round(float(ba["chi2"][-1]), 1)),
          "align err", err.max().round(3), "extent", np.linalg.norm(B, axis=1).max().round(3),
          "scale", round(s, 3), "step ms", round(info["ms"], 2), info.get("error", ""), flush=True)
