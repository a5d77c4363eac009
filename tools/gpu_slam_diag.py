"""B200: per-frame differences between DeviceSlam and tests/golden/slam_run.npz (diagnostics)."""
import sys

import numpy as np

sys.path.insert(0, ".")
from visual_odometry_b200.slam import DeviceSlam          # noqa: E402
from visual_odometry_b200.synth import make_sequence      # noqa: E402

z = np.load("tests/golden/slam_run.npz")
g = {k: z[k] for k in z.files}
F, N = int(g["F"]), int(g["N"])
seq = make_sequence(F, N, int(g["seed"]), descriptor="sift", K=g["K"])
slam = DeviceSlam(g["K"], N, F, descriptor="sift", max_cameras=int(g["max_cams"]))
p = slam.pipe
pairs = np.stack([np.arange(F - 1), np.arange(1, F)], 1).astype(np.int32)
p.run_host(seq.desc, seq.kp, seq.nkp, pairs)
for b in range(F - 1):
    info = slam.step(b)
    tag = f"f{b + 1}"
    m = slam.map
    n_cam = len(m.cam_frame)
    n_p, n_o = m.counts()
    poses = m.poses[:n_cam].cpu().numpy()
    ref = g[tag + "_cam_pose"]
    dp = np.abs(poses - ref).max(1) if poses.shape == ref.shape else None
    fid = g[tag + "_pt_fid"]
    ref_f = set((fid[:, 0] * N + fid[:, 1]).tolist())
    got_f = set(m.point_fid[:n_p].cpu().numpy().tolist())
    print(tag, "cams", info["cameras"], "pts", n_p, len(fid), "obs", n_o, len(g[tag + "_obs_pt"]),
          "pnp", info["pnp_points"], info["pnp_ok"], "pose diff", None if dp is None else np.round(dp, 8).tolist(),
          "only dev", sorted(got_f - ref_f)[:6], "only ref", sorted(ref_f - got_f)[:6],
          "ba", None if info["ba"] is None else (info["ba"]["iterations"], info["ba"]["chi2_initial"], float(info["ba"]["chi2"][-1])), flush=True)
    print("    golden ba", np.round(g[tag + "_ba_log"], 4).tolist(), "golden pnp", np.round(g[tag + "_pnp_log"], 6).tolist(),
          "dev pnp", info.get("pnp_pose"), flush=True)
