"""B200: achieved HBM bandwidth of the N2 kernels (reproj_kernel, observation pruning) at a size
far beyond L2 (20 M observations, 2 M points).  Algorithmic bytes per observation:
reproj 24 (indices + pixel) + 24 (point gather) in, 9 out = 57 B; prune 25 in + 24 out x keep rate."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from visual_odometry_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda")
M, P, C = 20_000_000, 2_000_000, 18
g = torch.Generator(device=dev).manual_seed(1)
points = torch.rand(P, 3, dtype=torch.float64, device=dev, generator=g) * 20 + torch.tensor([-10.0, -10, 10], device=dev, dtype=torch.float64)
poses = torch.eye(4, dtype=torch.float64, device=dev).reshape(1, 16).repeat(C, 1).contiguous()
obs_cam_random = torch.randint(0, C, (M,), dtype=torch.int32, device=dev, generator=g)
obs_cam_runs = ((torch.arange(M, device=dev) // 4096) % C).to(torch.int32)   # runs per camera, as a map grows
obs_cam = obs_cam_random
K = torch.tensor([1400.0, 0, 960, 0, 1400, 540, 0, 0, 1], dtype=torch.float64, device=dev)
sqerr = torch.empty(M, dtype=torch.float64, device=dev)
keep = torch.empty(M, dtype=torch.uint8, device=dev)
total = torch.zeros(1, dtype=torch.float64, device=dev)
n_keep = torch.zeros(1, dtype=torch.int32, device=dev)
wsb = max(lib.vo_reprojection_workspace_bytes(M), lib.vo_map_prune_workspace_bytes(M))
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
out_cam, out_pt = torch.empty_like(obs_cam), torch.empty_like(obs_cam)
out_xy = torch.empty(M, 2, dtype=torch.float64, device=dev)
for name, obs_pt, obs_cam in (
        ("random cameras, random points", torch.randint(0, P, (M,), dtype=torch.int32, device=dev, generator=g), obs_cam_random),
        ("camera runs, random points", torch.randint(0, P, (M,), dtype=torch.int32, device=dev, generator=g), obs_cam_runs),
        ("camera runs, points in observation order", (torch.arange(M, device=dev) // 10).to(torch.int32), obs_cam_runs)):
    X = points[obs_pt.long()]
    obs_xy = (torch.stack([X[:, 0] / X[:, 2] * 1400 + 960, X[:, 1] / X[:, 2] * 1400 + 540], 1)
              + torch.randn(M, 2, dtype=torch.float64, device=dev, generator=g)).contiguous()
    del X

    def reproj():
        _lib.check(lib.vo_reprojection_errors(
            _lib.ptr(points), P, _lib.ptr(poses), C, _lib.ptr(obs_cam), _lib.ptr(obs_pt), _lib.ptr(obs_xy), M,
            _lib.ptr(K), 1.0, _lib.ptr(sqerr), _lib.ptr(keep), _lib.ptr(total), _lib.ptr(n_keep), _lib.ptr(ws),
            wsb, _lib.cur_stream()))

    def prune():
        _lib.check(lib.vo_map_prune_observations(
            _lib.ptr(keep), M, _lib.ptr(obs_cam), _lib.ptr(obs_pt), _lib.ptr(obs_xy), _lib.ptr(out_cam),
            _lib.ptr(out_pt), _lib.ptr(out_xy), _lib.ptr(n_keep), _lib.ptr(ws), wsb, _lib.cur_stream()))

    for fn, label in ((reproj, "reproj"), (prune, "prune")):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        kept = int(n_keep.item())
        nbytes = M * 57 if label == "reproj" else M * 25 + kept * 24 + M   # prune reads keep twice
        print(f"{name}: {label} {ms:.3f} ms for {M/1e6:.0f} M observations = {nbytes/ms/1e6:.0f} GB/s algorithmic"
              f" (kept {kept})", flush=True)
