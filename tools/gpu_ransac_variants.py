"""Scratch GPU timing: ransac_solve_kernel register/occupancy variants on the bench workload."""
import sys, os, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200 import _lib
from visual_odometry_b200.pipeline import PairPipeline
from visual_odometry_b200.synth import make_sequence
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH)
F, N = 1000, 2000
seq = make_sequence(F, N, 20261105)
pipe = PairPipeline(N, F - 1, F, K=seq.K)
pipe.stage_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
pipe.run_staged(F, F - 1)
ref = (pipe.R.clone(), pipe.n_inl.clone(), pipe.n_iters.clone())
for minb in (2, 3, 4, 6, 8):
    raw.vo_debug_ransac(minb)
    for _ in range(2): pipe.run_device(F - 1)
    torch.cuda.synchronize(); lib.vo_profile_enable(1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): pipe.run_device(F - 1)
    e1.record(); torch.cuda.synchronize()
    out = {"minb": minb, "step_ms": e0.elapsed_time(e1) / 5}
    for kid, nm in [(0, "hamming"), (1, "score"), (3, "solve")]:
        ms, n = ctypes.c_double(0), ctypes.c_int(0)
        lib.vo_profile_read(kid, ctypes.byref(ms), ctypes.byref(n)); out[nm + "_ms_per_step"] = ms.value / 5
    lib.vo_profile_enable(0)
    out["same_R"] = bool(torch.equal(pipe.R, ref[0])); out["same_inl"] = bool(torch.equal(pipe.n_inl, ref[1]))
    out["mean_iters"] = float(pipe.n_iters.float().mean())
    print(json.dumps(out))
raw.vo_debug_ransac(4)
