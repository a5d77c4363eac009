"""GPU: K3 (five-point solve) time per step of config C2 for several residency targets
(vo_debug_ransac): python tools/gpu_k3_minb.py 4 5 6"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visual_odometry_b200 import _lib  # noqa: E402
from visual_odometry_b200.pipeline import PairPipeline  # noqa: E402
from visual_odometry_b200.synth import make_sequence  # noqa: E402

lib = _lib.load()
lib.vo_debug_ransac.argtypes = [C.c_int]
seq = make_sequence(1000, 2000, 20261105)
pipe = PairPipeline(2000, 999, 1000, K=seq.K, device="cuda:0")
pipe.stage_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
pin = pipe._pinned()
pipe.desc.copy_(pin["desc"]); pipe.kp.copy_(pin["kp"]); pipe.nkp.copy_(pin["nkp"]); pipe.pairs.copy_(pin["pairs"])
ref = None
for minb in [int(v) for v in sys.argv[1:]] or [4]:
    lib.vo_debug_ransac(minb)
    for _ in range(3):
        pipe.run_device(999)
    torch.cuda.synchronize()
    lib.vo_profile_enable(2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        pipe.run_device(999)
    e1.record()
    torch.cuda.synchronize()
    tot, cnt = C.c_double(), C.c_int()
    lib.vo_profile_read(3, C.byref(tot), C.byref(cnt))
    lib.vo_profile_enable(0)
    R = pipe.R[:999].clone()
    same = True if ref is None else bool(torch.equal(R, ref))
    ref = R if ref is None else ref
    print(f"minb {minb}: K3 {tot.value / 5:.3f} ms per step, step {e0.elapsed_time(e1) / 5:.3f} ms, poses identical to the first setting: {same}", flush=True)
lib.vo_debug_ransac(4)
