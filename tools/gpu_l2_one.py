"""One L2 matching workload for ncu captures: 200 pairs x 2000^2 integer SIFT, mode and debug flags
from the command line: python tools/gpu_l2_one.py [mode 0|1] [vo_debug_l2 flags] [reps]"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from visual_odometry_b200 import _lib
from visual_odometry_b200.synth import make_sequence

mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
lib = _lib.load()
lib.vo_debug_l2.argtypes = [ctypes.c_int]
lib.vo_debug_l2(flags)
N, B = 2000, 200
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
for _ in range(2): run()
torch.cuda.synchronize()
lib.vo_profile_enable(1)
for _ in range(reps): run()
torch.cuda.synchronize()
ms, n = ctypes.c_double(0), ctypes.c_int(0)
lib.vo_profile_read(2, ctypes.byref(ms), ctypes.byref(n)); lib.vo_profile_enable(0)
k = ms.value / max(n.value, 1)
print(json.dumps({"mode": mode, "flags": flags, "kernel_ms": k, "tflops_algorithmic": 2.0 * B * N * N * 128 / (k * 1e-3) / 1e12,
                  "n_out": float(n_out.float().mean())}))
