"""K1t timing with vo_debug_hamming flags: python tools/gpu_hamming_one.py <flags> [mode]  (999 pairs x 2000 kp)"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from visual_odometry_b200 import _lib
flags = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lib = _lib.load()
N, B = 2000, 999
F = B + 1
torch.manual_seed(0)
desc = torch.randint(0, 256, (F, N, 32), dtype=torch.uint8, device="cuda")
nkp = torch.full((F,), N, dtype=torch.int32, device="cuda")
kp = torch.rand(F, N, 2, device="cuda") * 1000
pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
q = torch.empty(B, N, dtype=torch.int32, device="cuda"); t = torch.empty_like(q)
dist = torch.empty(B, N, dtype=torch.float32, device="cuda")
n_out = torch.empty(B, dtype=torch.int32, device="cuda")
wsb = lib.vo_match_workspace_bytes(F, B, N); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
def run(fl):
    lib.vo_debug_hamming(fl)
    _lib.check(lib.vo_match_hamming_batch(_lib.ptr(desc), _lib.ptr(nkp), _lib.ptr(kp), N, F, _lib.ptr(pairs), B, mode, 0.8,
        _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), None, None, None, _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
run(2); torch.cuda.synchronize()
ref = (q.clone(), t.clone(), dist.clone(), n_out.clone())
for _ in range(2): run(flags)
torch.cuda.synchronize()
same = bool(torch.equal(n_out, ref[3]) and torch.equal(q[:, :500], ref[0][:, :500]) and torch.equal(t[:, :500], ref[1][:, :500]))
lib.vo_profile_enable(1)
for _ in range(5): run(flags)
torch.cuda.synchronize()
ms, n = ctypes.c_double(0), ctypes.c_int(0)
lib.vo_profile_read(0, ctypes.byref(ms), ctypes.byref(n)); lib.vo_profile_enable(0); lib.vo_debug_hamming(0)
print(json.dumps({"flags": flags, "kdbg": flags >> 2, "mode": mode, "kernel_ms": ms.value / max(n.value, 1), "same_as_default": same}))
