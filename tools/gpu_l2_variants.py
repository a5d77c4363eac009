import sys, os, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from visual_odometry_b200 import _lib
from visual_odometry_b200.synth import make_sequence
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH)
for N, B in [(2000, 200), (5000, 16)]:
    F = B + 1
    base = make_sequence(2, N, 5, descriptor="sift")
    desc = torch.from_numpy(base.desc).cuda().repeat((F + 1) // 2, 1, 1)[:F].contiguous()
    nkp = torch.full((F,), N, dtype=torch.int32, device="cuda")
    pairs = torch.stack([torch.arange(B), torch.arange(B) + 1], 1).to(torch.int32).cuda()
    q = torch.empty(B, N, dtype=torch.int32, device="cuda"); t = torch.empty_like(q)
    dist = torch.empty(B, N, dtype=torch.float32, device="cuda")
    n_out = torch.empty(B, dtype=torch.int32, device="cuda")
    wsb = lib.vo_match_l2_workspace_bytes(F, B, N); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    def run():
        _lib.check(lib.vo_match_l2_batch(_lib.ptr(desc), _lib.ptr(nkp), None, N, 128, F, _lib.ptr(pairs), B, 0, 0.0,
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), None, None, None, _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    for flags in (0, 1, 2, 3, 4, 5, 6, 7, 8, 64):   # 64 = fragment-layout epilogue (experimental)
        raw.vo_debug_l2(flags)
        for _ in range(2): run()
        torch.cuda.synchronize(); lib.vo_profile_enable(1)
        for _ in range(5): run()
        torch.cuda.synchronize()
        ms_k, nk = ctypes.c_double(0), ctypes.c_int(0); lib.vo_profile_read(2, ctypes.byref(ms_k), ctypes.byref(nk)); lib.vo_profile_enable(0)
        ms = ms_k.value / nk.value
        print(json.dumps({"N": N, "B": B, "flags": flags, "kernel_ms": ms, "tflops": 2.0 * B * 2 * N * N * 128 / (ms * 1e-3) / 1e12}))
raw.vo_debug_l2(0)
