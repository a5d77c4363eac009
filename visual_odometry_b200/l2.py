"""SIFT / NORM_L2 matching on the tcgen05 tensor cores (csrc/l2.cu): single-pair wrapper used by
B200Matcher; the batched form is called directly by PairPipeline."""
import numpy as np
import torch

from . import _lib


def _as_f32_cuda(a):
    if isinstance(a, torch.Tensor):
        return a.to(device="cuda", dtype=torch.float32).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def match_l2_single(lib, desc1, desc2, mode, ratio):
    """-> (q, t, dist, t2, dist2) torch tensors trimmed to the number of matches."""
    d1, d2 = _as_f32_cuda(desc1), _as_f32_cuda(desc2)
    n1, n2 = d1.shape[0], d2.shape[0]
    dev = d1.device
    empty_i = torch.empty(0, dtype=torch.int32, device=dev)
    empty_f = torch.empty(0, dtype=torch.float32, device=dev)
    if n1 == 0 or n2 == 0:
        return empty_i, empty_i, empty_f, empty_i, empty_f
    if d1.shape[1] != 128 or d2.shape[1] != 128:
        raise ValueError("L2 descriptors must be float32 [N,128] (SIFT)")
    cap = max(n1, n2)
    desc = torch.zeros(2, cap, 128, dtype=torch.float32, device=dev)
    desc[0, :n1], desc[1, :n2] = d1, d2
    nkp = torch.tensor([n1, n2], dtype=torch.int32, device=dev)
    pairs = torch.tensor([[0, 1]], dtype=torch.int32, device=dev)
    q = torch.empty(cap, dtype=torch.int32, device=dev)
    t, t2 = torch.empty_like(q), torch.empty_like(q)
    dist = torch.empty(cap, dtype=torch.float32, device=dev)
    dist2 = torch.empty_like(dist)
    n_out = torch.zeros(1, dtype=torch.int32, device=dev)
    wsb = lib.vo_match_l2_workspace_bytes(2, 1, cap)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    _lib.check(lib.vo_match_l2_batch(
        _lib.ptr(desc), _lib.ptr(nkp), None, cap, 128, 2, _lib.ptr(pairs), 1, int(mode),
        float(ratio), _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2), None,
        _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
    n = int(n_out.item())
    return q[:n], t[:n], dist[:n], t2[:n], dist2[:n]
