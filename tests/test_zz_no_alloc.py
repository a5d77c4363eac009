"""GPU: the no-allocation entry point vo_match_hamming_static (scratch in a file-scope device array
of libvo_b200, include/vo_b200.h) against the existing vo_match_hamming (caller workspace), both
through the C-ABI: bit-exact (integer / index work), on two calls in a row with DIFFERENT inputs —
the second call finds the first call's keys in the shared array and must not be affected by them.
Collected last on purpose."""
import numpy as np
import pytest
import torch

from visual_odometry_b200 import _lib

pytestmark = pytest.mark.gpu


def _call(lib, dev, d1, d2, mode, ratio, static):
    n1, n2 = len(d1), len(d2)
    n = max(n1, n2, 1)
    a = torch.from_numpy(d1).to(dev)
    b = torch.from_numpy(d2).to(dev)
    q = torch.full((n,), -7, dtype=torch.int32, device=dev)
    t, t2 = torch.full_like(q, -7), torch.full_like(q, -7)
    dist = torch.full((n,), -1.0, dtype=torch.float32, device=dev)
    dist2 = torch.full_like(dist, -1.0)
    n_out = torch.full((1,), -1, dtype=torch.int32, device=dev)
    if static:
        rc = lib.vo_match_hamming_static(_lib.ptr(a), n1, _lib.ptr(b), n2, mode, ratio, _lib.ptr(q),
                                         _lib.ptr(t), _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2),
                                         _lib.ptr(n_out), _lib.cur_stream())
    else:
        wsb = lib.vo_match_workspace_bytes(0, 1, n)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        rc = lib.vo_match_hamming(_lib.ptr(a), n1, _lib.ptr(b), n2, mode, ratio, _lib.ptr(q),
                                  _lib.ptr(t), _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2),
                                  _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream())
    _lib.check(rc)
    torch.cuda.synchronize()
    k = int(n_out.item())
    return k, [x[:k].cpu().numpy() for x in (q, t, dist, t2, dist2)]


def _descriptors(rng, n1, n2):
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    k = min(n1, n2) // 2
    d1[:k] = d2[rng.permutation(n2)[:k]]
    d1[:k] ^= np.packbits(rng.random((k, 256)) < 20 / 256, axis=1)
    if n1 > 10 and n2 > 10:          # exact duplicates: ties in both directions
        d1[5] = d1[3]
        d2[7] = d2[2]
        d1[9] = d2[2]
    return d1, d2


def test_static_scratch_entry_equals_workspace_entry_twice_in_a_row(cuda_dev):
    lib = _lib.load()
    assert lib.vo_match_hamming_static_max_keypoints() == 8192
    rng = np.random.default_rng(2026)
    # two calls in a row per mode, the second SMALLER than the first (stale keys of the first call
    # lie beyond and inside the rows the second one uses), then the largest size the array allows
    for sizes in ([(3000, 2777), (700, 1500)], [(8192, 8192), (129, 64)]):
        for mode, ratio in ((_lib.MATCH_MUTUAL, 0.0), (_lib.MATCH_KNN2_RATIO, 0.8)):
            for n1, n2 in sizes:
                d1, d2 = _descriptors(rng, n1, n2)
                k_s, out_s = _call(lib, cuda_dev, d1, d2, mode, ratio, static=True)
                k_w, out_w = _call(lib, cuda_dev, d1, d2, mode, ratio, static=False)
                assert k_s == k_w and k_s > 0
                for a, b in zip(out_s, out_w):
                    assert np.array_equal(a, b)          # bit-exact: index and integer work
    # the same input twice in a row gives the same answer (nothing accumulates in the array)
    d1, d2 = _descriptors(rng, 1000, 1000)
    first = _call(lib, cuda_dev, d1, d2, _lib.MATCH_MUTUAL, 0.0, static=True)
    second = _call(lib, cuda_dev, d1, d2, _lib.MATCH_MUTUAL, 0.0, static=True)
    assert first[0] == second[0] and all(np.array_equal(a, b) for a, b in zip(first[1], second[1]))


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum(cuda_dev):
    lib = _lib.load()
    big = np.zeros((8193, 32), np.uint8)
    small = np.zeros((10, 32), np.uint8)
    with pytest.raises(_lib.VoError) as e:
        _call(lib, cuda_dev, big, small, _lib.MATCH_MUTUAL, 0.0, static=True)
    assert e.value.code == _lib.VO_ERR_WORKSPACE and "8192" in str(e.value)
    # empty inputs: zero matches, like the workspace entry
    k, _ = _call(lib, cuda_dev, np.zeros((0, 32), np.uint8), small, _lib.MATCH_MUTUAL, 0.0, static=True)
    assert k == 0
