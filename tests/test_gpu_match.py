"""GPU: K1 Hamming matching through the C-ABI vs the oracle / cv2 / golden vectors (bit-exact)."""
import numpy as np
import pytest
import torch

import cv2
from oracle import cv_restate as CR
from visual_odometry_b200 import _lib
from visual_odometry_b200.matcher import B200Matcher, NORM_HAMMING
from visual_odometry_b200.synth import make_sequence
from conftest import load_golden

pytestmark = pytest.mark.gpu


def _triples(ms):
    return [(m.queryIdx, m.trainIdx, m.distance) for m in ms]


@pytest.mark.parametrize("name", ["orb_300", "orb_1000_c1", "orb_700_refK"])
def test_mutual_match_equals_golden(cuda_dev, name):
    g = load_golden(name)
    ms = B200Matcher(NORM_HAMMING, crossCheck=True).match(g["desc"][0], g["desc"][1])
    assert [m.queryIdx for m in ms] == g["q"].tolist()
    assert [m.trainIdx for m in ms] == g["t"].tolist()
    assert [m.distance for m in ms] == g["dist"].tolist()
    assert all(m.imgIdx == 0 for m in ms)


@pytest.mark.parametrize("n1,n2", [(1, 1), (5, 3), (129, 513), (1000, 1000), (2000, 1777),
                                   (37, 4000), (4000, 4000)])
def test_all_modes_equal_cv2(cuda_dev, n1, n2):
    rng = np.random.default_rng(n1 * 7919 + n2)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    k = min(n1, n2) // 2
    if k:
        d1[:k] = d2[rng.permutation(n2)[:k]]
        d1[:k] ^= np.packbits(rng.random((k, 256)) < 20 / 256, axis=1)
    if n1 > 10 and n2 > 10:  # exact duplicates: ties in both directions
        d1[5] = d1[3]
        d2[7] = d2[2]
        d1[9] = d2[2]
    for cc in (True, False):
        ref = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=cc).match(d1, d2)
        got = B200Matcher(NORM_HAMMING, crossCheck=cc).match(d1, d2)
        assert _triples(ref) == _triples(got)
    D = CR.hamming_matrix(d1, d2)
    q, t, d = CR.match_from_matrix(D, True)   # the numpy restatement agrees too
    got = B200Matcher(NORM_HAMMING, crossCheck=True).match(d1, d2)
    assert [m.queryIdx for m in got] == q.tolist() and [m.trainIdx for m in got] == t.tolist()
    if n2 >= 2:
        refk = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2)
        gotk = B200Matcher(NORM_HAMMING).knnMatch(d1, d2, k=2)
        assert [_triples(r) for r in refk] == [_triples(r) for r in gotk]
        for ratio in (0.3, 0.8):
            good = [m for m, n in refk if m.distance < ratio * n.distance]
            mine = B200Matcher(NORM_HAMMING).ratio_match(d1, d2, ratio)
            assert _triples(good) == _triples(mine)


def test_empty_inputs(cuda_dev):
    e = np.zeros((0, 32), np.uint8)
    d = np.random.default_rng(0).integers(0, 256, (10, 32), dtype=np.uint8)
    m = B200Matcher(NORM_HAMMING, crossCheck=True)
    assert m.match(e, d) == [] and m.match(d, e) == [] and m.match(e, e) == []


def test_all_identical_descriptors_tie_break(cuda_dev):
    d = np.tile(np.arange(32, dtype=np.uint8), (50, 1))
    ref = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d, d)
    got = B200Matcher(NORM_HAMMING, crossCheck=True).match(d, d)
    assert _triples(ref) == _triples(got) == [(0, 0, 0.0)]


def _batch(lib, dev, desc, kp, nkp, pairs, N, mode=0, ratio=0.0, path=0):
    """vo_match_hamming_batch on host arrays; path 1 = popc kernel, 2 = int8 tensor-core kernel
    (K1t), 3 = fp4 tensor-core kernel (K1f, what large batches run by default)"""
    B, F = len(pairs), desc.shape[0]
    d_desc, d_kp = torch.from_numpy(desc).to(dev), torch.from_numpy(kp).to(dev)
    q = torch.full((B, N), -7, dtype=torch.int32, device=dev)
    t, t2 = torch.full_like(q, -7), torch.full_like(q, -7)
    dist = torch.empty(B, N, dtype=torch.float32, device=dev)
    dist2 = torch.empty_like(dist)
    pts = torch.empty(B, N, 4, dtype=torch.float64, device=dev)
    n_out = torch.empty(B, dtype=torch.int32, device=dev)
    wsb = lib.vo_match_workspace_bytes(F, B, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    nkp_d = torch.from_numpy(np.asarray(nkp, np.int32)).to(dev)
    pairs_d = torch.from_numpy(np.asarray(pairs, np.int32)).to(dev)
    assert lib.vo_debug_hamming(path) == 0
    try:
        _lib.check(lib.vo_match_hamming_batch(
            _lib.ptr(d_desc), _lib.ptr(nkp_d), _lib.ptr(d_kp), N, F, _lib.ptr(pairs_d), B, mode, ratio,
            _lib.ptr(q), _lib.ptr(t), _lib.ptr(dist), _lib.ptr(t2), _lib.ptr(dist2), _lib.ptr(pts),
            _lib.ptr(n_out), _lib.ptr(ws), wsb, _lib.cur_stream()))
        torch.cuda.synchronize()
    finally:
        lib.vo_debug_hamming(0)
    return [x.cpu().numpy() for x in (q, t, dist, t2, dist2, pts, n_out)]


@pytest.mark.parametrize("path", [1, 2, 3])
def test_batched_ragged_frames_and_gathered_points(cuda_dev, path):
    """batched entry: ragged keypoint counts, arbitrary pair list, fused point gather; the popc
    kernel and the two tcgen05 kernels (row-block / tile boundaries at 128, 208, 256, 416, 512)."""
    lib = _lib.load()
    seq = make_sequence(5, 600, 42)
    nkp = np.array([600, 512, 1, 333, 600], np.int32)
    pairs = np.array([[0, 1], [1, 2], [2, 3], [3, 4], [4, 0], [0, 0]], np.int32)
    N = 600
    q, t, dist, _, _, pts, n_out = _batch(lib, cuda_dev, seq.desc, seq.kp, nkp, pairs, N, path=path)
    for b, (fa, fb) in enumerate(pairs):
        da, db = seq.desc[fa, :nkp[fa]], seq.desc[fb, :nkp[fb]]
        rq, rt, rd = CR.match_from_matrix(CR.hamming_matrix(da, db), True)
        n = n_out[b]
        assert n == len(rq)
        assert np.array_equal(q[b, :n], rq) and np.array_equal(t[b, :n], rt)
        assert np.array_equal(dist[b, :n], rd)
        want = np.concatenate([seq.kp[fa][rq], seq.kp[fb][rt]], 1).astype(np.float64)
        assert np.array_equal(pts[b, :n], want)


def _tie_heavy_frames(rng, F, N):
    """descriptors with planted duplicates, near-duplicates, complements (distance 256) and rows of
    all-zero / all-one bits: every tie-break and the extreme distances"""
    desc = rng.integers(0, 256, (F, N, 32), dtype=np.uint8)
    for f in range(1, F):
        k = N // 2
        src = rng.permutation(N)[:k]
        desc[f, :k] = desc[f - 1, src]
        desc[f, :k] ^= np.packbits(rng.random((k, 256)) < 12 / 256, axis=1)
        desc[f, k:k + 8] = desc[f - 1, src[0]]            # eight copies of one train row
        desc[f, k + 8] = ~desc[f - 1, src[1]]             # complement: Hamming 256
    desc[:, -1] = 0
    desc[:, -2] = 255
    desc[0, 3] = desc[0, 300 % N]
    return desc


@pytest.mark.parametrize("tc_path", [2, 3])
@pytest.mark.parametrize("N,nk", [(700, [700, 699, 257, 256, 255, 1, 700, 128, 129, 207, 208, 209, 416, 417, 624]),
                                  (2000, [2000, 1777, 2000, 1025, 1872, 1873])])
def test_tensor_core_path_equals_cv2_and_popc_path(cuda_dev, N, nk, tc_path):
    """K1t (tcgen05 kind::i8) and K1f (kind::mxf4, e2m1) against live cv2 AND against the popc
    kernel: mutual matches and knn(k=2) + ratio, indices and distances bit-exact, ragged counts on
    every tile boundary of both kernels (256-row blocks; 256- / 208-column tiles)."""
    lib = _lib.load()
    rng = np.random.default_rng(N)
    F = len(nk)
    desc = _tie_heavy_frames(rng, F, N)
    kp = rng.random((F, N, 2), dtype=np.float32) * 1000
    pairs = np.array([[f, f + 1] for f in range(F - 1)] + [[F - 1, 0], [0, 0]], np.int32)
    out_tc = _batch(lib, cuda_dev, desc, kp, nk, pairs, N, path=tc_path)
    out_pc = _batch(lib, cuda_dev, desc, kp, nk, pairs, N, path=1)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)
    for b, (fa, fb) in enumerate(pairs):
        n = out_tc[6][b]
        assert n == out_pc[6][b]
        for k in (0, 1, 2, 5):
            assert np.array_equal(out_tc[k][b, :n], out_pc[k][b, :n]), (b, k)
        ref = bf.match(desc[fa, :nk[fa]], desc[fb, :nk[fb]])
        assert [m.queryIdx for m in ref] == out_tc[0][b, :n].tolist()
        assert [m.trainIdx for m in ref] == out_tc[1][b, :n].tolist()
        assert [m.distance for m in ref] == out_tc[2][b, :n].tolist()
    # knn(k=2) + ratio 0.8, and ratio < 0 = every query with both neighbours
    for ratio in (0.8, -1.0):
        k_tc = _batch(lib, cuda_dev, desc, kp, nk, pairs, N, mode=1, ratio=ratio, path=tc_path)
        k_pc = _batch(lib, cuda_dev, desc, kp, nk, pairs, N, mode=1, ratio=ratio, path=1)
        for b, (fa, fb) in enumerate(pairs):
            n = k_tc[6][b]
            assert n == k_pc[6][b]
            for k in range(5):
                assert np.array_equal(k_tc[k][b, :n], k_pc[k][b, :n]), (b, k, ratio)
            if nk[fb] >= 2 and ratio < 0:
                refk = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(desc[fa, :nk[fa]], desc[fb, :nk[fb]], k=2)
                assert [r[0].trainIdx for r in refk] == k_tc[1][b, :n].tolist()
                assert [r[1].trainIdx for r in refk] == k_tc[3][b, :n].tolist()
                assert [r[0].distance for r in refk] == k_tc[2][b, :n].tolist()
                assert [r[1].distance for r in refk] == k_tc[4][b, :n].tolist()


@pytest.mark.parametrize("tc_path", [2, 3])
def test_tensor_core_path_full_size_c4_properties(cuda_dev, tc_path):
    """config C4 size through K1t / K1f: 6 pairs x 4000 kp, results equal to the popc kernel's and
    every distance equal to the recomputed Hamming distance."""
    lib = _lib.load()
    seq = make_sequence(7, 4000, 77)
    pairs = seq.consecutive_pairs()
    a = _batch(lib, cuda_dev, seq.desc, seq.kp, seq.nkp, pairs, 4000, path=tc_path)
    b = _batch(lib, cuda_dev, seq.desc, seq.kp, seq.nkp, pairs, 4000, path=1)
    assert np.array_equal(a[6], b[6])
    for p in range(len(pairs)):
        n = a[6][p]
        assert n > 0.7 * 4000
        for k in (0, 1, 2, 5):
            assert np.array_equal(a[k][p, :n], b[k][p, :n])
        da, db = seq.desc[p][a[0][p, :n]], seq.desc[p + 1][a[1][p, :n]]
        assert np.array_equal(np.unpackbits(da ^ db, axis=1).sum(1).astype(np.float32), a[2][p, :n])


def test_full_size_property_c4(cuda_dev):
    """BASELINE config C4 size (4000 kp): mutual matches form a partial permutation and every
    reported distance equals the recomputed Hamming distance."""
    seq = make_sequence(2, 4000, 99)
    m = B200Matcher(NORM_HAMMING, crossCheck=True)
    q, t, d, _, _ = m.match_arrays(seq.desc[0], seq.desc[1])
    q, t, d = q.cpu().numpy(), t.cpu().numpy(), d.cpu().numpy()
    assert np.all(np.diff(q) > 0) and len(set(t.tolist())) == len(t)
    a, b = seq.desc[0][q], seq.desc[1][t]
    assert np.array_equal(np.unpackbits(a ^ b, axis=1).sum(1).astype(np.float32), d)
    true = (seq.point_id[0][q] == seq.point_id[1][t]).mean()
    assert true > 0.9 and len(q) > 0.7 * 4000


def test_descriptor_buffer_refilled_in_place_is_not_served_stale(cuda_dev):
    """cv2.BFMatcher reads the array it is given, every time.  A video loop that refills ONE
    ndarray in place must get matches for the new contents (the device-copy cache only keeps
    arrays the caller has frozen with flags.writeable = False)."""
    rng = np.random.default_rng(5)
    buf = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    other = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    m = B200Matcher(NORM_HAMMING, crossCheck=True)
    first = _triples(m.match(buf, other))
    buf[:] = rng.integers(0, 256, (300, 32), dtype=np.uint8)          # same object, new frame
    second = _triples(m.match(buf, other))
    ref = _triples(cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(buf, other))
    assert second == ref and second != first
    frozen = buf.copy()
    frozen.flags.writeable = False                                    # opt in to the cache
    assert _triples(m.match(frozen, other)) == ref == _triples(m.match(frozen, other))


def test_fp4_accumulator_is_the_exact_key(cuda_dev):
    """K1f's raw fp32 accumulator (vo_debug_hamming_f4_acc) of one 128 x 208 block equals
    36 * (256 - 2 * Hamming) + (207 - column) / 4 exactly: the e2m1 operand forms, the all-ones
    block scales and the index step of the kind::mxf4 contraction, checked against numpy."""
    lib = _lib.load()
    rng = np.random.default_rng(3)
    N = 1000
    d = rng.integers(0, 256, (2, N, 32), dtype=np.uint8)
    d[0, :64] = d[1, rng.permutation(208)[:64]]
    d[0, :64, :2] ^= rng.integers(0, 256, (64, 2), dtype=np.uint8)
    d[0, 100] = 0; d[1, 7] = 255; d[0, 101] = 255                      # distances 256 and 0
    desc = torch.from_numpy(d).to(cuda_dev)
    nkp = torch.tensor([N, N], dtype=torch.int32, device=cuda_dev)
    pairs = torch.tensor([[0, 1]], dtype=torch.int32, device=cuda_dev)
    acc = torch.full((128, 208), -12345.0, dtype=torch.float32, device=cuda_dev)
    wsb = lib.vo_match_workspace_bytes(2, 1, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device=cuda_dev)
    _lib.check(lib.vo_debug_hamming_f4_acc(_lib.ptr(desc), _lib.ptr(nkp), N, _lib.ptr(pairs), _lib.ptr(acc),
                                           _lib.ptr(ws), wsb, _lib.cur_stream()))
    torch.cuda.synchronize()
    bits = np.unpackbits(d, axis=2).astype(np.int32)
    H = (bits[0, :128, None, :] != bits[1, None, :208, :]).sum(-1)
    want = 36.0 * (256 - 2 * H) + 0.25 * (207 - np.arange(208))[None, :]
    assert np.array_equal(acc.cpu().numpy().astype(np.float64), want)
