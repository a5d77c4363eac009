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


def test_batched_ragged_frames_and_gathered_points(cuda_dev):
    """batched entry: ragged keypoint counts, arbitrary pair list, fused point gather."""
    lib = _lib.load()
    seq = make_sequence(5, 600, 42)
    nkp = np.array([600, 512, 1, 333, 600], np.int32)
    pairs = np.array([[0, 1], [1, 2], [2, 3], [3, 4], [4, 0], [0, 0]], np.int32)
    B, N = len(pairs), 600
    dev = cuda_dev
    desc = torch.from_numpy(seq.desc).to(dev)
    kp = torch.from_numpy(seq.kp).to(dev)
    q = torch.empty(B, N, dtype=torch.int32, device=dev)
    t = torch.empty_like(q)
    dist = torch.empty(B, N, dtype=torch.float32, device=dev)
    pts = torch.empty(B, N, 4, dtype=torch.float64, device=dev)
    n_out = torch.empty(B, dtype=torch.int32, device=dev)
    wsb = lib.vo_match_workspace_bytes(B, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    nkp_d, pairs_d = torch.from_numpy(nkp).to(dev), torch.from_numpy(pairs).to(dev)
    _lib.check(lib.vo_match_hamming_batch(
        _lib.ptr(desc), _lib.ptr(nkp_d), _lib.ptr(kp), N, _lib.ptr(pairs_d), B, 0, 0.0, _lib.ptr(q), _lib.ptr(t),
        _lib.ptr(dist), None, None, _lib.ptr(pts), _lib.ptr(n_out), _lib.ptr(ws), wsb,
        _lib.cur_stream()))
    n_out = n_out.cpu().numpy()
    for b, (fa, fb) in enumerate(pairs):
        da, db = seq.desc[fa, :nkp[fa]], seq.desc[fb, :nkp[fb]]
        rq, rt, rd = CR.match_from_matrix(CR.hamming_matrix(da, db), True)
        n = n_out[b]
        assert n == len(rq)
        assert np.array_equal(q[b, :n].cpu().numpy(), rq) and np.array_equal(t[b, :n].cpu().numpy(), rt)
        assert np.array_equal(dist[b, :n].cpu().numpy(), rd)
        want = np.concatenate([seq.kp[fa][rq], seq.kp[fb][rt]], 1).astype(np.float64)
        assert np.array_equal(pts[b, :n].cpu().numpy(), want)


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
