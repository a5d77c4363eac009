"""CPU: pins the oracle (oracle/*.py) against the golden vectors produced by the REAL reference
ImagePair (tests/golden/make_golden.py) and against cv2 itself."""
import numpy as np
import pytest

import cv2
from oracle import cv_geometry as CG
from oracle import cv_restate as CR
from oracle import reference_path as RP
from visual_odometry_b200.synth import make_sequence


def _dist_matrix(g):
    if g["descriptor"] == "orb":
        return CR.hamming_matrix(g["desc"][0], g["desc"][1])
    return CR.l2_matrix(g["desc"][0], g["desc"][1])


def test_matching_restatement_equals_reference(golden):
    q, t, d = CR.match_from_matrix(_dist_matrix(golden), cross_check=True)
    assert np.array_equal(q, golden["q"]) and np.array_equal(t, golden["t"])
    assert np.array_equal(d, golden["dist"])  # bit-exact float32 distances


def test_reference_path_equals_reference(golden):
    g = golden
    norm = cv2.NORM_HAMMING if g["descriptor"] == "orb" else cv2.NORM_L2
    o = RP.pair_cv2(g["desc"][0], g["kp"][0], g["desc"][1], g["kp"][1], g["K"], norm)
    assert np.array_equal(o["q"], g["q"]) and np.array_equal(o["t"], g["t"])
    assert np.array_equal(o["E"], g["E"])
    assert np.array_equal(o["q"][o["mask"].ravel() == 1], g["inlier_q"])
    assert np.array_equal(o["R"], g["R"]) and np.array_equal(o["tvec"], g["tvec"])
    assert np.allclose(o["X"], g["X"], rtol=1e-12, atol=0)


def test_ransac_restatement_equals_reference(golden):
    g = golden
    p1 = g["kp"][0][g["q"]].astype(np.float64)
    p2 = g["kp"][1][g["t"]].astype(np.float64)
    E, mask = CG.find_essential_ransac(p1, p2, g["K"])
    assert np.array_equal(g["q"][mask.ravel() == 1], g["inlier_q"])  # inlier set exact
    dE = min(np.abs(E - g["E"]).max(), np.abs(E + g["E"]).max())
    assert dE < 1e-7
    inl = mask.ravel() == 1
    n, R, t, pm = CG.recover_pose(g["E"], p1[inl], p2[inl], g["K"])
    assert np.abs(R - g["R"]).max() < 1e-12 and np.abs(t - g["tvec"]).max() < 1e-12
    X = CG.triangulate_points(g["projection_matrix"], g["null_projection_matrix"],
                              p1[inl].T, p2[inl].T)
    X = X / X[3]
    rel = np.abs(X - g["X"]).max(0) / np.abs(g["X"]).max(0)
    assert np.median(rel) < 1e-9 and rel.max() < 1e-4  # north_star: points within 1e-4 relative


def test_rng_replay_matches_cv2_winning_sample():
    """the restated cv::RNG stream contains cv2's winning minimal sample."""
    seq = make_sequence(2, 400, 7)
    q, t, _ = RP.match_cv2(seq.desc[0], seq.desc[1], cv2.NORM_HAMMING)
    p1 = seq.kp[0][q].astype(np.float64)
    p2 = seq.kp[1][t].astype(np.float64)
    E, _ = cv2.findEssentialMat(p1, p2, seq.K, cv2.FM_RANSAC, 0.99, 1)
    x1, x2 = CG.normalize_points(p1, seq.K), CG.normalize_points(p2, seq.K)
    res = np.abs(np.einsum("ni,ij,nj->n", np.c_[x2, np.ones(len(x2))], E,
                           np.c_[x1, np.ones(len(x1))]))
    on = set(np.argsort(res)[:5].tolist())
    samples = CR.ransac_samples(len(p1), 200)
    assert any(set(s.tolist()) == on for s in samples)


def test_knn_and_ratio_restatement_equal_cv2():
    rng = np.random.default_rng(3)
    d1 = rng.integers(0, 256, (150, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (170, 32), dtype=np.uint8)
    d1[:40] = d2[:40]
    d1[5] = d1[3]
    d2[9] = d2[3]
    D = CR.hamming_matrix(d1, d2)
    ref = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2)
    idx, dist = CR.knn2_from_matrix(D)
    assert [[m.trainIdx for m in r] for r in ref] == idx.tolist()
    assert [[m.distance for m in r] for r in ref] == dist.tolist()
    q, t, d = CR.ratio_from_matrix(D, 0.3)
    good = [m for m, n in ref if m.distance < 0.3 * n.distance]
    assert [m.queryIdx for m in good] == q.tolist() and [m.trainIdx for m in good] == t.tolist()


def test_update_num_iters_matches_cv2_behaviour():
    assert CR.ransac_update_num_iters(0.99, 0.3, 5, 1000) == 25
    assert CR.ransac_update_num_iters(0.99, 1.0, 5, 1000) == 1000
    assert CR.ransac_update_num_iters(0.99, 0.0, 5, 1000) == 0


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fivept_restatement_same_solution_count_as_cv2(seed):
    seq = make_sequence(2, 200, seed, inlier_ratio=0.9)
    q, t, _ = RP.match_cv2(seq.desc[0], seq.desc[1], cv2.NORM_HAMMING)
    p1 = seq.kp[0][q].astype(np.float64)
    p2 = seq.kp[1][t].astype(np.float64)
    rng = np.random.default_rng(seed)
    for _ in range(15):
        idx = rng.choice(len(p1), 5, replace=False)
        Ecv, _ = cv2.findEssentialMat(p1[idx], p2[idx], seq.K, cv2.FM_RANSAC, 0.99, 1)
        mine = CG.fivept_solve(CG.normalize_points(p1[idx], seq.K),
                               CG.normalize_points(p2[idx], seq.K))
        assert (0 if Ecv is None else Ecv.shape[0] // 3) == len(mine)


def test_fp4_key_arithmetic_of_the_tensor_core_hamming_kernel():
    """The arithmetic csrc/hamming_f4.cuh relies on, restated in numpy (no GPU): bits as +-6 in
    e2m1 give 36 * (256 - 2 * Hamming); the index step (ones group 4, 4, 4, 4, 4, 1, 0.5 against
    the base-4 digits of x = 207 - column, d0 stored as d0 / 2) adds exactly x / 4; every key is
    a multiple of 1/4 below 2^24 (exact in fp32), one Hamming step (72) outweighs the index part
    (<= 51.75), and 4 * key + 36864 = 288 * (256 - H) + x decodes both."""
    e2m1 = {0: 0.0, 1: 0.5, 2: 1.0, 3: 1.5, 4: 2.0, 5: 3.0, 6: 4.0, 7: 6.0}
    val = lambda c: -e2m1[c & 7] if c & 8 else e2m1[c & 7]
    ones = [6, 6, 6, 6, 6, 2, 1]                       # nibble codes of 0x01266666
    code = lambda d: 2 * d - (d == 3)                  # values 0, 1, 2, 3 -> codes 0, 2, 4, 5
    for x in range(208):
        d3, d2, d1, d0 = x >> 6, (x >> 4) & 3, (x >> 2) & 3, x & 3
        idx = [code(d3)] * 4 + [code(d2), code(d1), d0]
        assert sum(val(a) * val(b) for a, b in zip(ones, idx)) == x / 4
    rng = np.random.default_rng(0)
    a = rng.integers(0, 2, (64, 256)); b = rng.integers(0, 2, (208, 256))
    a[0] = b[5]; a[1] = 1 - b[7]                        # Hamming 0 and 256
    enc = lambda bits: np.where(bits == 1, val(0x7), val(0xF))
    H = (a[:, None, :] != b[None, :, :]).sum(-1)
    key = (enc(a) @ enc(b).T + 0.25 * (207 - np.arange(208))[None, :]).astype(np.float32)
    assert np.array_equal(key.astype(np.float64), 36.0 * (256 - 2 * H) + 0.25 * (207 - np.arange(208))[None, :])
    y = (key * 4).astype(np.int64) + 36864
    assert np.array_equal(256 - y // 288, H) and np.array_equal(207 - y % 288, np.broadcast_to(np.arange(208), H.shape))
    # the largest key of a row = its nearest column, lowest column on ties
    want = np.array([np.flatnonzero(H[i] == H[i].min())[0] for i in range(64)])
    assert np.array_equal(key.argmax(1), want)
