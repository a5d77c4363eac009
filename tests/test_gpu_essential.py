"""GPU: cv::RNG replay, five-point solver, Sampson scoring and batched E-RANSAC vs oracle/cv2."""
import numpy as np
import pytest
import torch

import cv2
from oracle import cv_geometry as CG
from oracle import cv_restate as CR
from oracle import reference_path as RP
from visual_odometry_b200 import geometry as G
from visual_odometry_b200 import image_pair as IP
from visual_odometry_b200.synth import make_sequence
from conftest import record_parity, load_golden, GOLDEN_CASES

pytestmark = pytest.mark.gpu


def _pair_points(seed, n=600, inl=0.7):
    seq = make_sequence(2, n, seed, inlier_ratio=inl)
    q, t, _ = RP.match_cv2(seq.desc[0], seq.desc[1], cv2.NORM_HAMMING)
    return seq.kp[0][q].astype(np.float64), seq.kp[1][t].astype(np.float64), seq.K


@pytest.mark.parametrize("count", [5, 6, 17, 300, 1500])
def test_rng_replay_bit_exact(cuda_dev, count):
    got = G.ransac_samples(count, 300).cpu().numpy()
    assert np.array_equal(got, CR.ransac_samples(count, 300))


def test_fivept_matches_cv2_minimal_solver(cuda_dev):
    p1, p2, K = _pair_points(11)
    ptsn = np.c_[CG.normalize_points(p1, K), CG.normalize_points(p2, K)]
    rng = np.random.default_rng(0)
    S = 200
    samples = np.stack([rng.choice(len(p1), 5, replace=False) for _ in range(S)]).astype(np.int32)
    E, n = G.fivept_solve(torch.from_numpy(ptsn).cuda(), torch.from_numpy(samples).cuda())
    E, n = E.cpu().numpy(), n.cpu().numpy()
    n_close, n_total, count_mismatch = 0, 0, 0
    for s in range(S):
        idx = samples[s]
        Ecv, _ = cv2.findEssentialMat(p1[idx], p2[idx], K, cv2.FM_RANSAC, 0.99, 1)
        ncv = 0 if Ecv is None else Ecv.shape[0] // 3
        if ncv != n[s]:
            count_mismatch += 1
            continue
        for k in range(n[s]):
            e = E[s, k]
            assert abs(np.linalg.norm(e) - 1) < 1e-12
            # a true essential matrix through the 5 points
            x1 = np.c_[ptsn[idx, :2], np.ones(5)]
            x2 = np.c_[ptsn[idx, 2:], np.ones(5)]
            assert np.abs(np.einsum("ni,ij,nj->n", x2, e, x1)).max() < 1e-10
            assert np.abs(2 * e @ e.T @ e - np.trace(e @ e.T) * e).max() < 1e-8
            d = min(min(np.abs(e - Ecv[3 * j:3 * j + 3]).max(), np.abs(e + Ecv[3 * j:3 * j + 3]).max())
                    for j in range(ncv))
            n_total += 1
            n_close += d < 1e-6
    record_parity("fivept_vs_cv2_minimal_solver", samples=S, solution_count_mismatches=count_mismatch,
                  solutions_compared=n_total, solutions_within_1e_6=n_close)
    assert count_mismatch == 0            # same number of real solutions as OpenCV (measured: 0 / 200)
    assert n_close >= 0.995 * n_total     # and the same matrices (cv2's own roots are ~1e-6 off
    #                                       on ill-conditioned samples; ours satisfy the constraints)


def test_sampson_masks_exact_with_identical_hypotheses(cuda_dev):
    """north_star: with identical hypotheses, inlier masks exact (vs the restated computeError)."""
    p1, p2, K = _pair_points(12, n=1500)
    x1, x2 = CG.normalize_points(p1, K), CG.normalize_points(p2, K)
    Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.FM_RANSAC, 0.99, 1)
    rng = np.random.default_rng(1)
    hyps = [Ecv] + [CG.fivept_solve(x1[i], x2[i]) for i in
                    (rng.choice(len(p1), 5, replace=False) for _ in range(30))]
    H = np.concatenate([h.reshape(-1, 3, 3) for h in hyps])
    thr = 1.0 / ((K[0, 0] + K[1, 1]) / 2)
    thr2 = np.float32(thr * thr)
    counts, masks = G.sampson_score(torch.from_numpy(H).cuda(),
                                    torch.from_numpy(np.c_[x1, x2]).cuda(), float(thr2), True)
    masks = masks.cpu().numpy().astype(bool)
    for h in range(len(H)):
        assert np.array_equal(masks[h], CG.sampson_mask(H[h], x1, x2, thr))
    assert np.array_equal(masks[0], mcv.ravel() == 1)      # and vs cv2's own mask for its E
    assert np.array_equal(counts.cpu().numpy(), masks.sum(1))


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_essential_equals_golden(cuda_dev, name):
    g = load_golden(name)
    p1 = g["kp"][0][g["q"]].astype(np.float64)
    p2 = g["kp"][1][g["t"]].astype(np.float64)
    E, mask = IP.find_essential_mat(p1, p2, g["K"], 0.99, 1)
    assert mask.shape == (len(p1), 1) and mask.dtype == np.uint8 and set(np.unique(mask)) <= {0, 1}
    assert np.array_equal(g["q"][mask.ravel() == 1], g["inlier_q"])   # inlier mask exact
    assert min(np.abs(E - g["E"]).max(), np.abs(E + g["E"]).max()) < 1e-6
    assert abs(np.linalg.norm(E) - 1) < 1e-12


def test_batch_vs_cv2_sequence(cuda_dev):
    seq = make_sequence(25, 1200, 5)
    K, B, N = seq.K, 24, 1200
    pts = np.zeros((B, N, 4))
    m = np.zeros(B, np.int32)
    ref = []
    for b in range(B):
        q, t, _ = RP.match_cv2(seq.desc[b], seq.desc[b + 1], cv2.NORM_HAMMING)
        m[b] = len(q)
        pts[b, :m[b], :2], pts[b, :m[b], 2:] = seq.kp[b][q], seq.kp[b + 1][t]
        ref.append(cv2.findEssentialMat(pts[b, :m[b], :2], pts[b, :m[b], 2:], K, cv2.FM_RANSAC, 0.99, 1))
    out = G.find_essential_batch(torch.from_numpy(pts).cuda(), torch.from_numpy(m).cuda(),
                                 torch.from_numpy(K).cuda())
    E, M = out["E"].cpu().numpy(), out["mask"].cpu().numpy()
    same, diff_bits, e_dev = 0, [], 0.0
    for b in range(B):
        Ecv, mcv = ref[b]
        eq = np.array_equal(M[b, :m[b]], mcv.ravel())
        same += eq
        if eq:    # same consensus set: E itself agrees up to sign and root-finder accuracy
            e_dev = max(e_dev, min(np.abs(E[b] - Ecv).max(), np.abs(E[b] + Ecv).max()))
        else:
            diff_bits.append(int((M[b, :m[b]] != mcv.ravel()).sum()))
        assert np.all(M[b, m[b]:] == 0)
    record_parity("essential_batch_vs_cv2_sequence", pairs=B, masks_identical=same,
                  mismatching_mask_bits=diff_bits, max_abs_E_dev_on_identical=e_dev)
    assert same == B                          # identical masks to cv2 end to end (measured: 24/24)
    assert e_dev < 1e-5
    assert np.array_equal(out["n_inl"].cpu().numpy(), M.sum(1))
    assert np.all(out["status"].cpu().numpy() == 0)


def test_fixed_hypothesis_count_c3_mode(cuda_dev):
    """config C3 flavour: adaptive stop off, exactly max_iters sample sets are evaluated; the
    result equals the oracle RANSAC run with the same setting."""
    p1, p2, K = _pair_points(21, n=500, inl=0.5)
    pts = torch.from_numpy(np.c_[p1, p2][None]).cuda().contiguous()
    m = torch.tensor([len(p1)], dtype=torch.int32, device="cuda")
    out = G.find_essential_batch(pts, m, torch.from_numpy(K).cuda(), max_iters=300, adaptive=False)
    Eo, mo = CG.find_essential_ransac(p1, p2, K, max_iters=300, adaptive=False)
    assert int(out["n_iters"][0]) == 300
    assert np.array_equal(out["mask"][0, :len(p1)].cpu().numpy(), mo.ravel())


def test_too_few_points_raises_cv2_error(cuda_dev):
    p = np.random.default_rng(0).uniform(0, 1000, (4, 2))
    with pytest.raises(cv2.error):
        IP.find_essential_mat(p, p + 1, np.eye(3))
    # batched: status flags instead of exceptions, other pairs unaffected
    p1, p2, K = _pair_points(3, n=300)
    pts = np.zeros((2, len(p1), 4))
    pts[0, :, :2], pts[0, :, 2:] = p1, p2
    out = G.find_essential_batch(torch.from_numpy(pts).cuda(),
                                 torch.tensor([len(p1), 3], dtype=torch.int32, device="cuda"),
                                 torch.from_numpy(K).cuda())
    st = out["status"].cpu().numpy()
    assert st[0] == 0 and st[1] == -2 and int(out["n_inl"][1]) == 0
