"""GPU: K7 PnP-RANSAC through the C-ABI vs cv2.solvePnPRansac (reference visual_slam.py:231)."""
import numpy as np
import pytest

import cv2
from oracle import cv_pnp as P
from visual_odometry_b200 import pnp
from test_oracle_pnp import scene, K
from conftest import record_parity

pytestmark = pytest.mark.gpu


def test_pnp_ransac_vs_cv2(cuda_dev):
    same, N, off = 0, 16, []
    for seed in range(N):
        X, x, _, _ = scene(seed, n=200 + 150 * seed, out=0.1 + 0.02 * (seed % 12))
        okc, r, t, inl = cv2.solvePnPRansac(X, x, K, np.zeros(4))
        okg, r2, t2, inl2 = pnp.solve_pnp_ransac(X, x, K, np.zeros(4))
        assert okc and okg
        assert r2.shape == (3, 1) and t2.shape == (3, 1) and inl2.dtype == np.int32 and inl2.shape[1] == 1
        assert np.all(np.diff(inl2.ravel()) > 0)
        if np.array_equal(inl, inl2):
            same += 1
            assert np.abs(r - r2).max() < 1e-8 and np.abs(t - t2).max() < 1e-7
        else:   # borderline 8 px points may flip; the pose stays within the north_star tolerance
            off.append({"seed": seed, "inliers": [len(inl2), len(inl)],
                        "set_difference": len(set(inl.ravel()) ^ set(inl2.ravel())),
                        "rvec_abs": float(np.abs(r - r2).max()), "tvec_abs": float(np.abs(t - t2).max())})
            assert len(set(inl.ravel()) ^ set(inl2.ravel())) <= 0.02 * len(inl)
        assert np.abs(r - r2).max() < 1e-4 and np.abs(t - t2).max() / max(1, np.abs(t).max()) < 1e-3
    record_parity("pnp_ransac_vs_cv2", scenes=N, inlier_sets_identical=same, scenes_with_different_set=off)
    assert same >= 0.75 * N                   # measured: 14 / 16 identical, the other two differ by 1 and 11 points


def test_pnp_four_points_is_p3p_like_cv2(cuda_dev):
    """npoints == 4: cv2.solvePnPRansac = solvePnP(SOLVEPNP_P3P) on the four points, every point
    an inlier, no refit (solvepnp.cpp).  Any exact P3P has the same solutions: pose within 1e-4."""
    rng = np.random.default_rng(3)
    n_cmp, n_none, worst = 0, 0, 0.0
    for trial in range(60):
        X = np.c_[rng.uniform(-5, 5, 4), rng.uniform(-3, 3, 4), rng.uniform(8, 20, 4)]
        rv, tv = rng.normal(0, 0.1, 3), rng.normal(0, 0.5, 3)
        x = cv2.projectPoints(X, rv, tv, K, None)[0].reshape(-1, 2) + rng.normal(0, 0.3, (4, 2))
        okc, r, t, inl = cv2.solvePnPRansac(X, x, K, np.zeros(4))
        okg, r2, t2, inl2 = pnp.solve_pnp_ransac(X, x, K, np.zeros(4))
        if not (okc and np.all(np.isfinite(t))):
            # no real P3P solution: cv2 4.13 answers True with a NaN translation; here: False
            assert not okg
            n_none += 1
            continue
        assert okg and inl2.ravel().tolist() == [0, 1, 2, 3] == inl.ravel().tolist()
        R1, R2 = cv2.Rodrigues(r)[0], cv2.Rodrigues(r2)[0]
        ang = np.arccos(np.clip((np.trace(R1.T @ R2) - 1) / 2, -1, 1))
        worst = max(worst, ang, np.abs(t - t2).max())
        assert ang < 1e-4 and np.abs(t - t2).max() < 1e-4
        n_cmp += 1
    record_parity("pnp_four_points_p3p", compared=n_cmp, no_real_solution=n_none, worst_dev=worst)
    assert n_cmp >= 50


def test_pnp_five_points_is_plain_epnp_like_cv2(cuda_dev):
    """npoints == 5: cv2 runs EPnP once on all five points (no RANSAC, no refit).  With five
    points EPnP's 10 x 12 system has a two-dimensional null space whose basis is decided by
    rounding inside the eigen-solver, so two correct EPnP implementations
    can settle in different minima of the Gauss-Newton stage: the check is that both explain the
    five observations equally well, the pose deviation is recorded."""
    rng = np.random.default_rng(4)
    worst_t, worst_ratio = 0.0, 0.0
    for trial in range(30):
        X = np.c_[rng.uniform(-5, 5, 5), rng.uniform(-3, 3, 5), rng.uniform(8, 20, 5)]
        rv, tv = rng.normal(0, 0.1, 3), rng.normal(0, 0.5, 3)
        x = cv2.projectPoints(X, rv, tv, K, None)[0].reshape(-1, 2) + rng.normal(0, 0.3, (5, 2))
        okc, r, t, inl = cv2.solvePnPRansac(X, x, K, np.zeros(4))
        okg, r2, t2, inl2 = pnp.solve_pnp_ransac(X, x, K, np.zeros(4))
        assert okc and okg and inl2.ravel().tolist() == [0, 1, 2, 3, 4] == inl.ravel().tolist()
        e1 = np.linalg.norm(cv2.projectPoints(X, r, t, K, None)[0].reshape(-1, 2) - x, axis=1).mean()
        e2 = np.linalg.norm(cv2.projectPoints(X, r2, t2, K, None)[0].reshape(-1, 2) - x, axis=1).mean()
        worst_t = max(worst_t, float(np.abs(t - t2).max()))
        worst_ratio = max(worst_ratio, float(e2 / max(e1, 1e-9)))
        assert e2 < 2.0 * e1 + 0.5
    record_parity("pnp_five_points_epnp", cases=30, worst_tvec_abs_vs_cv2=worst_t,
                  worst_reprojection_ratio_vs_cv2=worst_ratio)


def test_pnp_batch_c5_64_frames_vs_cv2(cuda_dev):
    """BASELINE config C5: 64 new frames x 5000 observations of a 100 k-point map in ONE launch
    group (vo_pnp_ransac_batch, CSR offsets) vs cv2.solvePnPRansac frame by frame; ragged frame
    sizes incl. the 4- and 5-point cases and an empty frame; batch == single-call results."""
    rng = np.random.default_rng(11)
    n_map = 100_000
    M = np.c_[rng.uniform(-12, 12, n_map), rng.uniform(-7, 7, n_map), rng.uniform(8, 40, n_map)]
    sizes = [5000] * 60 + [4, 5, 0, 37]
    Xs, xs = [], []
    for n in sizes:
        rv, tv = rng.normal(0, 0.03, 3), rng.normal(0, 0.4, 3)
        X = M[rng.permutation(n_map)[:n]]
        if n:
            img = cv2.projectPoints(X, rv, tv, K, None)[0].reshape(-1, 2) + rng.normal(0, 0.5, (n, 2))
            if n > 5:
                bad = rng.random(n) < 0.2
                img[bad] = rng.uniform([0, 0], [1920, 1080], (int(bad.sum()), 2))
        else:
            img = np.zeros((0, 2))
        Xs.append(X)
        xs.append(img)
    res = pnp.solve_pnp_ransac_batch(Xs, xs, K)
    same, off = 0, []
    for f, n in enumerate(sizes):
        ok, r2, t2, inl2 = res[f]
        if n < 4:
            assert not ok
            continue
        okc, r, t, inl = cv2.solvePnPRansac(Xs[f], xs[f], K, np.zeros(4))
        assert ok == okc
        if n > 5:
            if np.array_equal(inl, inl2):
                same += 1
                assert np.abs(r - r2).max() < 1e-7 and np.abs(t - t2).max() < 1e-6
            else:
                off.append({"frame": f, "set_difference": len(set(inl.ravel()) ^ set(inl2.ravel())),
                            "rvec_abs": float(np.abs(r - r2).max()), "tvec_abs": float(np.abs(t - t2).max())})
                # a different RANSAC winner (5-point EPnP hypotheses are reproducible only to ~1e-4,
                # DESIGN.md section 2): whichever consensus set is larger, the refit poses agree
                assert len(set(inl.ravel()) ^ set(inl2.ravel())) <= 0.10 * len(inl)
            assert np.abs(r - r2).max() < 1e-4 and np.abs(t - t2).max() < 1e-3
        if f in (0, 60, 61, 63):   # the batch entry and the single entry give the same bits
            ok1, r1, t1, inl1 = pnp.solve_pnp_ransac(Xs[f], xs[f], K, np.zeros(4))
            assert ok1 == ok and np.array_equal(r1, r2) and np.array_equal(t1, t2) and np.array_equal(inl1, inl2)
    record_parity("pnp_batch_c5_64_frames", frames_ransac=61, inlier_sets_identical=same,
                  frames_with_different_set=off)
    assert same >= 45


def test_pnp_inliers_are_those_of_the_ransac_model(cuda_dev):
    X, x, _, _ = scene(3, n=800, out=0.3)
    ok, r, t, inl = pnp.solve_pnp_ransac(X, x, K, np.zeros(4))
    ok2, r2, t2, inl2, model = P.solve_pnp_ransac(X, x, K, return_model=True)
    assert np.array_equal(inl, inl2)


def test_rodrigues_and_errors(cuda_dev):
    r = np.array([[0.1], [-0.2], [0.3]])
    R, _ = pnp.rodrigues(r)
    assert np.abs(R - cv2.Rodrigues(r)[0]).max() < 1e-14
    assert np.abs(pnp.rodrigues(R)[0] - r).max() < 1e-12
    with pytest.raises(cv2.error):
        pnp.solve_pnp_ransac(np.zeros((3, 3)), np.zeros((3, 2)), K, np.zeros(4))
    # all-outlier input: no model with more than 4 inliers is not guaranteed, but the call
    # must return cleanly with cv2's (retval, rvec, tvec, inliers) shape
    rng = np.random.default_rng(0)
    out = pnp.solve_pnp_ransac(rng.uniform(-5, 5, (50, 3)) + [0, 0, 20], rng.uniform(0, 1000, (50, 2)), K, None)
    assert len(out) == 4


def test_cv2_shim_routes_pnp_to_gpu(cuda_dev):
    from visual_odometry_b200.cv2_shim import cv2 as shim
    X, x, _, _ = scene(5, n=400, out=0.2)
    ok, r, t, inl = shim.solvePnPRansac(X, x, K, np.zeros(4))
    okc, rc, tc, inlc = cv2.solvePnPRansac(X, x, K, np.zeros(4))
    assert ok and np.abs(r - rc).max() < 1e-4
    assert np.abs(shim.Rodrigues(r)[0] - cv2.Rodrigues(rc)[0]).max() < 1e-4
    assert shim.FM_RANSAC == cv2.FM_RANSAC and shim.imshow("w", np.zeros((2, 2))) is None
