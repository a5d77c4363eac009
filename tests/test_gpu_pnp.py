"""GPU: K7 PnP-RANSAC through the C-ABI vs cv2.solvePnPRansac (reference visual_slam.py:231)."""
import numpy as np
import pytest

import cv2
from oracle import cv_pnp as P
from visual_odometry_b200 import pnp
from test_oracle_pnp import scene, K

pytestmark = pytest.mark.gpu


def test_pnp_ransac_vs_cv2(cuda_dev):
    same, N = 0, 16
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
            assert len(set(inl.ravel()) ^ set(inl2.ravel())) <= 0.02 * len(inl)
        assert np.abs(r - r2).max() < 1e-4 and np.abs(t - t2).max() / max(1, np.abs(t).max()) < 1e-3
    assert same >= 0.75 * N


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
