"""CPU: oracle/ba_restate.py (restatement of g2o's Levenberg + Schur + EdgeProjectXYZ2UV + Huber,
the arithmetic of the reference's Map.optimize_map).  g2o is absent, so the restatement is checked
for self-consistency and against independent numerics: analytic Jacobians vs central differences of
the error under g2o's own update rule, LM descent, first-order optimality of the result, and
scipy.optimize.least_squares being unable to improve on it."""
import numpy as np
import pytest

from oracle import ba_restate as BA
from ba_problem import make_ba


def test_jacobians_match_numerical_derivatives_of_the_update_rule():
    d = make_ba(3, 12, 3, n_fixed=0)
    pr = BA.BAProblem(**d)
    s = pr.state
    Hpp, bp, Hll, bl, Hpl = pr.build_system(s)
    e0 = pr.errors(s)
    w = BA.huber(e0[:, 0] ** 2 + e0[:, 1] ** 2)[1]
    h = 1e-6
    # b = -J^T w e, so check J through b and H = J^T w J on one camera and one point
    for c in range(3):
        J = np.zeros((len(e0), 2, 6))
        for k in range(6):
            dx = np.zeros((3, 6)); dx[c, k] = h
            ep = pr.errors(pr.apply(s, dx, np.zeros_like(s.points)))
            em = pr.errors(pr.apply(s, -dx, np.zeros_like(s.points)))
            J[:, :, k] = (ep - em) / (2 * h)
        sel = pr.obs_cam == c
        Hn = np.einsum("m,mki,mkj->ij", w[sel], J[sel], J[sel])
        bn = -np.einsum("m,mki,mk->i", w[sel], J[sel], e0[sel])
        assert np.allclose(Hn, Hpp[c], rtol=1e-5, atol=1e-5 * np.abs(Hpp[c]).max())
        assert np.allclose(bn, bp[c], rtol=1e-5, atol=1e-5 * np.abs(bp[c]).max())
    for p in range(4):
        J = np.zeros((len(e0), 2, 3))
        for k in range(3):
            dl = np.zeros_like(s.points); dl[p, k] = h
            ep = pr.errors(pr.apply(s, np.zeros((3, 6)), dl))
            em = pr.errors(pr.apply(s, np.zeros((3, 6)), -dl))
            J[:, :, k] = (ep - em) / (2 * h)
        sel = pr.obs_pt == p
        Hn = np.einsum("m,mki,mkj->ij", w[sel], J[sel], J[sel])
        assert np.allclose(Hn, Hll[p], rtol=1e-5, atol=1e-5 * np.abs(Hll[p]).max())


def test_schur_solution_solves_the_full_damped_system():
    d = make_ba(4, 30, 5)
    pr = BA.BAProblem(**d)
    sysm = pr.build_system(pr.state)
    Hpp, bp, Hll, bl, Hpl = sysm
    lam = 3.7
    xp, xl, ok = pr.solve(sysm, lam)
    assert ok
    # residual of (H + lam I) x = b, block row by block row
    rp = (Hpp + lam * np.eye(6)) @ xp[:, :, None]
    rp = rp[:, :, 0].copy()
    np.add.at(rp, pr.obs_cam, np.einsum("mij,mj->mi", Hpl, xl[pr.obs_pt]))
    free = ~pr.fixed
    assert np.abs(rp[free] - bp[free]).max() < 1e-8 * np.abs(bp).max()
    rl = np.einsum("pij,pj->pi", Hll + lam * np.eye(3), xl)
    np.add.at(rl, pr.obs_pt, np.einsum("mji,mj->mi", Hpl, xp[pr.obs_cam]))
    assert np.abs(rl - bl).max() < 1e-8 * np.abs(bl).max()
    assert np.all(xp[pr.fixed] == 0)


def test_lm_descends_and_reaches_a_stationary_point():
    d = make_ba(4, 60, 7)
    pr = BA.BAProblem(**d)
    chi0 = pr.robust_chi2(pr.state)
    with np.errstate(all="ignore"):
        log = pr.optimize(40)
    chis = [chi0] + [l[1] for l in log]
    assert all(b <= a for a, b in zip(chis, chis[1:]))
    assert chis[-1] < 0.2 * chi0
    # numerical gradient of the robust cost w.r.t. g2o's local parameters, before and after
    free = np.nonzero(~pr.fixed)[0]

    def grad(s):
        h, g = 1e-6, []
        for c in free:
            for k in range(6):
                dx = np.zeros((len(pr.fixed), 6)); dx[c, k] = h
                g.append((pr.robust_chi2(pr.apply(s, dx, np.zeros_like(s.points))) -
                          pr.robust_chi2(pr.apply(s, -dx, np.zeros_like(s.points)))) / (2 * h))
        return np.array(g)
    g0 = grad(BA.BAProblem(**d).state)
    g1 = grad(pr.state)
    assert np.linalg.norm(g1) < 1e-3 * np.linalg.norm(g0)


def test_scipy_cannot_improve_on_the_result():
    scipy_opt = pytest.importorskip("scipy.optimize")
    import cv2
    d = make_ba(3, 40, 11)
    pr = BA.BAProblem(**d)
    with np.errstate(all="ignore"):
        pr.optimize(120)
    s = pr.state
    best = pr.robust_chi2(s)
    free = [c for c in range(3) if not d["fixed"][c]]

    def unpack(x):
        R, t = s.R.copy(), s.t.copy()
        for k, c in enumerate(free):
            R[c] = cv2.Rodrigues(x[6 * k:6 * k + 3])[0]
            t[c] = x[6 * k + 3:6 * k + 6]
        return BA.BAState(R, t, x[6 * len(free):].reshape(-1, 3))

    def fun(x):
        e = pr.errors(unpack(x))
        return np.sqrt(BA.huber(e[:, 0] ** 2 + e[:, 1] ** 2)[0])
    x0 = np.concatenate([np.concatenate([cv2.Rodrigues(s.R[c])[0].ravel(), s.t[c]]) for c in free]
                        + [s.points.ravel()])
    r = scipy_opt.least_squares(fun, x0, method="trf", x_scale="jac", max_nfev=30)
    assert 2 * r.cost > best * (1 - 1e-5)
