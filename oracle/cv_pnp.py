"""oracle/cv_pnp.py — numpy restatement of cv2.solvePnPRansac with default arguments, the call
the reference makes at src/visual_slam.py:231-235 (then cv2.Rodrigues at :243).
TEST INFRASTRUCTURE ONLY (see cv_restate.py header).

Upstream files restated: modules/calib3d/src/solvepnp.cpp (solvePnPRansac, PnPRansacCallback),
modules/calib3d/src/epnp.cpp (EPnP minimal solver on 5-point samples), ptsetreg.cpp (RANSAC loop,
same cv::RNG replay as findEssentialMat), calibration.cpp (iterative refit = Levenberg-Marquardt on
the reprojection error starting from the RANSAC model).
Behaviour pinned by tests/test_oracle_pnp.py against cv2 itself:
  * points are converted to float32 first (solvePnPRansac does opoints.convertTo(CV_32F));
  * RANSAC: 5-point samples, EPnP model, err = float32 squared reprojection distance,
    inlier iff err <= 64 (8 px), 100 iterations max with the adaptive bound, strict-greater update;
  * the returned inliers are those of the RANSAC model, the returned pose is the refit on them.
"""
import numpy as np

from .cv_restate import CvRNG, ransac_update_num_iters


def rodrigues_to_R(rvec):
    r = np.asarray(rvec, np.float64).ravel()
    th = np.linalg.norm(r)
    if th < 2.220446049250313e-16:
        return np.eye(3)
    k = r / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * K


def R_to_rodrigues(R):
    """cv2.Rodrigues(R) for a proper rotation (matrix -> vector)."""
    R = np.asarray(R, np.float64)
    U, _, Vt = np.linalg.svd(R)
    R = U @ Vt
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = np.sqrt((r @ r) * 0.25)
    c = np.clip((np.trace(R) - 1) * 0.5, -1.0, 1.0)
    th = np.arccos(c)
    if s < 1e-5:
        if c > 0:
            return np.zeros(3)
        t = np.sqrt(np.maximum((np.diag(R) + 1) * 0.5, 0))
        t[1] *= -1 if R[0, 1] < 0 else 1
        t[2] *= -1 if R[0, 2] < 0 else 1
        if abs(t[0]) < abs(t[1]) and abs(t[0]) < abs(t[2]) and (R[1, 2] > 0) != (t[1] * t[2] > 0):
            t[2] = -t[2]
        return t * (th / np.linalg.norm(t))
    return r * (0.5 / s) * th


def project(X, rvec, tvec, K):
    R = rodrigues_to_R(rvec)
    Xc = X @ R.T + np.asarray(tvec, np.float64).ravel()
    return np.stack([K[0, 0] * Xc[:, 0] / Xc[:, 2] + K[0, 2],
                     K[1, 1] * Xc[:, 1] / Xc[:, 2] + K[1, 2]], 1)


# ------------------------------------------------------------------ EPnP (epnp.cpp)
def epnp(X, x, K):
    """EPnP pose from n >= 4 correspondences: returns (R, t)."""
    X = np.asarray(X, np.float64)
    x = np.asarray(x, np.float64)
    n = len(X)
    fu, fv, uc, vc = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    # control points: centroid + principal directions
    c0 = X.mean(0)
    PW0 = X - c0
    w, V = np.linalg.eigh(PW0.T @ PW0)
    order = np.argsort(w)[::-1]
    w, V = w[order], V[:, order]
    cws = np.vstack([c0] + [c0 + np.sqrt(max(w[i], 0) / n) * V[:, i] for i in range(3)])
    CC = (cws[1:] - cws[0]).T
    alphas = np.empty((n, 4))
    alphas[:, 1:] = (np.linalg.pinv(CC) @ (X - cws[0]).T).T
    alphas[:, 0] = 1 - alphas[:, 1:].sum(1)
    M = np.zeros((2 * n, 12))
    for j in range(4):
        M[0::2, 3 * j] = alphas[:, j] * fu
        M[0::2, 3 * j + 2] = alphas[:, j] * (uc - x[:, 0])
        M[1::2, 3 * j + 1] = alphas[:, j] * fv
        M[1::2, 3 * j + 2] = alphas[:, j] * (vc - x[:, 1])
    ew, ev = np.linalg.eigh(M.T @ M)
    v = [ev[:, i] for i in range(4)]           # ascending: v[0] is the smallest
    pairs = [(0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)]
    dv = [[vi[3 * a:3 * a + 3] - vi[3 * b:3 * b + 3] for (a, b) in pairs] for vi in v]
    L = np.zeros((6, 10))
    for i in range(6):
        d = [dv[k][i] for k in range(4)]
        L[i] = [d[0] @ d[0], 2 * d[0] @ d[1], d[1] @ d[1], 2 * d[0] @ d[2], 2 * d[1] @ d[2],
                d[2] @ d[2], 2 * d[0] @ d[3], 2 * d[1] @ d[3], 2 * d[2] @ d[3], d[3] @ d[3]]
    rho = np.array([np.sum((cws[a] - cws[b]) ** 2) for (a, b) in pairs])

    def lstsq(A, b):
        return np.linalg.lstsq(A, b, rcond=None)[0]

    def approx1():
        b4 = lstsq(L[:, [0, 1, 3, 6]], rho)
        if b4[0] < 0:
            b0 = np.sqrt(-b4[0])
            return np.array([b0, -b4[1] / b0, -b4[2] / b0, -b4[3] / b0])
        b0 = np.sqrt(b4[0])
        return np.array([b0, b4[1] / b0, b4[2] / b0, b4[3] / b0])

    def approx2():
        b3 = lstsq(L[:, [0, 1, 2]], rho)
        if b3[0] < 0:
            b = [np.sqrt(-b3[0]), np.sqrt(-b3[2]) if b3[2] < 0 else 0.0]
        else:
            b = [np.sqrt(b3[0]), np.sqrt(b3[2]) if b3[2] > 0 else 0.0]
        if b3[1] < 0:
            b[0] = -b[0]
        return np.array([b[0], b[1], 0.0, 0.0])

    def approx3():
        b5 = lstsq(L[:, [0, 1, 2, 3, 4]], rho)
        if b5[0] < 0:
            b = [np.sqrt(-b5[0]), np.sqrt(-b5[2]) if b5[2] < 0 else 0.0]
        else:
            b = [np.sqrt(b5[0]), np.sqrt(b5[2]) if b5[2] > 0 else 0.0]
        if b5[1] < 0:
            b[0] = -b[0]
        return np.array([b[0], b[1], b5[3] / b[0], 0.0])

    def gauss_newton(b):
        b = b.copy()
        for _ in range(5):
            A = np.stack([2 * L[:, 0] * b[0] + L[:, 1] * b[1] + L[:, 3] * b[2] + L[:, 6] * b[3],
                          L[:, 1] * b[0] + 2 * L[:, 2] * b[1] + L[:, 4] * b[2] + L[:, 7] * b[3],
                          L[:, 3] * b[0] + L[:, 4] * b[1] + 2 * L[:, 5] * b[2] + L[:, 8] * b[3],
                          L[:, 6] * b[0] + L[:, 7] * b[1] + L[:, 8] * b[2] + 2 * L[:, 9] * b[3]], 1)
            r = rho - (L[:, 0] * b[0] * b[0] + L[:, 1] * b[0] * b[1] + L[:, 2] * b[1] * b[1] +
                       L[:, 3] * b[0] * b[2] + L[:, 4] * b[1] * b[2] + L[:, 5] * b[2] * b[2] +
                       L[:, 6] * b[0] * b[3] + L[:, 7] * b[1] * b[3] + L[:, 8] * b[2] * b[3] +
                       L[:, 9] * b[3] * b[3])
            b = b + lstsq(A, r)
        return b

    def R_and_t(b):
        ccs = sum(b[j] * v[j] for j in range(4)).reshape(4, 3)
        pcs = alphas @ ccs
        if pcs[0, 2] < 0:
            ccs, pcs = -ccs, -pcs
        pc0, pw0 = pcs.mean(0), X.mean(0)
        ABt = (pcs - pc0).T @ (X - pw0)
        U, _, Vt = np.linalg.svd(ABt)
        R = U @ Vt
        if np.linalg.det(R) < 0:
            R[2] = -R[2]
        t = pc0 - R @ pw0
        Xc = X @ R.T + t
        ue = uc + fu * Xc[:, 0] / Xc[:, 2]
        ve = vc + fv * Xc[:, 1] / Xc[:, 2]
        return R, t, np.mean(np.sqrt((x[:, 0] - ue) ** 2 + (x[:, 1] - ve) ** 2))

    sols = [R_and_t(gauss_newton(f())) for f in (approx1, approx2, approx3)]
    N = 0
    if sols[1][2] < sols[0][2]:
        N = 1
    if sols[2][2] < sols[N][2]:
        N = 2
    return sols[N][0], sols[N][1]


# ------------------------------------------------------------------ iterative refit (LM)
def refine_pose(X, x, K, rvec, tvec, iters=30):
    """solvePnP(SOLVEPNP_ITERATIVE, useExtrinsicGuess=True): minimise the squared reprojection
    error over (rvec, tvec) from the given start (Levenberg-Marquardt; run to convergence)."""
    p = np.concatenate([np.asarray(rvec, np.float64).ravel(), np.asarray(tvec, np.float64).ravel()])

    def resid(p):
        return (project(X, p[:3], p[3:], K) - x).ravel()

    lam = 1e-3
    r = resid(p)
    for _ in range(iters):
        J = np.empty((len(r), 6))
        for k in range(6):
            h = 1e-6 * max(1.0, abs(p[k]))
            d = np.zeros(6)
            d[k] = h
            J[:, k] = (resid(p + d) - resid(p - d)) / (2 * h)
        A, g = J.T @ J, J.T @ r
        while True:
            step = np.linalg.solve(A + lam * np.diag(np.diag(A)), -g)
            r2 = resid(p + step)
            if r2 @ r2 <= r @ r or lam > 1e12:
                break
            lam *= 10
        if r2 @ r2 <= r @ r:
            p, r = p + step, r2
            lam = max(lam / 10, 1e-12)
        if np.linalg.norm(step) < 1e-14 * max(1.0, np.linalg.norm(p)):
            break
    return p[:3].copy(), p[3:].copy()


def p3p_solutions(X3, y3):
    """All real solutions (R, t) of the perspective-three-point problem: X3 [3,3] object points,
    y3 [3,3] unit bearing vectors of their images.  Grunert's quartic in v = s3 / s1 (Haralick et
    al. 1994, eq. 6-9; OpenCV's calib3d/src/p3p.cpp solves the same system after Gao et al., any
    exact P3P has the same solution set), roots by numpy, depths polished by Newton on the three
    cosine-law equations, pose from the two point triads."""
    X3, y3 = np.asarray(X3, np.float64), np.asarray(y3, np.float64)
    a2 = ((X3[1] - X3[2]) ** 2).sum()
    b2 = ((X3[0] - X3[2]) ** 2).sum()
    c2 = ((X3[0] - X3[1]) ** 2).sum()
    ca, cb, cg = y3[1] @ y3[2], y3[0] @ y3[2], y3[0] @ y3[1]
    if min(a2, b2, c2) <= 0:
        return []
    q = (a2 - c2) / b2
    p_ = (a2 + c2) / b2
    A4 = (q - 1) ** 2 - 4 * c2 / b2 * ca * ca
    A3 = 4 * (q * (1 - q) * cb - (1 - p_) * ca * cg + 2 * c2 / b2 * ca * ca * cb)
    A2 = 2 * (q * q - 1 + 2 * q * q * cb * cb + 2 * (b2 - c2) / b2 * ca * ca
              - 4 * p_ * ca * cb * cg + 2 * (b2 - a2) / b2 * cg * cg)
    A1 = 4 * (-q * (1 + q) * cb + 2 * a2 / b2 * cg * cg * cb - (1 - p_) * ca * cg)
    A0 = (1 + q) ** 2 - 4 * a2 / b2 * cg * cg
    sols = []
    for v in np.roots([A4, A3, A2, A1, A0]):
        if abs(v.imag) > 1e-7 * max(1.0, abs(v.real)) or v.real <= 0:
            continue
        v = v.real
        den = 2 * (cg - v * ca)
        if abs(den) < 1e-14:
            continue
        u = ((q - 1) * v * v - 2 * q * cb * v + 1 + q) / den
        d1 = 1 + v * v - 2 * v * cb
        if u <= 0 or d1 <= 0:
            continue
        s = np.array([1.0, u, v]) * np.sqrt(b2 / d1)
        for _ in range(10):                        # Newton polish on the cosine-law system
            f = np.array([s[1] ** 2 + s[2] ** 2 - 2 * s[1] * s[2] * ca - a2,
                          s[0] ** 2 + s[2] ** 2 - 2 * s[0] * s[2] * cb - b2,
                          s[0] ** 2 + s[1] ** 2 - 2 * s[0] * s[1] * cg - c2])
            J = np.array([[0, 2 * s[1] - 2 * s[2] * ca, 2 * s[2] - 2 * s[1] * ca],
                          [2 * s[0] - 2 * s[2] * cb, 0, 2 * s[2] - 2 * s[0] * cb],
                          [2 * s[0] - 2 * s[1] * cg, 2 * s[1] - 2 * s[0] * cg, 0]])
            try:
                s = s - np.linalg.solve(J, f)
            except np.linalg.LinAlgError:
                break
        if np.any(s <= 0) or np.abs(f).max() > 1e-6 * max(a2, b2, c2):
            continue
        C = y3 * s[:, None]                        # the three points in the camera frame

        def frame(P):
            e1 = P[1] - P[0]
            e1 = e1 / np.linalg.norm(e1)
            e3 = np.cross(e1, P[2] - P[0])
            e3 = e3 / np.linalg.norm(e3)
            return np.stack([e1, np.cross(e3, e1), e3], 1)
        R = frame(C) @ frame(X3).T
        t = C[0] - R @ X3[0]
        if not any(np.abs(R - R0).max() < 1e-9 and np.abs(t - t0).max() < 1e-9 for R0, t0 in sols):
            sols.append((R, t))
    return sols


def solve_p3p_4points(X, x, K):
    """cv2.solvePnP(..., SOLVEPNP_P3P) on exactly 4 points = what solvePnPRansac does when
    npoints == 4 (solvepnp.cpp: model_points == npoints -> solvePnP with the RANSAC kernel
    method): P3P on the first three points, the solution with the smallest squared reprojection
    error of the fourth wins."""
    Ki = np.linalg.inv(K)
    y = (Ki @ np.c_[x, np.ones(len(x))].T).T
    y3 = y[:3] / np.linalg.norm(y[:3], axis=1, keepdims=True)
    best, best_e = None, np.inf
    for R, t in p3p_solutions(X[:3], y3):
        pc = R @ X[3] + t
        if pc[2] == 0:
            continue
        pr = (K @ (pc / pc[2]))[:2]
        e = ((pr - x[3]) ** 2).sum()
        if e < best_e:
            best, best_e = (R, t), e
    return best



def solve_pnp_ransac(X, x, K, iterations=100, reproj_error=8.0, confidence=0.99,
                     return_model=False):
    """cv2.solvePnPRansac(X, x, K, zeros) with default arguments.
    Returns (retval, rvec [3,1], tvec [3,1], inliers int32 [k,1])."""
    X = np.asarray(X, np.float64).reshape(-1, 3).astype(np.float32).astype(np.float64)
    x = np.asarray(x, np.float64).reshape(-1, 2).astype(np.float32).astype(np.float64)
    n = len(X)
    if n < 4:
        raise ValueError("solvePnPRansac needs >= 4 points")
    if n == 4 or n == 5:
        # solvepnp.cpp: model_points == npoints -> plain solvePnP with the RANSAC kernel method
        # (P3P for 4 points, EPnP for 5), no RANSAC, no iterative refit, every point an inlier
        try:
            sol = solve_p3p_4points(X, x, K) if n == 4 else epnp(X, x, K)
        except np.linalg.LinAlgError:
            sol = None
        if sol is None:
            return False, np.zeros((3, 1)), np.zeros((3, 1)), np.zeros((0, 1), np.int32)
        out = (True, R_to_rodrigues(sol[0]).reshape(3, 1), np.asarray(sol[1], np.float64).reshape(3, 1),
               np.arange(n, dtype=np.int32).reshape(-1, 1))
        return out + ((out[1].ravel(), out[2].ravel()),) if return_model else out
    rng = CvRNG()
    thr = np.float32(reproj_error * reproj_error)
    niters = iterations
    best, best_mask, max_good = None, None, 0
    it = 0
    while it < niters:
        idx = []
        while len(idx) < 5:
            v = rng.uniform(0, n)
            if v not in idx:
                idx.append(v)
        try:
            R, t = epnp(X[idx], x[idx], K)
            rv = R_to_rodrigues(R)
            proj = project(X, rv, t, K).astype(np.float32)
            d = x.astype(np.float32) - proj
            err = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32)
            mask = err <= thr
            good = int(mask.sum())
            if good > max(max_good, 4):
                best, best_mask, max_good = (rv, t), mask, good
                niters = ransac_update_num_iters(confidence, (n - good) / n, 5, niters)
        except np.linalg.LinAlgError:
            pass
        it += 1
    if best is None:
        return False, np.zeros((3, 1)), np.zeros((3, 1)), np.zeros((0, 1), np.int32)
    inl = np.nonzero(best_mask)[0]
    rv, tv = refine_pose(X[inl], x[inl], K, best[0], best[1])
    out = (True, rv.reshape(3, 1), tv.reshape(3, 1), inl.astype(np.int32).reshape(-1, 1))
    return out + (best,) if return_model else out
