"""oracle/cv_geometry.py — numpy restatement of OpenCV's two-view geometry used by the reference
(src/image_pair.py:280-339).  TEST INFRASTRUCTURE ONLY (see cv_restate.py header).

Upstream: modules/calib3d/src/five-point.cpp, ptsetreg.cpp, triangulate.cpp.
"""
import numpy as np

from .cv_restate import CvRNG, ransac_update_num_iters


# ------------------------------------------------------------------ normalisation + Sampson
def normalize_points(pts, K):
    """findEssentialMat: x' = (x - c) / f per axis."""
    pts = np.asarray(pts, np.float64)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    return np.stack([(pts[:, 0] - cx) / fx, (pts[:, 1] - cy) / fy], 1)


def sampson_error(E, x1, x2):
    """EMEstimatorCallback::computeError: float32((x2'Ex1)^2 / (a+b+c+d)), no FMA, OpenCV's
    summation order (Matx product accumulates k = 0,1,2)."""
    E = np.asarray(E, np.float64).reshape(3, 3)
    X1 = np.c_[x1, np.ones(len(x1))]
    X2 = np.c_[x2, np.ones(len(x2))]
    Ex1 = np.stack([(E[r, 0] * X1[:, 0] + E[r, 1] * X1[:, 1]) + E[r, 2] * X1[:, 2]
                    for r in range(3)], 1)
    Etx2 = np.stack([(E[0, c] * X2[:, 0] + E[1, c] * X2[:, 1]) + E[2, c] * X2[:, 2]
                     for c in range(3)], 1)
    x2tEx1 = (X2[:, 0] * Ex1[:, 0] + X2[:, 1] * Ex1[:, 1]) + X2[:, 2] * Ex1[:, 2]
    a = Ex1[:, 0] * Ex1[:, 0]
    b = Ex1[:, 1] * Ex1[:, 1]
    c = Etx2[:, 0] * Etx2[:, 0]
    d = Etx2[:, 1] * Etx2[:, 1]
    with np.errstate(divide="ignore", invalid="ignore"):
        return (x2tEx1 * x2tEx1 / (a + b + c + d)).astype(np.float32)


def sampson_mask(E, x1, x2, thr):
    """findInliers: err <= float32(thr^2)."""
    return sampson_error(E, x1, x2) <= np.float32(thr * thr)


# ------------------------------------------------------------------ five-point minimal solver
# monomials of total degree <= 3 in (x,y,z), as exponent triples, in Nister's elimination order
_MONO = [(3, 0, 0), (0, 3, 0), (2, 1, 0), (1, 2, 0), (2, 0, 1), (2, 0, 0), (0, 2, 1), (0, 2, 0),
         (1, 1, 1), (1, 1, 0),
         (1, 0, 2), (1, 0, 1), (1, 0, 0), (0, 1, 2), (0, 1, 1), (0, 1, 0), (0, 0, 3), (0, 0, 2),
         (0, 0, 1), (0, 0, 0)]
_MIDX = {m: i for i, m in enumerate(_MONO)}


def _pmul(p, q):
    out = {}
    for (a, ca) in p.items():
        for (b, cb) in q.items():
            k = (a[0] + b[0], a[1] + b[1], a[2] + b[2])
            out[k] = out.get(k, 0.0) + ca * cb
    return out


def _padd(p, q, s=1.0):
    out = dict(p)
    for k, v in q.items():
        out[k] = out.get(k, 0.0) + s * v
    return out


def fivept_constraints(X, Y, Z, W):
    """10x20 coefficient matrix of det(E)=0 and 2EE'E - tr(EE')E = 0 for E = xX+yY+zZ+W."""
    Ep = [[{(1, 0, 0): X[i, j], (0, 1, 0): Y[i, j], (0, 0, 1): Z[i, j], (0, 0, 0): W[i, j]}
           for j in range(3)] for i in range(3)]
    EEt = [[None] * 3 for _ in range(3)]
    for i in range(3):
        for j in range(3):
            acc = {}
            for k in range(3):
                acc = _padd(acc, _pmul(Ep[i][k], Ep[j][k]))
            EEt[i][j] = acc
    tr = _padd(_padd(EEt[0][0], EEt[1][1]), EEt[2][2])
    L = [[_padd(EEt[i][j], tr, -0.5) if i == j else EEt[i][j] for j in range(3)]
         for i in range(3)]
    rows = []
    det = {}
    det = _padd(det, _pmul(Ep[0][0], _padd(_pmul(Ep[1][1], Ep[2][2]), _pmul(Ep[1][2], Ep[2][1]), -1)))
    det = _padd(det, _pmul(Ep[0][1], _padd(_pmul(Ep[1][0], Ep[2][2]), _pmul(Ep[1][2], Ep[2][0]), -1)), -1)
    det = _padd(det, _pmul(Ep[0][2], _padd(_pmul(Ep[1][0], Ep[2][1]), _pmul(Ep[1][1], Ep[2][0]), -1)))
    rows.append(det)
    for i in range(3):
        for j in range(3):
            acc = {}
            for k in range(3):
                acc = _padd(acc, _pmul(L[i][k], Ep[k][j]))
            rows.append(acc)
    A = np.zeros((10, 20))
    for r, p in enumerate(rows):
        for m, v in p.items():
            A[r, _MIDX[m]] = v
    return A


def _polish(A, x, y, z, iters=2):
    """Gauss-Newton on the ten cubic constraints (restores accuracy lost in the degree-10
    polynomial for ill-conditioned samples)."""
    ex = np.array(_MONO, dtype=np.float64)
    for _ in range(iters):
        v = np.array([x, y, z])
        mono = np.prod(v[None, :] ** ex, axis=1)
        f = A @ mono
        J = np.zeros((10, 3))
        for d in range(3):
            e2 = ex.copy()
            coef = e2[:, d].copy()
            e2[:, d] = np.maximum(e2[:, d] - 1, 0)
            J[:, d] = A @ (coef * np.prod(v[None, :] ** e2, axis=1))
        try:
            step = np.linalg.lstsq(J, -f, rcond=None)[0]
        except np.linalg.LinAlgError:
            break
        if not np.all(np.isfinite(step)):
            break
        x, y, z = x + step[0], y + step[1], z + step[2]
    return x, y, z


def fivept_solve(x1, x2):
    """All real essential matrices through 5 normalised correspondences (Nister's method as in
    OpenCV five-point.cpp runKernel): returns [n,3,3], ||E||_F = 1, sorted by the hidden
    variable z."""
    x1 = np.asarray(x1, np.float64)
    x2 = np.asarray(x2, np.float64)
    Q = np.stack([x2[:, 0] * x1[:, 0], x2[:, 0] * x1[:, 1], x2[:, 0],
                  x2[:, 1] * x1[:, 0], x2[:, 1] * x1[:, 1], x2[:, 1],
                  x1[:, 0], x1[:, 1], np.ones(5)], 1)
    _, _, Vt = np.linalg.svd(Q)
    X, Y, Z, W = (Vt[k].reshape(3, 3) for k in (5, 6, 7, 8))
    A = fivept_constraints(X, Y, Z, W)
    try:
        Rm = np.linalg.solve(A[:, :10], A[:, 10:])   # reduced rows: [I | Rm]
    except np.linalg.LinAlgError:
        return np.zeros((0, 3, 3))

    def zpoly(row, lo, n):  # coefficients highest power first
        return np.asarray(Rm[row, lo:lo + n])

    B = [[None] * 3 for _ in range(3)]
    for bi, (p, q) in enumerate([(4, 5), (6, 7), (8, 9)]):
        for c, (lo, n) in enumerate([(0, 3), (3, 3), (6, 4)]):
            a = zpoly(p, lo, n)
            b = zpoly(q, lo, n)
            B[bi][c] = np.polysub(a, np.polymul(b, [1.0, 0.0]))
    pm = np.polymul
    det = np.polysub(
        np.polyadd(pm(B[0][0], np.polysub(pm(B[1][1], B[2][2]), pm(B[1][2], B[2][1]))),
                   pm(B[0][2], np.polysub(pm(B[1][0], B[2][1]), pm(B[1][1], B[2][0])))),
        pm(B[0][1], np.polysub(pm(B[1][0], B[2][2]), pm(B[1][2], B[2][0]))))
    roots = np.roots(det)
    out = []
    for r in sorted(roots[np.abs(roots.imag) <= 1e-10].real):
        Bz = np.array([[np.polyval(B[i][j], r) for j in range(3)] for i in range(3)])
        _, _, vt = np.linalg.svd(Bz)
        v = vt[2]
        if abs(v[2]) < 1e-10:
            continue
        x, y, r = _polish(A, v[0] / v[2], v[1] / v[2], r)
        E = x * X + y * Y + r * Z + W
        out.append(E / np.linalg.norm(E))
    return np.array(out).reshape(-1, 3, 3)


# ------------------------------------------------------------------ RANSAC (ptsetreg.cpp)
def find_essential_ransac(p1, p2, K, prob=0.99, threshold=1.0, max_iters=1000, adaptive=True,
                          solver=fivept_solve, return_trace=False):
    """cv2.findEssentialMat(p1, p2, K, FM_RANSAC, prob, threshold): replayed cv::RNG samples,
    strict-greater best update, adaptive iteration bound, no refit.
    Returns (E[3,3] or None, mask uint8 [M,1] in {0,1})."""
    p1 = np.asarray(p1, np.float64).reshape(-1, 2)
    p2 = np.asarray(p2, np.float64).reshape(-1, 2)
    count = len(p1)
    if count < 5:
        raise ValueError("findEssentialMat needs >= 5 points")
    x1, x2 = normalize_points(p1, K), normalize_points(p2, K)
    thr = threshold / ((K[0, 0] + K[1, 1]) / 2.0)
    if count == 5:
        Es = solver(x1, x2)
        if len(Es) == 0:
            return None, np.zeros((count, 1), np.uint8)
        return Es.reshape(-1, 3), np.ones((count, 1), np.uint8)
    rng = CvRNG()
    niters = max_iters
    best_E, best_mask, max_good = None, None, 0
    it = 0
    trace = []
    while it < niters:
        idx = []
        while len(idx) < 5:
            v = rng.uniform(0, count)
            if v not in idx:
                idx.append(v)
        Es = solver(x1[idx], x2[idx])
        for E in Es:
            mask = sampson_mask(E, x1, x2, thr)
            good = int(mask.sum())
            trace.append((it, good))
            if good > max(max_good, 4):
                best_E, best_mask, max_good = E, mask, good
                if adaptive:
                    niters = ransac_update_num_iters(prob, (count - good) / count, 5, niters)
        it += 1
    if best_E is None:
        out = (None, np.zeros((count, 1), np.uint8))
    else:
        out = (best_E, best_mask.astype(np.uint8).reshape(-1, 1))
    return out + ((it, trace),) if return_trace else out


# ------------------------------------------------------------------ triangulation (triangulate.cpp)
def triangulate_points(P1, P2, x1, x2):
    """cv2.triangulatePoints: per point the right singular vector of the smallest singular value
    of the 4x4 DLT matrix (unit norm, sign arbitrary).  x1,x2 [2,N] -> X [4,N]."""
    P1 = np.asarray(P1, np.float64)
    P2 = np.asarray(P2, np.float64)
    x1 = np.asarray(x1, np.float64)
    x2 = np.asarray(x2, np.float64)
    n = x1.shape[1]
    A = np.empty((n, 4, 4))
    A[:, 0] = x1[0][:, None] * P1[2] - P1[0]
    A[:, 1] = x1[1][:, None] * P1[2] - P1[1]
    A[:, 2] = x2[0][:, None] * P2[2] - P2[0]
    A[:, 3] = x2[1][:, None] * P2[2] - P2[1]
    _, _, Vt = np.linalg.svd(A)
    return Vt[:, 3, :].T.copy()


# ------------------------------------------------------------------ recoverPose (five-point.cpp)
def decompose_essential(E):
    U, _, Vt = np.linalg.svd(np.asarray(E, np.float64))
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    Wm = np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1.0]])
    return U @ Wm @ Vt, U @ Wm.T @ Vt, U[:, 2].copy()


def recover_pose(E, p1, p2, K, dist_thresh=50.0):
    """cv2.recoverPose(E, p1, p2, K): 4 candidates [R1|t],[R2|t],[R1|-t],[R2|-t], DLT on
    normalised points, cheirality masks, first max count wins.
    Returns (n_good, R, t[3,1], mask uint8 [N,1] in {0,255})."""
    x1 = normalize_points(p1, K).T
    x2 = normalize_points(p2, K).T
    R1, R2, t = decompose_essential(E)
    P0 = np.eye(3, 4)
    cands = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    masks = []
    for R, tt in cands:
        P = np.hstack([R, tt.reshape(3, 1)])
        Q = triangulate_points(P0, P, x1, x2)
        m = (Q[2] * Q[3]) > 0
        with np.errstate(divide="ignore", invalid="ignore"):
            Q = Q / Q[3]
            m &= Q[2] < dist_thresh
            Q2 = P @ Q
            m &= Q2[2] > 0
            m &= Q2[2] < dist_thresh
        masks.append(m)
    goods = [int(m.sum()) for m in masks]
    # OpenCV's if-chain: good1 >= good2,good3,good4 first, then good2, good3, good4
    g1, g2, g3, g4 = goods
    if g1 >= g2 and g1 >= g3 and g1 >= g4:
        k = 0
    elif g2 >= g1 and g2 >= g3 and g2 >= g4:
        k = 1
    elif g3 >= g1 and g3 >= g2 and g3 >= g4:
        k = 2
    else:
        k = 3
    R, tt = cands[k]
    return goods[k], R, tt.reshape(3, 1), (masks[k].astype(np.uint8) * 255).reshape(-1, 1)
