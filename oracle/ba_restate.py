"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the reference's local bundle adjustment.

Reference: Map.optimize_map (src/map.py:104-186): g2o SparseOptimizer, BlockSolverSE3 +
LinearSolverEigenSE3, OptimizationAlgorithmLevenberg, one CameraParameters(focal = K[0,0],
principal point, baseline 0), VertexSE3Expmap per camera (fixed flag), marginalised VertexPointXYZ
per point, EdgeProjectXYZ2UV + identity information + RobustKernelHuber (default delta 1) per
observation, optimize(40).

The arithmetic lives in the third-party dependency g2o (Pipfile.lock: g2o-python ==0.0.11), which
is ABSENT from /root/reference and not installable here, so this file restates its published
algorithm (g2o/core/optimization_algorithm_levenberg.cpp, block_solver.hpp (Schur complement),
base_binary_edge.hpp (robustified quadratic form: first-order weight only), robust_kernel_impl.cpp
(Huber), types/sba/types_six_dof_expmap.cpp (EdgeProjectXYZ2UV error + Jacobians,
CameraParameters::cam_map uses ONE focal length for both axes), types/slam3d/se3quat.h (exp map,
update = exp(dx) * T with dx = (omega, upsilon))).
PARITY UNPINNED against g2o itself (the reference holds no BA fixture: test.g2o is a one-line
parameter dump).  Independent checks in tests/test_oracle_ba.py: analytic Jacobians against central
differences of the error under the update rule, the Schur solution against the full damped system,
first-order optimality of the result, and scipy.optimize.least_squares started from the result
being unable to lower the same robust cost.  Only tests/ may import this module.
"""
import numpy as np

HUBER_DELTA = 1.0
TAU = 1e-5
GOOD_STEP_LOWER, GOOD_STEP_UPPER = 1.0 / 3.0, 2.0 / 3.0
MAX_TRIALS = 10
DBL_MAX = np.finfo(np.float64).max


def skew(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], np.float64)


def se3_exp(u):
    """SE3Quat::exp: u = (omega[3], upsilon[3]) -> (R, t)"""
    omega, ups = np.asarray(u[:3], np.float64), np.asarray(u[3:], np.float64)
    theta = np.sqrt(omega @ omega)
    Om = skew(omega)
    Om2 = Om @ Om
    if theta < 0.00001:
        R = np.eye(3) + Om + 0.5 * Om2
        V = np.eye(3) + 0.5 * Om + (1.0 / 6.0) * Om2
    else:
        R = np.eye(3) + np.sin(theta) / theta * Om + (1 - np.cos(theta)) / (theta * theta) * Om2
        V = np.eye(3) + (1 - np.cos(theta)) / (theta * theta) * Om \
            + (theta - np.sin(theta)) / (theta ** 3) * Om2
    return R, V @ ups


def huber(e2, delta=HUBER_DELTA):
    """RobustKernelHuber::robustify -> rho[0], rho[1]"""
    e2 = np.asarray(e2, np.float64)
    d2 = delta * delta
    s = np.sqrt(np.maximum(e2, 1e-300))
    out = e2 > d2
    return np.where(out, 2 * s * delta - d2, e2), np.where(out, delta / s, 1.0)


def inv3(V):
    """3x3 inverses by cofactors / determinant (what Eigen's fixed-size inverse() does): a
    singular block gives inf / nan instead of raising, and the LM loop then rejects the step."""
    a, b, c = V[:, 0, 0], V[:, 0, 1], V[:, 0, 2]
    d, e, g = V[:, 1, 1], V[:, 1, 2], V[:, 2, 2]
    c00, c01, c02 = d * g - e * e, c * e - b * g, b * e - c * d
    with np.errstate(all="ignore"):
        idet = 1.0 / (a * c00 + b * c01 + c * c02)
        out = np.empty_like(V)
        out[:, 0, 0], out[:, 0, 1], out[:, 0, 2] = c00 * idet, c01 * idet, c02 * idet
        out[:, 1, 1], out[:, 1, 2], out[:, 2, 2] = (a * g - c * c) * idet, (b * c - a * e) * idet, \
            (a * d - b * b) * idet
    out[:, 1, 0], out[:, 2, 0], out[:, 2, 1] = out[:, 0, 1], out[:, 0, 2], out[:, 1, 2]
    return out


class BAState:
    def __init__(self, R, t, points):
        self.R, self.t, self.points = R.copy(), t.copy(), points.copy()

    def copy(self):
        return BAState(self.R, self.t, self.points)


class BAProblem:
    def __init__(self, R, t, fixed, points, obs_cam, obs_pt, obs_xy, f, cx, cy):
        self.state = BAState(np.asarray(R, np.float64).reshape(-1, 3, 3),
                             np.asarray(t, np.float64).reshape(-1, 3),
                             np.asarray(points, np.float64).reshape(-1, 3))
        self.fixed = np.asarray(fixed, bool).ravel()
        self.obs_cam = np.asarray(obs_cam, np.int64).ravel()
        self.obs_pt = np.asarray(obs_pt, np.int64).ravel()
        self.obs_xy = np.asarray(obs_xy, np.float64).reshape(-1, 2)
        self.f, self.c = float(f), np.array([cx, cy], np.float64)
        C, P = len(self.state.R), len(self.state.points)
        self.cam_active = np.bincount(self.obs_cam, minlength=C) > 0
        self.pt_active = np.bincount(self.obs_pt, minlength=P) > 0

    # ---- EdgeProjectXYZ2UV
    def transformed(self, s):
        X = s.points[self.obs_pt]
        return np.einsum("mij,mj->mi", s.R[self.obs_cam], X) + s.t[self.obs_cam]

    def errors(self, s):
        Xc = self.transformed(s)
        return self.obs_xy - (Xc[:, :2] / Xc[:, 2:3] * self.f + self.c)

    def robust_chi2(self, s):
        e = self.errors(s)
        return float(np.sum(huber(e[:, 0] ** 2 + e[:, 1] ** 2)[0]))

    def build_system(self, s):
        Xc = self.transformed(s)
        x, y, z = Xc[:, 0], Xc[:, 1], Xc[:, 2]
        z2 = z * z
        f = self.f
        e = self.obs_xy - (Xc[:, :2] / Xc[:, 2:3] * f + self.c)
        M = len(x)
        tmp = np.zeros((M, 2, 3))
        tmp[:, 0, 0] = f
        tmp[:, 0, 2] = -x / z * f
        tmp[:, 1, 1] = f
        tmp[:, 1, 2] = -y / z * f
        A = (-1.0 / z)[:, None, None] * np.einsum("mij,mjk->mik", tmp, s.R[self.obs_cam])   # d e / d point
        B = np.zeros((M, 2, 6))                                                             # d e / d pose
        B[:, 0, 0] = x * y / z2 * f
        B[:, 0, 1] = -(1 + x * x / z2) * f
        B[:, 0, 2] = y / z * f
        B[:, 0, 3] = -1.0 / z * f
        B[:, 0, 5] = x / z2 * f
        B[:, 1, 0] = (1 + y * y / z2) * f
        B[:, 1, 1] = -x * y / z2 * f
        B[:, 1, 2] = -x / z * f
        B[:, 1, 4] = -1.0 / z * f
        B[:, 1, 5] = y / z2 * f
        w = huber(e[:, 0] ** 2 + e[:, 1] ** 2)[1]
        C, P = len(s.R), len(s.points)
        free = ~self.fixed[self.obs_cam]
        Hll, bl = np.zeros((P, 3, 3)), np.zeros((P, 3))
        Hpp, bp = np.zeros((C, 6, 6)), np.zeros((C, 6))
        np.add.at(Hll, self.obs_pt, w[:, None, None] * np.einsum("mki,mkj->mij", A, A))
        np.add.at(bl, self.obs_pt, -w[:, None] * np.einsum("mki,mk->mi", A, e))
        np.add.at(Hpp, self.obs_cam[free], (w[:, None, None] * np.einsum("mki,mkj->mij", B, B))[free])
        np.add.at(bp, self.obs_cam[free], (-w[:, None] * np.einsum("mki,mk->mi", B, e))[free])
        Hpl = w[:, None, None] * np.einsum("mki,mkj->mij", B, A)       # [M, 6, 3], pose x point
        Hpl[~free] = 0.0
        return Hpp, bp, Hll, bl, Hpl

    def solve(self, sysm, lam):
        """BlockSolver::solve with the Schur complement; -> (xp [C,6], xl [P,3], ok)"""
        Hpp, bp, Hll, bl, Hpl = sysm
        C, P = len(Hpp), len(Hll)
        cam_free = self.cam_active & ~self.fixed
        V = Hll + lam * np.eye(3)
        V[~self.pt_active] = np.eye(3)
        Vinv = inv3(V)
        Y = np.einsum("mij,mjk->mik", Hpl, Vinv[self.obs_pt])          # Hpl * Hll^-1
        S = np.zeros((C, 6, C, 6))
        for c in range(C):
            S[c, :, c, :] = Hpp[c] + lam * np.eye(6) if cam_free[c] else np.eye(6)
        bs = np.where(cam_free[:, None], bp, 0.0).copy()
        np.add.at(bs, self.obs_cam, -np.einsum("mij,mj->mi", Y, bl[self.obs_pt]))
        order = np.argsort(self.obs_pt, kind="stable")
        start = np.searchsorted(self.obs_pt[order], np.arange(P + 1))
        for p in range(P):
            idx = order[start[p]:start[p + 1]]
            for i in idx:
                if self.fixed[self.obs_cam[i]]:
                    continue
                for j in idx:
                    if self.fixed[self.obs_cam[j]]:
                        continue
                    S[self.obs_cam[i], :, self.obs_cam[j], :] -= Y[i] @ Hpl[j].T
        S2 = S.reshape(6 * C, 6 * C)
        if not np.all(np.isfinite(S2)):
            return None, None, False
        try:
            L = np.linalg.cholesky(S2)
        except np.linalg.LinAlgError:
            return None, None, False
        xp = np.linalg.solve(L.T, np.linalg.solve(L, bs.reshape(-1))).reshape(C, 6)
        xp[~cam_free] = 0.0
        cp = bl.copy()
        np.add.at(cp, self.obs_pt, -np.einsum("mji,mj->mi", Hpl, xp[self.obs_cam]))
        xl = np.einsum("pij,pj->pi", Vinv, cp)
        xl[~self.pt_active] = 0.0
        return xp, xl, True

    def apply(self, s, xp, xl):
        n = s.copy()
        for c in range(len(s.R)):
            if self.cam_active[c] and not self.fixed[c]:
                dR, dt = se3_exp(xp[c])
                n.R[c] = dR @ s.R[c]
                n.t[c] = dR @ s.t[c] + dt
        n.points = s.points + xl
        return n

    def lambda_init(self, sysm):
        Hpp, _bp, Hll, _bl, _ = sysm
        cam_free = self.cam_active & ~self.fixed
        m = 0.0
        if cam_free.any():
            m = max(m, np.abs(np.einsum("cii->ci", Hpp[cam_free])).max())
        if self.pt_active.any():
            m = max(m, np.abs(np.einsum("pii->pi", Hll[self.pt_active])).max())
        return TAU * m

    def optimize(self, iterations=40):
        """SparseOptimizer::optimize + OptimizationAlgorithmLevenberg::solve.  Returns a log:
        list of (iteration, chi2 after, lambda after, trials)."""
        s = self.state
        lam, ni = 0.0, 2.0
        log = []
        for it in range(iterations):
            current = self.robust_chi2(s)
            temp = current
            sysm = self.build_system(s)
            if it == 0:
                lam, ni = self.lambda_init(sysm), 2.0
            rho, qmax = 0.0, 0
            while True:
                xp, xl, ok = self.solve(sysm, lam)
                if ok:
                    trial = self.apply(s, xp, xl)
                    temp = self.robust_chi2(trial)
                    scale = float(np.sum(xp * (lam * xp + sysm[1])) + np.sum(xl * (lam * xl + sysm[3])))
                else:
                    temp, scale = DBL_MAX, 0.0
                scale += 1e-3
                rho = (current - temp) / scale
                if rho > 0 and np.isfinite(temp):
                    alpha = 1.0 - (2 * rho - 1) ** 3
                    alpha = min(alpha, GOOD_STEP_UPPER)
                    lam *= max(GOOD_STEP_LOWER, alpha)
                    ni = 2.0
                    current = temp
                    s = trial
                else:
                    lam *= ni
                    ni *= 2
                    if not np.isfinite(lam):
                        break
                qmax += 1
                if not (rho < 0 and qmax < MAX_TRIALS):
                    break
            log.append((it, current, lam, qmax))
            if qmax == MAX_TRIALS or rho == 0 or not np.isfinite(lam):
                break
        self.state = s
        return log
