"""TEST INFRASTRUCTURE -- CPU restatement of two-view triangulation + bundle adjustment.  PARITY UNPINNED.

Only tests/ may import this module; the product (gtsfm_b200/) never does.

What it restates (numpy, dense normal equations, sized for a few hundred tracks):
  TwoViewEstimator.bundle_adjust                           gtsfm/two_view_estimator.py:137-208
  Point3dInitializer.triangulate (NO_RANSAC)               gtsfm/data_association/point3d_initializer.py:225-295
  TwoViewBundleAdjustment: graph, priors, LM, filtering    gtsfm/bundle/bundle_adjustment.py:106-357
The arithmetic of the reference sits in gtsam==4.2 (triangulatePoint3, GeneralSFMFactor2Cal3Bundler, PriorFactorPose3,
PriorFactorPoint3, LevenbergMarquardtOptimizer), which is neither vendored under /root/reference nor installed here, so this
file could not be checked against the reference's own output: it follows gtsam's PUBLISHED algorithm.

On purpose this is the reference's FORMULATION, not the kernel's: both camera poses are 6-DOF variables in a world frame, the
gauge is held by the soft priors the reference adds (PriorFactorPose3 on camera 0, sigma 0.1; PriorFactorPoint3 on the first
landmark, sigma 0.1), the calibrations are constants (their prior has sigma 1e-5), and the Jacobians come from finite
differences.  The CUDA kernel optimises the gauge-fixed 5-DOF relative pose with analytic Jacobians and a Schur complement; the
tests compare the two at convergence (relative rotation, translation direction, valid tracks).
"""
from typing import Optional, Tuple

import numpy as np

HUBER_K = 1.345
POSE_PRIOR_SIGMA = 0.1     # bundle_adjustment.py:47 (cam_pose3_prior_noise_sigma)
POINT_PRIOR_SIGMA = 0.1    # bundle_adjustment.py:242-246


def _skew(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], np.float64)


def expmap(w) -> np.ndarray:
    th = float(np.linalg.norm(w))
    K = _skew(w)
    if th < 1e-8:
        return np.eye(3) + K + 0.5 * K @ K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * K @ K


def logmap(R) -> np.ndarray:
    c = np.clip(0.5 * (np.trace(R) - 1.0), -1.0, 1.0)
    th = np.arccos(c)
    v = 0.5 * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    return v if th < 1e-8 else v * th / np.sin(th)


def project(cal, pc) -> Tuple[np.ndarray, np.ndarray]:
    """gtsam PinholeCamera<Cal3Bundler>.projectSafe of camera-frame points (N, 3): pixels (N, 2), in-front flags (N,)."""
    f, k1, k2, u0, v0 = cal
    z = pc[:, 2]
    ok = z > 1e-12
    zs = np.where(ok, z, 1.0)
    a, b = pc[:, 0] / zs, pc[:, 1] / zs
    r2 = a * a + b * b
    g = 1.0 + (k1 + k2 * r2) * r2
    return np.stack([f * g * a + u0, f * g * b + v0], 1), ok


def calibrate(cal, uv) -> np.ndarray:
    """gtsam Cal3Bundler.calibrate: fixed-point undistortion (exact when k1 = k2 = 0)."""
    f, k1, k2, u0, v0 = cal
    pn = (uv - np.array([u0, v0])) / f
    p = pn.copy()
    if k1 != 0.0 or k2 != 0.0:
        for _ in range(20):
            r2 = np.sum(p * p, 1, keepdims=True)
            p = pn / (1.0 + (k1 + k2 * r2) * r2)
    return p


def triangulate(cal1, cal2, R, t, uv1, uv2, reproj_thr: float, min_angle_deg: float = 0.0):
    """Every correspondence: DLT on the calibrated rays (cameras [I|0] and [R|t]), Gauss-Newton on the pixel error
    (triangulatePoint3(optimize=True)), cheirality in both cameras, both reprojection errors < reproj_thr.
    Returns points (N, 3) in camera-1 coordinates and the success flags."""
    n = len(uv1)
    x1, x2 = calibrate(cal1, uv1), calibrate(cal2, uv2)
    X = np.zeros((n, 3))
    good = np.zeros(n, bool)
    P2 = np.hstack([R, t.reshape(3, 1)])
    for i in range(n):
        A = np.array([[-1, 0, x1[i, 0], 0], [0, -1, x1[i, 1], 0], x2[i, 0] * P2[2] - P2[0], x2[i, 1] * P2[2] - P2[1]])
        q = np.linalg.svd(A)[2][-1]
        if abs(q[3]) < 1e-300:
            continue
        p = q[:3] / q[3]
        for _ in range(3):
            def res(pp):
                a, oa = project(cal1, pp[None])
                b, ob = project(cal2, (R @ pp + t)[None])
                return np.concatenate([(a[0] - uv1[i]) * oa[0], (b[0] - uv2[i]) * ob[0]])
            r0 = res(p)
            J = np.zeros((4, 3))
            for c in range(3):
                d = np.zeros(3); d[c] = 1e-6 * max(1.0, abs(p[c]))
                J[:, c] = (res(p + d) - res(p - d)) / (2 * d[c])
            try:
                p = p - np.linalg.solve(J.T @ J, J.T @ r0)
            except np.linalg.LinAlgError:
                break
        pc2 = R @ p + t
        if not (np.all(np.isfinite(p)) and p[2] > 0 and pc2[2] > 0):
            continue
        a, _ = project(cal1, p[None]); b, _ = project(cal2, pc2[None])
        if not (np.linalg.norm(a[0] - uv1[i]) < reproj_thr and np.linalg.norm(b[0] - uv2[i]) < reproj_thr):
            continue
        if min_angle_deg > 0:
            r1, r2 = p, R.T @ pc2
            cs = np.clip(r1 @ r2 / (np.linalg.norm(r1) * np.linalg.norm(r2)), -1, 1)
            if np.degrees(np.arccos(cs)) < min_angle_deg:
                continue
        X[i], good[i] = p, True
    return X, good


def _huber(e2: np.ndarray, k: float):
    """weight and loss of gtsam's mEstimator::Huber on the norm of a whitened 2-D factor error."""
    e = np.sqrt(e2)
    w = np.where(e <= k, 1.0, k / np.maximum(e, 1e-300))
    rho = np.where(e <= k, 0.5 * e2, k * (e - 0.5 * k))
    return w, rho


class _State:
    def __init__(self, R0, c0, R1, c1, X):
        self.R0, self.c0, self.R1, self.c1, self.X = R0, c0, R1, c1, X

    def retract(self, d):
        """Pose3 retract T <- T * (Exp(w), v) (rotation and translation in the body frame), points additive."""
        return _State(self.R0 @ expmap(d[0:3]), self.c0 + self.R0 @ d[3:6], self.R1 @ expmap(d[6:9]), self.c1 + self.R1 @ d[9:12],
                      self.X + d[12:].reshape(-1, 3))


def _reproj(cal, R, c, X, uv):
    p, ok = project(cal, (X - c) @ R)              # R^T (X - c) per row
    return (p - uv) * ok[:, None], ok              # a cheirality failure is a zero error with zero Jacobian (GeneralSFMFactor2)


def _errors(s: _State, cal1, cal2, uv1, uv2, X0_prior, k):
    e0, _ = _reproj(cal1, s.R0, s.c0, s.X, uv1)
    e1, _ = _reproj(cal2, s.R1, s.c1, s.X, uv2)
    pp = np.concatenate([logmap(s.R0), s.c0]) / POSE_PRIOR_SIGMA            # prior pose = identity (camera 0 of the initial data)
    pq = (s.X[0] - X0_prior) / POINT_PRIOR_SIGMA
    w0, r0 = _huber(np.sum(e0 * e0, 1), k)
    w1, r1 = _huber(np.sum(e1 * e1, 1), k)
    cost = float(r0.sum() + r1.sum() + 0.5 * pp @ pp + 0.5 * pq @ pq)
    return e0, e1, w0, w1, pp, pq, cost


def _linearise(s: _State, cal1, cal2, uv1, uv2, X0_prior, k):
    n = len(s.X)
    e0, e1, w0, w1, pp, pq, cost = _errors(s, cal1, cal2, uv1, uv2, X0_prior, k)
    J = np.zeros((4 * n + 9, 12 + 3 * n))
    r = np.concatenate([(e0 * np.sqrt(w0)[:, None]).ravel(), (e1 * np.sqrt(w1)[:, None]).ravel(), pp, pq])
    h = 1e-6
    rows0 = np.arange(2 * n).reshape(n, 2)
    rows1 = 2 * n + rows0
    for a in range(6):                              # central differences, every track at once
        d = np.zeros(6); d[a] = h
        Rp, cp = (s.R0 @ expmap(d[:3]), s.c0 + s.R0 @ d[3:]); Rm, cm = (s.R0 @ expmap(-d[:3]), s.c0 - s.R0 @ d[3:])
        g = (_reproj(cal1, Rp, cp, s.X, uv1)[0] - _reproj(cal1, Rm, cm, s.X, uv1)[0]) / (2 * h)
        J[rows0.ravel(), a] = (g * np.sqrt(w0)[:, None]).ravel()
        Rp, cp = (s.R1 @ expmap(d[:3]), s.c1 + s.R1 @ d[3:]); Rm, cm = (s.R1 @ expmap(-d[:3]), s.c1 - s.R1 @ d[3:])
        g = (_reproj(cal2, Rp, cp, s.X, uv2)[0] - _reproj(cal2, Rm, cm, s.X, uv2)[0]) / (2 * h)
        J[rows1.ravel(), 6 + a] = (g * np.sqrt(w1)[:, None]).ravel()
    for a in range(3):
        d = np.zeros(3); d[a] = h
        g0 = (_reproj(cal1, s.R0, s.c0, s.X + d, uv1)[0] - _reproj(cal1, s.R0, s.c0, s.X - d, uv1)[0]) / (2 * h) * np.sqrt(w0)[:, None]
        g1 = (_reproj(cal2, s.R1, s.c1, s.X + d, uv2)[0] - _reproj(cal2, s.R1, s.c1, s.X - d, uv2)[0]) / (2 * h) * np.sqrt(w1)[:, None]
        cols = 12 + 3 * np.arange(n) + a
        for c in range(2):
            J[rows0[:, c], cols] = g0[:, c]
            J[rows1[:, c], cols] = g1[:, c]
    # priors: d log(R0 Exp(w)) / dw ~ I near the prior, d(c0 + R0 v)/dv = R0
    J[4 * n:4 * n + 3, 0:3] = np.eye(3) / POSE_PRIOR_SIGMA
    J[4 * n + 3:4 * n + 6, 3:6] = s.R0 / POSE_PRIOR_SIGMA
    J[4 * n + 6:4 * n + 9, 12:15] = np.eye(3) / POINT_PRIOR_SIGMA
    return J, r, cost


def bundle_adjust(cal1, cal2, R, t, uv1, uv2, tri_thr: float = 100.0, ba_thr: Optional[float] = 0.5, max_iterations: int = 100,
                  huber_k: float = HUBER_K, min_angle_deg: float = 0.0, rel_tol: float = 1e-5, abs_tol: float = 1e-5):
    """TwoViewEstimator.bundle_adjust for one pair: (R, t) = i2Ri1, i2Ui1 from the verifier, uv1 / uv2 the verified
    correspondences in pixels.  Returns dict(R, t (unit), valid (N,) bool, triangulated (N,) bool, iterations, cost, status):
    status 0 = BA ran, 2 = nothing triangulated, 3 = no track passed the filter."""
    R = np.asarray(R, np.float64).reshape(3, 3); t = np.asarray(t, np.float64).reshape(3)
    uv1 = np.asarray(uv1, np.float64).reshape(-1, 2); uv2 = np.asarray(uv2, np.float64).reshape(-1, 2)
    X, tri = triangulate(cal1, cal2, R, t, uv1, uv2, tri_thr, min_angle_deg)
    n_all = len(uv1)
    if not tri.any():
        return dict(R=R, t=t, valid=np.zeros(n_all, bool), triangulated=tri, iterations=0, cost=0.0, status=2)
    a1, a2 = uv1[tri], uv2[tri]
    s = _State(np.eye(3), np.zeros(3), R.T.copy(), -R.T @ t, X[tri].copy())      # wTi1 = identity, wTi2 = i2Ti1^-1
    X0 = s.X[0].copy()
    lam, iters = 1e-5, 0                             # LevenbergMarquardtParams defaults: lambdaInitial, lambdaFactor 10, bound 1e5
    J, r, cost = _linearise(s, cal1, cal2, a1, a2, X0, huber_k)
    done = False
    while iters < max_iterations and not done:
        H, g = J.T @ J, J.T @ r
        while True:
            try:
                d = -np.linalg.solve(H + lam * np.eye(len(H)), g)
                trial = s.retract(d)
                new_cost = _errors(trial, cal1, cal2, a1, a2, X0, huber_k)[-1]
            except np.linalg.LinAlgError:
                new_cost = np.inf
            if np.isfinite(new_cost) and new_cost < cost:
                dec = cost - new_cost
                done = dec < abs_tol or dec < rel_tol * cost
                s, lam, iters = trial, lam / 10.0, iters + 1
                J, r, cost = _linearise(s, cal1, cal2, a1, a2, X0, huber_k)
                break
            lam *= 10.0
            if lam > 1e5:
                done = True
                break
    # filter_landmarks: both reprojection errors below the threshold, projections valid
    p0, ok0 = project(cal1, (s.X - s.c0) @ s.R0); p1, ok1 = project(cal2, (s.X - s.c1) @ s.R1)
    thr = np.inf if ba_thr is None else ba_thr
    v = ok0 & ok1 & (np.linalg.norm(p0 - a1, axis=1) < thr) & (np.linalg.norm(p1 - a2, axis=1) < thr)
    valid = np.zeros(n_all, bool); valid[np.flatnonzero(tri)[v]] = True
    if not v.any():
        return dict(R=R, t=t, valid=valid, triangulated=tri, iterations=iters, cost=cost, status=3)
    Rr = s.R1.T @ s.R0                               # wTi2.between(wTi1)
    tr = s.R1.T @ (s.c0 - s.c1)
    return dict(R=Rr, t=tr / np.linalg.norm(tr), valid=valid, triangulated=tri, iterations=iters, cost=cost, status=0)
