"""ctypes front for the CPU oracle (oracle/liboracle.so).

ORACLE = test infrastructure.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package gtsfm_b200 never does.

Each function cites the reference code it restates in oracle_match.c / oracle_verify.c.
"""
import ctypes
import os
import subprocess
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

_dp = ctypes.POINTER(ctypes.c_double)
_fp = ctypes.POINTER(ctypes.c_float)
_ip = ctypes.POINTER(ctypes.c_int32)
_up = ctypes.POINTER(ctypes.c_uint32)
_bp = ctypes.POINTER(ctypes.c_uint8)


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (gcc, a second or two)."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle_match.c", "oracle_verify.c", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_cv_l2sqr.restype = ctypes.c_float
        _lib.oracle_sampson_sq.restype = ctypes.c_double
        _lib.oracle_lo_refine.restype = ctypes.c_double
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def l2sqr(a: np.ndarray, b: np.ndarray) -> np.float32:
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return np.float32(lib().oracle_cv_l2sqr(a.ctypes.data_as(_fp), b.ctypes.data_as(_fp), int(a.size)))


def knn2(q: np.ndarray, t: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """cv2.BFMatcher(NORM_L2).knnMatch(q, t, k=2) as (best idx, best dist, second dist)."""
    q = np.ascontiguousarray(q, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    nq, dim = q.shape
    idx = np.empty(nq, np.int32)
    d1 = np.empty(nq, np.float32)
    d2 = np.empty(nq, np.float32)
    lib().oracle_knn2(q.ctypes.data_as(_fp), nq, t.ctypes.data_as(_fp), int(t.shape[0]), dim,
                      idx.ctypes.data_as(_ip), d1.ctypes.data_as(_fp), d2.ctypes.data_as(_fp))
    return idx, d1, d2


def twoway_match(desc1: np.ndarray, desc2: np.ndarray, ratio: Optional[float] = None) -> np.ndarray:
    """TwoWayMatcher(EUCLIDEAN, ratio).match (twoway_matcher.py:42-144). Empty -> np.array([])."""
    if desc1.size == 0 or desc2.size == 0:
        return np.array([])
    a = np.ascontiguousarray(desc1, np.float32)
    b = np.ascontiguousarray(desc2, np.float32)
    n1, dim = a.shape
    n2 = b.shape[0]
    out = np.empty((min(n1, n2), 2), np.uint32)
    k = lib().oracle_twoway_match(a.ctypes.data_as(_fp), n1, b.ctypes.data_as(_fp), n2, dim,
                                  int(ratio is not None), ctypes.c_double(0.0 if ratio is None else ratio),
                                  out.ctypes.data_as(_up))
    if k < 0:
        raise ValueError("not enough values to unpack (expected 2, got 1)")
    if k == 0:
        return np.array([])
    return out[:k].copy()


def twoway_match_hamming(desc1: np.ndarray, desc2: np.ndarray, ratio: Optional[float] = None) -> np.ndarray:
    """TwoWayMatcher(HAMMING, ratio).match on uint8 descriptors (twoway_matcher.py:42-144 with cv2.NORM_HAMMING)."""
    if desc1.size == 0 or desc2.size == 0:
        return np.array([])
    a = np.ascontiguousarray(desc1, np.uint8)
    b = np.ascontiguousarray(desc2, np.uint8)
    out = np.empty((min(a.shape[0], b.shape[0]), 2), np.uint32)
    k = lib().oracle_twoway_match_hamming(a.ctypes.data_as(_bp), a.shape[0], b.ctypes.data_as(_bp), b.shape[0], a.shape[1],
                                          int(ratio is not None), ctypes.c_double(0.0 if ratio is None else ratio),
                                          out.ctypes.data_as(_up))
    if k < 0:
        raise ValueError("not enough values to unpack (expected 2, got 1)")
    return out[:k].copy() if k else np.array([])


def sampson_sq(F: np.ndarray, c1: np.ndarray, c2: np.ndarray) -> np.ndarray:
    """gtsfm/utils/verification.py:170-214."""
    F = np.ascontiguousarray(F, np.float64)
    c1 = np.asarray(c1, np.float64)
    c2 = np.asarray(c2, np.float64)
    n = c1.shape[0]
    x1, y1, x2, y2 = (np.ascontiguousarray(v) for v in (c1[:, 0], c1[:, 1], c2[:, 0], c2[:, 1]))
    out = np.empty(n)
    lib().oracle_sampson_sq_batch(_d(F), _d(x1), _d(y1), _d(x2), _d(y2), n, _d(out))
    return out


def fivept(p1: np.ndarray, p2: np.ndarray) -> np.ndarray:
    """Nister 5-point on five normalised correspondences -> (k,3,3) unit-norm essential matrices."""
    p1 = np.asarray(p1, np.float64)
    p2 = np.asarray(p2, np.float64)
    x1, y1, x2, y2 = (np.ascontiguousarray(v) for v in (p1[:, 0], p1[:, 1], p2[:, 0], p2[:, 1]))
    out = np.zeros(90)
    n = lib().oracle_fivept(_d(x1), _d(y1), _d(x2), _d(y2), _d(out))
    return out[: 9 * n].reshape(n, 3, 3).copy()


def recover_pose(E: np.ndarray, n1: np.ndarray, n2: np.ndarray):
    """cv2.recoverPose(E, n1, n2) on normalised coordinates (verification.py:81) -> R, t, good[4]."""
    E = np.ascontiguousarray(E, np.float64)
    n1 = np.asarray(n1, np.float64)
    n2 = np.asarray(n2, np.float64)
    x1, y1, x2, y2 = (np.ascontiguousarray(v) for v in (n1[:, 0], n1[:, 1], n2[:, 0], n2[:, 1]))
    R = np.empty(9)
    t = np.empty(3)
    good = np.zeros(4, np.int32)
    lib().oracle_recover_pose(_d(E), _d(x1), _d(y1), _d(x2), _d(y2), int(n1.shape[0]), _d(R), _d(t),
                              good.ctypes.data_as(_ip))
    return R.reshape(3, 3), t, good


def calibrate(coords: np.ndarray, cal) -> np.ndarray:
    """normalize_coordinates (features.py:40-50); cal = (f, k1, k2, u0, v0)."""
    coords = np.asarray(coords, np.float64)
    out = np.empty((coords.shape[0], 2))
    x = ctypes.c_double()
    y = ctypes.c_double()
    f = lib().oracle_calibrate
    f.argtypes = [ctypes.c_double] * 7 + [_dp, _dp]
    for i in range(coords.shape[0]):
        f(coords[i, 0], coords[i, 1], cal[0], cal[1], cal[2], cal[3], cal[4], ctypes.byref(x), ctypes.byref(y))
        out[i] = (x.value, y.value)
    return out


def verify_pair(kp1: np.ndarray, kp2: np.ndarray, match_indices: np.ndarray, cal1, cal2, thr_px: float,
                samples: np.ndarray, lo_iters: int = 20) -> dict:
    """OpencvVerifierBase.verify + Ransac.estimate_E with an explicit (H,5) sample array.

    Returns a dict: status, E_min, E, mask (M,) uint8, R, t, ninl, best_id, ncand, cost_min, cost.
    """
    kp1 = np.ascontiguousarray(kp1, np.float64).reshape(-1, 2)
    kp2 = np.ascontiguousarray(kp2, np.float64).reshape(-1, 2)
    mi = np.asarray(match_indices).reshape(-1, 2)
    M = mi.shape[0]
    m1 = np.ascontiguousarray(mi[:, 0], np.int32)
    m2 = np.ascontiguousarray(mi[:, 1], np.int32)
    c1 = np.asarray(cal1, np.float64)
    c2 = np.asarray(cal2, np.float64)
    samples = np.ascontiguousarray(samples, np.int32).reshape(-1, 5)
    E_min = np.zeros(9)
    E = np.zeros(9)
    mask = np.zeros(max(M, 1), np.uint8)
    R = np.zeros(9)
    t = np.zeros(3)
    info = np.zeros(3, np.int32)
    cost = np.zeros(2)
    st = lib().oracle_verify_pair(_d(kp1), _d(kp2), m1.ctypes.data_as(_ip), m2.ctypes.data_as(_ip), M,
                                  _d(c1), _d(c2), ctypes.c_double(thr_px), samples.ctypes.data_as(_ip),
                                  int(samples.shape[0]), int(lo_iters), _d(E_min), _d(E),
                                  mask.ctypes.data_as(_bp), _d(R), _d(t), info.ctypes.data_as(_ip), _d(cost))
    return dict(status=int(st), E_min=E_min.reshape(3, 3), E=E.reshape(3, 3), mask=mask[:M], R=R.reshape(3, 3),
                t=t, ninl=int(info[0]), best_id=int(info[1]), ncand=int(info[2]), cost_min=float(cost[0]),
                cost=float(cost[1]))


# ---------------------------------------------------------------- fundamental-matrix mode (ransac.py:84-111)
def sevenpt(p1: np.ndarray, p2: np.ndarray) -> np.ndarray:
    """7-point solver on seven pixel correspondences (cv2.findFundamentalMat(FM_7POINT)) -> (k,3,3) unit-norm F."""
    p1 = np.asarray(p1, np.float64)
    p2 = np.asarray(p2, np.float64)
    x1, y1, x2, y2 = (np.ascontiguousarray(v) for v in (p1[:, 0], p1[:, 1], p2[:, 0], p2[:, 1]))
    out = np.zeros(27)
    n = lib().oracle_sevenpt(_d(x1), _d(y1), _d(x2), _d(y2), _d(out))
    return out[: 9 * n].reshape(n, 3, 3).copy()


def fm_error(F: np.ndarray, c1: np.ndarray, c2: np.ndarray) -> np.ndarray:
    """cv2's FM_RANSAC error: max of the two squared point-to-epipolar-line distances, float32 (double arithmetic on the
    float32-rounded pixel coordinates)."""
    F = np.ascontiguousarray(F, np.float64)
    c1 = np.asarray(c1, np.float32)
    c2 = np.asarray(c2, np.float32)
    x1, y1, x2, y2 = (np.ascontiguousarray(v) for v in (c1[:, 0], c1[:, 1], c2[:, 0], c2[:, 1]))
    out = np.empty(x1.shape[0], np.float32)
    lib().oracle_fm_error_batch(_d(F), x1.ctypes.data_as(_fp), y1.ctypes.data_as(_fp), x2.ctypes.data_as(_fp),
                                y2.ctypes.data_as(_fp), int(x1.shape[0]), out.ctypes.data_as(_fp))
    return out


def verify_pair_F(kp1: np.ndarray, kp2: np.ndarray, match_indices: np.ndarray, cal1, cal2, thr_px: float,
                  samples: np.ndarray) -> dict:
    """OpencvVerifierBase.verify with use_intrinsics_in_verification=False and an explicit (H,7) sample array.

    Returns a dict: status, F, E, mask (M,) uint8, R, t, ninl, best_id, ncand.
    """
    kp1 = np.ascontiguousarray(kp1, np.float64).reshape(-1, 2)
    kp2 = np.ascontiguousarray(kp2, np.float64).reshape(-1, 2)
    mi = np.asarray(match_indices).reshape(-1, 2)
    M = mi.shape[0]
    m1 = np.ascontiguousarray(mi[:, 0], np.int32)
    m2 = np.ascontiguousarray(mi[:, 1], np.int32)
    c1 = np.asarray(cal1, np.float64)
    c2 = np.asarray(cal2, np.float64)
    samples = np.ascontiguousarray(samples, np.int32).reshape(-1, 7)
    F = np.zeros(9)
    E = np.zeros(9)
    mask = np.zeros(max(M, 1), np.uint8)
    R = np.zeros(9)
    t = np.zeros(3)
    info = np.zeros(3, np.int32)
    st = lib().oracle_verify_pair_F(_d(kp1), _d(kp2), m1.ctypes.data_as(_ip), m2.ctypes.data_as(_ip), M, _d(c1), _d(c2),
                                    ctypes.c_double(thr_px), samples.ctypes.data_as(_ip), int(samples.shape[0]),
                                    _d(F), _d(E), mask.ctypes.data_as(_bp), _d(R), _d(t), info.ctypes.data_as(_ip))
    return dict(status=int(st), F=F.reshape(3, 3), E=E.reshape(3, 3), mask=mask[:M], R=R.reshape(3, 3), t=t,
                ninl=int(info[0]), best_id=int(info[1]), ncand=int(info[2]))


# ---------------------------------------------------------------- post-verification report
def essential_to_fundamental(R: np.ndarray, t: np.ndarray, cal1, cal2) -> np.ndarray:
    """essential_to_fundamental_matrix (utils/verification.py:113-126) with i2Ei1 = EssentialMatrix(R, Unit3(t)) = [t/|t|]x R."""
    t = np.asarray(t, np.float64) / np.linalg.norm(t)
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    K1 = np.array([[cal1[0], 0, cal1[3]], [0, cal1[0], cal1[4]], [0, 0, 1.0]])
    K2 = np.array([[cal2[0], 0, cal2[3]], [0, cal2[0], cal2[4]], [0, 0, 1.0]])
    return np.linalg.inv(K2.T) @ (tx @ np.asarray(R, np.float64)) @ np.linalg.inv(K1)


def sampson_sq_numpy(c1: np.ndarray, c2: np.ndarray, F: np.ndarray) -> np.ndarray:
    """compute_epipolar_distances_sq_sampson (utils/verification.py:170-214), line for line in numpy."""
    h1 = np.c_[c1, np.ones(len(c1))]
    h2 = np.c_[c2, np.ones(len(c2))]
    lines_i2 = h1 @ F.T                      # F x1
    lines_i1 = h2 @ F                        # F^T x2
    num = np.square(np.sum(h1 * lines_i1, axis=1))
    den = np.sum(np.square(lines_i1[:, :2]), axis=1) + np.sum(np.square(lines_i2[:, :2]), axis=1)
    return num / den


def two_view_report(kp1, kp2, v_corr_idxs, n_putative, R_est, t_est, cal1, cal2, gt_R, gt_t, dist_threshold,
                    min_num_inliers, min_inlier_ratio) -> dict:
    """epipolar_inlier_correspondences (utils/metrics.py:100-128) + compute_relative_pose_metrics
    (utils/geometry_comparisons.py:266-309) + generate_two_view_report aggregates (two_view_estimator.py:354-393) +
    InlierSupportProcessor.run_inlier_support (frontend/inlier_support_processor.py:39-95) for one pair."""
    from scipy.spatial.transform import Rotation
    v = np.asarray(v_corr_idxs).reshape(-1, 2)
    n_est = len(v) if R_est is not None else 0
    ratio = n_est / n_putative if (R_est is not None and n_putative) else 0.0
    keep = not (ratio < min_inlier_ratio) and not (n_est > 0 and n_est < min_num_inliers) and R_est is not None
    out = dict(keep=keep, dist_sq=None, gt_mask=None, num_inliers_gt=0, inlier_avg=float("nan"), outlier_avg=float("nan"),
               R_error_deg=None, U_error_deg=None)
    if gt_R is None or R_est is None:
        return out
    out["R_error_deg"] = float(np.rad2deg(np.linalg.norm(Rotation.from_matrix(np.asarray(R_est).T @ np.asarray(gt_R)).as_rotvec())))
    u = np.asarray(gt_t, np.float64) / np.linalg.norm(gt_t)
    out["U_error_deg"] = float(np.rad2deg(np.arccos(np.clip(np.dot(np.asarray(t_est), u), -1, 1))))
    if n_est == 0:
        return out
    F = essential_to_fundamental(gt_R, gt_t, cal1, cal2)
    d2 = sampson_sq_numpy(np.asarray(kp1, np.float64)[v[:, 0]], np.asarray(kp2, np.float64)[v[:, 1]], F)
    gm = d2 < dist_threshold ** 2
    out.update(dist_sq=d2, gt_mask=gm, num_inliers_gt=int(np.count_nonzero(gm)))
    with np.errstate(all="ignore"):
        out["inlier_avg"] = float(np.mean(d2[gm])) if gm.any() else float("nan")
        rest = d2[~gm]
        out["outlier_avg"] = float(np.nanmean(rest)) if np.any(~np.isnan(rest)) else float("nan")
    return out


# ---------------------------------------------------------------------------------------------------- retriever (8(f) rank 4)
def retriever_similarity(global_descriptors: np.ndarray, blocksize: int = 50) -> np.ndarray:
    """gtsfm/retriever/netvlad_retriever.py:77-153: blocked G G^T, only blocks with block_j >= block_i, zeros elsewhere.
    float64 accumulation here (the reference's einsum order is BLAS-defined), rounded to float32."""
    g = np.asarray(global_descriptors, np.float64)
    n = g.shape[0]
    sim = g @ g.T
    bi = np.arange(n) // blocksize
    sim[bi[None, :] < bi[:, None]] = 0.0
    return sim.astype(np.float32)


def pairs_from_score_matrix(scores: np.ndarray, num_select: int, min_score: Optional[float] = None):
    """gtsfm/retriever/netvlad_retriever.py:166-168,196-228 in numpy: upper triangle without the diagonal, scores below
    min_score (compared in float32) dropped, the num_select best per row (ties: lower index), row-major pair list."""
    s = np.array(scores, np.float32)
    n = s.shape[0]
    invalid = ~np.triu(np.ones((n, n), bool))
    np.fill_diagonal(invalid, True)
    if min_score is not None:
        invalid |= s < np.float32(min_score)
    s[invalid] = -np.inf
    k = min(num_select, n)
    pairs = []
    for i in range(n):
        order = np.lexsort((np.arange(n), -s[i]))[:k]
        pairs += [(i, int(j)) for j in order if np.isfinite(s[i, j])]
    return pairs
