"""ORACLE / CPU BASELINE (test infrastructure only -- never imported by the product package gtsfm_b200).

The reference's CPU path for one image pair, restated call for call on the library the reference itself uses
(opencv-python, un-vendored; the container and the GPU box ship cv2 4.13):

  TwoWayMatcher.match          gtsfm/frontend/matcher/twoway_matcher.py:42-144   cv2.BFMatcher(NORM_L2).knnMatch / .match,
                                                                                 ratio test in Python doubles, stable sort,
                                                                                 mutual check
  OpencvVerifierBase.verify    gtsfm/frontend/verifier/opencv_verifier_base.py:45-109
  Ransac.estimate_E            gtsfm/frontend/verifier/ransac.py:52-82           cv2.findEssentialMat(USAC_ACCURATE, prob 0.999999)
  recover_relative_pose...     gtsfm/utils/verification.py:52-94                 cv2.recoverPose

It exists because /root/reference (and gtsam, dask) are absent on the GPU box: bench.py times THIS as the reference
arm, one task per pair on a process pool with cv2 pinned to one thread per worker -- the reference's own execution
model (one Dask task per pair, gtsfm_runner_base.py:288-296).  normalize_coordinates is vectorised here (the reference
loops in Python over gtsam calls, features.py:40-50), so this baseline is, if anything, faster than the reference.
Its outputs are checked against the golden vectors recorded from the unmodified reference (tests/test_ref_cv2.py).
"""
import os
from concurrent.futures import ProcessPoolExecutor
from typing import Optional

import numpy as np


def _oneway(d_q: np.ndarray, d_t: np.ndarray, ratio: Optional[float]):
    import cv2
    bf = cv2.BFMatcher(cv2.NORM_L2, crossCheck=False)
    if ratio is not None:
        matches = [m1 for m1, m2 in bf.knnMatch(d_q, d_t, k=2) if m1.distance <= ratio * m2.distance]
    else:
        matches = bf.match(d_q, d_t)
    matches = sorted(matches, key=lambda r: r.distance)
    return {m.queryIdx: m.trainIdx for m in matches}


def match_pair(d1: np.ndarray, d2: np.ndarray, ratio: Optional[float]) -> np.ndarray:
    if d1.size == 0 or d2.size == 0:
        return np.array([])
    v1 = np.flatnonzero(~np.isnan(d1).any(axis=1))
    v2 = np.flatnonzero(~np.isnan(d2).any(axis=1))
    if v1.size == 0 or v2.size == 0:
        return np.array([])
    a, b = np.ascontiguousarray(d1[v1], np.float32), np.ascontiguousarray(d2[v2], np.float32)
    m12 = _oneway(a, b, ratio)
    m21 = _oneway(b, a, ratio)
    out = np.array([(i1, i2) for i1, i2 in m12.items() if m21.get(i2, -1) == i1], dtype=np.uint32)
    if out.size == 0:
        return np.array([])
    return np.stack([v1[out[:, 0]], v2[out[:, 1]]], axis=1).astype(np.uint32)


def _normalize(kp: np.ndarray, cal) -> np.ndarray:
    f, k1, k2, u0, v0 = cal
    assert k1 == 0 and k2 == 0
    return (np.asarray(kp, np.float64) - [u0, v0]) / f


def verify_pair_F(kp1: np.ndarray, kp2: np.ndarray, match_indices: np.ndarray, cal1, cal2, thr_px: float):
    """Ransac(use_intrinsics_in_verification=False): ransac.py:84-111 + opencv_verifier_base.py:90-108."""
    import cv2
    fail = (None, None, np.array([], dtype=np.uint64), 0.0)
    if match_indices.ndim != 2 or match_indices.shape[0] < 8:
        return fail
    F, mask = cv2.findFundamentalMat(np.asarray(kp1, np.float64)[match_indices[:, 0]], np.asarray(kp2, np.float64)[match_indices[:, 1]],
                                     method=cv2.FM_RANSAC, ransacReprojThreshold=thr_px, confidence=0.999999, maxIters=1000000)
    if F is None or mask is None:
        return fail
    K1 = np.array([[cal1[0], 0, cal1[3]], [0, cal1[0], cal1[4]], [0, 0, 1.0]])
    K2 = np.array([[cal2[0], 0, cal2[3]], [0, cal2[0], cal2[4]], [0, 0, 1.0]])
    E = K2.T @ F @ K1                                                          # utils/verification.py:110
    inl = np.flatnonzero(mask.ravel() == 1)
    v = match_indices[inl]
    if len(inl) == 0:
        return fail
    uv1, uv2 = _normalize(kp1, cal1), _normalize(kp2, cal2)
    _, R, t, _ = cv2.recoverPose(E, uv1[v[:, 0]], uv2[v[:, 1]])
    return R, t.ravel(), v, float(mask.mean())


def verify_pair(kp1: np.ndarray, kp2: np.ndarray, match_indices: np.ndarray, cal1, cal2, thr_px: float):
    """Returns (R, t, v_corr_idxs, inlier_ratio) or (None, None, empty, 0.0)."""
    import cv2
    fail = (None, None, np.array([], dtype=np.uint64), 0.0)
    if match_indices.ndim != 2 or match_indices.shape[0] < 6:
        return fail
    uv1, uv2 = _normalize(kp1, cal1), _normalize(kp2, cal2)
    fx = max(cal1[0], cal2[0])
    E, mask = cv2.findEssentialMat(uv1[match_indices[:, 0]], uv2[match_indices[:, 1]], np.eye(3), method=cv2.USAC_ACCURATE,
                                   threshold=thr_px / fx, prob=0.999999)
    if E is None or mask is None:
        return fail
    if E.shape[0] > 3:
        E = E[:3]
    inl = np.flatnonzero(mask.ravel() == 1)
    v = match_indices[inl]
    if len(inl) == 0:
        return fail
    _, R, t, _ = cv2.recoverPose(E, uv1[v[:, 0]], uv2[v[:, 1]])
    return R, t.ravel(), v, float(mask.mean())


_SCENE = None
_KEEP_POSES = False


def _init_worker(scene, keep_poses=False):
    global _SCENE, _KEEP_POSES
    import cv2
    cv2.setNumThreads(1)
    _SCENE = scene
    _KEEP_POSES = keep_poses


def _task(p: int):
    s = _SCENE
    i, j = s.pairs[p]
    m = match_pair(s.descriptors[i], s.descriptors[j], s.ratio)
    if m.size == 0:
        return 0, 0
    fn = verify_pair_F if getattr(s, "fundamental", False) else verify_pair
    R, t, v, ratio = fn(s.keypoints[i], s.keypoints[j], m, s.cal, s.cal, s.threshold_px)
    if _KEEP_POSES:                    # bench.py's agreement report: cv2's pose and inlier set of the sampled pair
        return len(m), len(v), R, t, np.asarray(v, np.uint32).reshape(-1, 2)
    return len(m), len(v)


class PairPool:
    """One task per pair on n_workers single-threaded processes (the reference's Dask LocalCluster granularity)."""

    def __init__(self, scene, n_workers: Optional[int] = None, keep_poses: bool = False):
        self.n_workers = n_workers or len(os.sched_getaffinity(0))
        import multiprocessing as mp
        self._ex = ProcessPoolExecutor(self.n_workers, mp_context=mp.get_context("fork"), initializer=_init_worker,
                                       initargs=(scene, keep_poses))

    def run(self, pair_ids):
        return list(self._ex.map(_task, [int(p) for p in pair_ids]))

    def close(self):
        self._ex.shutdown()
