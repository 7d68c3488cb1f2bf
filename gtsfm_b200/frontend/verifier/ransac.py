"""B200 Ransac verifier: drop-in for the reference's gtsfm/frontend/verifier/ransac.py:25-111 on top of
OpencvVerifierBase.verify (opencv_verifier_base.py:45-109): same constructor kwargs, gates, outputs.

use_intrinsics_in_verification=True: essential matrix (ransac.py:52-82) -- normalisation, 5-point hypotheses,
Sampson/MSAC scoring, local optimisation, inlier mask and cv2.recoverPose-style pose recovery.
use_intrinsics_in_verification=False: fundamental matrix (ransac.py:84-111) -- 7-point hypotheses on float32 pixel
coordinates, cv2's FM_RANSAC error / inlier rule / model choice, E = K2^T F K1, the same pose recovery.
Both run entirely in libgtsfm_b200.so.  Hypothesis sample sets come from a counter-based
sampler (seed, pair key), exported by gtsfm_b200.ransac_samples so a CPU oracle can consume the identical set.
"""
from typing import List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200 import _native
from gtsfm_b200.common.types import Cal3Bundler, Keypoints, Rot3, Unit3, cal_to_tuple
from gtsfm_b200.frontend.verifier.verifier_base import VerifierBase

DEFAULT_NUM_HYPOTHESES = 1024      # the reference relies on cv2's default maxIters = 1000 (ransac.py:74-81)
DEFAULT_LO_ITERS = 20
# fundamental-matrix mode: the reference passes confidence 0.999999 and maxIters 1,000,000 to cv2.findFundamentalMat
# (ransac.py:22-23,103-110), i.e. the budget adapts to the inlier ratio; here: rounds of power-of-two budgets up to 2^20
RANSAC_SUCCESS_PROB = 0.999999
RANSAC_MAX_ITERS = 1 << 20


class Ransac(VerifierBase):
    def __init__(self, use_intrinsics_in_verification: bool, estimation_threshold_px: float,
                 num_hypotheses: int = DEFAULT_NUM_HYPOTHESES, seed: int = 0, lo_iters: int = DEFAULT_LO_ITERS, device: int = 0,
                 max_hypotheses: int = RANSAC_MAX_ITERS) -> None:
        super().__init__(use_intrinsics_in_verification, estimation_threshold_px)
        self._num_hypotheses = int(num_hypotheses)
        self._max_hypotheses = int(max_hypotheses)       # F mode only: cap of the adaptive budget (0 = fixed num_hypotheses)
        self._seed = int(seed)
        self._lo_iters = int(lo_iters)
        self._device = device

    def _params(self, pair_keys: Optional[np.ndarray] = None) -> "_native.VerifyParams":
        p = _native.VerifyParams()
        p.threshold_px = float(self._estimation_threshold_px)
        p.n_hypotheses = self._num_hypotheses
        p.lo_iters = self._lo_iters
        p.seed = self._seed
        p.pair_keys = None if pair_keys is None else pair_keys.ctypes.data
        p.samples = None
        p.estimate_fundamental = 0 if self._use_intrinsics_in_verification else 1
        p.adaptive_max_hypotheses = 0 if self._use_intrinsics_in_verification else self._max_hypotheses
        p.confidence = RANSAC_SUCCESS_PROB
        return p

    def verify(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, match_indices: np.ndarray,
               camera_intrinsics_i1: Cal3Bundler, camera_intrinsics_i2: Cal3Bundler
               ) -> Tuple[Optional[Rot3], Optional[Unit3], np.ndarray, float]:
        match_indices = np.asarray(match_indices)
        # opencv_verifier_base.py:69-78: fewer than 5 (E) / 8 (F) matches, then fewer than 6 in the E mode
        min_matches = 6 if self._use_intrinsics_in_verification else self._min_matches
        if match_indices.ndim != 2 or match_indices.shape[0] < min_matches:
            return self._failure_result
        ctx = _native.default_context(self._device)
        out = ctx.verify_pair(keypoints_i1.coordinates, keypoints_i2.coordinates, match_indices.astype(np.uint32),
                              cal_to_tuple(camera_intrinsics_i1), cal_to_tuple(camera_intrinsics_i2), self._params())
        return self._to_result(out["status"], out["R"], out["t"], out["mask"], match_indices)

    def _to_result(self, status, R, t, mask, match_indices):
        if status != _native.V_OK:
            return self._failure_result
        inlier_idxs = np.flatnonzero(mask == 1)
        v_corr_idxs = match_indices[inlier_idxs]                                   # opencv_verifier_base.py:98-100
        inlier_ratio = float(mask.mean()) if mask.size else 0.0                    # opencv_verifier_base.py:101
        return Rot3(np.array(R)), Unit3(np.array(t)), v_corr_idxs, inlier_ratio

    def verify_batch_raw(self, keypoints: Sequence[Keypoints], intrinsics: Sequence[Cal3Bundler], pairs: np.ndarray,
                         match_indices: Sequence[np.ndarray], device: Optional[int] = None, pair_ids: Optional[np.ndarray] = None):
        """All pairs in one device batch; returns (raw output arrays of gtsfm_verify_batch, row offsets, per-pair (M, 2) uint32
        match arrays, the context that holds the batch -- for gtsfm_two_view_report_batch).  pairs index `keypoints`; pair_ids
        (P, 2) are the caller's global image ids (sampler keys, so a pair draws the same hypotheses in any shard)."""
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        ctx = _native.default_context(self._device if device is None else device)
        ctx.clear()
        for i, (kp, cal) in enumerate(zip(keypoints, intrinsics)):
            ctx.upload_image(i, None, kp.coordinates, cal_to_tuple(cal))
        ms = [np.asarray(m).reshape(-1, 2).astype(np.uint32) if np.asarray(m).size else np.zeros((0, 2), np.uint32) for m in match_indices]
        off = np.zeros(len(ms) + 1, np.int64)
        np.cumsum([m.shape[0] for m in ms], out=off[1:])
        allm = np.concatenate(ms) if off[-1] else np.zeros((0, 2), np.uint32)
        ids = pairs if pair_ids is None else np.asarray(pair_ids).reshape(-1, 2)
        keys = (ids[:, 0].astype(np.uint64) << np.uint64(32)) | ids[:, 1].astype(np.uint64)
        out = ctx.verify_batch(pairs, self._params(keys), allm, off)
        return out, off, ms, ctx

    def verify_batch(self, keypoints: Sequence[Keypoints], intrinsics: Sequence[Cal3Bundler], pairs: np.ndarray,
                     match_indices: Sequence[np.ndarray]) -> List[Tuple[Optional[Rot3], Optional[Unit3], np.ndarray, float]]:
        """All pairs in one device batch (replaces one Dask task per pair, two_view_estimator.py:569-584)."""
        out, off, ms, _ = self.verify_batch_raw(keypoints, intrinsics, pairs, match_indices)
        src = [np.asarray(m).reshape(-1, 2) if np.asarray(m).size else np.zeros((0, 2), np.uint32) for m in match_indices]
        return [self._to_result(out["status"][p], out["R"][p], out["t"][p], out["mask"][off[p]:off[p + 1]], src[p])
                for p in range(len(ms))]
