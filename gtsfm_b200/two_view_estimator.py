"""TwoViewEstimator and run_two_view_estimator_as_futures: drop-ins for the reference's gtsfm/two_view_estimator.py:54-349
and :531-587 -- same constructor, same run_2view(...) / run_two_view_estimator_as_futures(client, ...) signatures, same
TWO_VIEW_OUTPUT 6-tuples -- with the per-pair Dask submission loop (two_view_estimator.py:567-587) replaced by ONE task per
GPU shard of the pair list.

A shard task uploads the keypoints of the images its pairs name, verifies all pairs in one device batch
(gtsfm_verify_batch), triangulates and bundle-adjusts every pair with enough verified correspondences in one launch when
bundle_adjust_2view is True (gtsfm_two_view_ba_batch), computes the ground-truth metrics, the report aggregates and the
InlierSupportProcessor decision of every pair in one more launch per report (gtsfm_two_view_report_batch, pre- and post-BA)
and assembles the reference's structures on the host.  run_2view is a batch of one through the same code.

Not offered (loud errors, no CPU fallback): relative pose priors together with bundle adjustment (two_view_estimator.py:
165-167,189: the prior as the initial pose and a BetweenFactor), triangulation modes other than NO_RANSAC, and mesh-based
correspondence metrics (`gt_scene_mesh`, utils/metrics.py:75-83: ray casting against a trimesh).
"""
import dataclasses
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200 import _native
from gtsfm_b200.common.two_view_estimation_report import TwoViewEstimationReport
from gtsfm_b200.common.types import Keypoints, Rot3, Unit3, cal_to_tuple
from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
from gtsfm_b200.frontend.verifier.verifier_base import VerifierBase
from gtsfm_b200.two_view_batch import TwoViewBatchResult, num_gpu_devices, num_gpu_shards, shard_pairs, two_view_reports

PRE_BA_REPORT_TAG = "PRE_BA_2VIEW_REPORT"
POST_BA_REPORT_TAG = "POST_BA_2VIEW_REPORT"
POST_ISP_REPORT_TAG = "POST_INLIER_SUPPORT_PROCESSOR_2VIEW_REPORT"
VIEWGRAPH_REPORT_TAG = "VIEWGRAPH_2VIEW_REPORT"

TWO_VIEW_OUTPUT = Tuple[Optional[Rot3], Optional[Unit3], np.ndarray, TwoViewEstimationReport, TwoViewEstimationReport,
                        TwoViewEstimationReport]


def _gt_relative_pose(gt_camera_i1, gt_camera_i2) -> Tuple[np.ndarray, np.ndarray]:
    """i2Ti1 = gt_camera_i2.pose().between(gt_camera_i1.pose()) (utils/metrics.py:86) as (R (3,3), t (3,))."""
    w1, w2 = gt_camera_i1.pose(), gt_camera_i2.pose()
    R1, t1 = np.asarray(w1.rotation().matrix(), np.float64), np.asarray(w1.translation(), np.float64).reshape(3)
    R2, t2 = np.asarray(w2.rotation().matrix(), np.float64), np.asarray(w2.translation(), np.float64).reshape(3)
    return R2.T @ R1, R2.T @ (t1 - t2)


class TwoViewEstimator:
    """Two-view relative pose estimation on image pairs, batched on the GPU."""

    def __init__(self, verifier: VerifierBase, inlier_support_processor: InlierSupportProcessor, bundle_adjust_2view: bool,
                 eval_threshold_px: float, triangulation_options: Any = None, bundle_adjust_2view_maxiters: int = 100,
                 ba_reproj_error_thresholds: Sequence[Optional[float]] = (0.5,)) -> None:
        self._verifier = verifier
        self.processor = inlier_support_processor
        self._bundle_adjust_2view = bundle_adjust_2view
        self._corr_metric_dist_threshold = eval_threshold_px
        self._triangulation_options = triangulation_options
        self._ba_reproj_error_thresholds = list(ba_reproj_error_thresholds)
        self._bundle_adjust_2view_maxiters = bundle_adjust_2view_maxiters

    def __repr__(self) -> str:
        return f"""
        TwoViewEstimator:
            Verifier: {self._verifier}
            Bundle adjust 2-view: {self._bundle_adjust_2view}
            Correspondence metric dist. threshold: {self._corr_metric_dist_threshold}
            BA reproj. error thresholds: {self._ba_reproj_error_thresholds}
            BA 2-view max. iters: {self._bundle_adjust_2view_maxiters}
        """

    def get_corr_metric_dist_threshold(self) -> float:
        return self._corr_metric_dist_threshold

    # ---- the reference's per-pair entry point (two_view_estimator.py:272-349)
    def run_2view(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, putative_corr_idxs: np.ndarray, camera_intrinsics_i1,
                  camera_intrinsics_i2, i2Ti1_prior=None, gt_camera_i1=None, gt_camera_i2=None, gt_scene_mesh: Optional[Any] = None
                  ) -> TWO_VIEW_OUTPUT:
        out = self.run_2view_batch({0: keypoints_i1, 1: keypoints_i2}, {(0, 1): putative_corr_idxs},
                                   {0: camera_intrinsics_i1, 1: camera_intrinsics_i2}, {(0, 1): i2Ti1_prior},
                                   {0: gt_camera_i1, 1: gt_camera_i2}, gt_scene_mesh)
        return out[(0, 1)]

    # ---- one GPU shard
    def run_2view_batch(self, keypoints, putative_corr_idxs_dict: Dict[Tuple[int, int], np.ndarray], camera_intrinsics,
                        relative_pose_priors=None, gt_cameras=None, gt_scene_mesh: Optional[Any] = None,
                        device: Optional[int] = None) -> Dict[Tuple[int, int], TWO_VIEW_OUTPUT]:
        """run_2view for every pair of the dictionary, in one device batch.  keypoints / camera_intrinsics / gt_cameras are
        indexable by image id (lists, or dicts holding only the images this shard names)."""
        if self._bundle_adjust_2view:
            mode = getattr(self._triangulation_options, "mode", None)
            if self._triangulation_options is None or getattr(mode, "value", mode) != "NO_RANSAC":
                raise NotImplementedError("two-view bundle adjustment on the B200 path triangulates with TriangulationSamplingMode.NO_RANSAC "
                                          "(configs/sift_front_end.yaml:50-55); pass such TriangulationOptions")
            if relative_pose_priors is not None and any(relative_pose_priors.get(k) is not None for k in putative_corr_idxs_dict):
                raise NotImplementedError("relative pose priors in two-view bundle adjustment (two_view_estimator.py:165-167,189) are not "
                                          "part of the B200 two-view path")
        if gt_scene_mesh is not None:
            raise NotImplementedError("mesh-based correspondence metrics (utils/metrics.py:75-83) are not part of the B200 two-view path")
        if not hasattr(self._verifier, "verify_batch_raw"):
            raise NotImplementedError(f"{type(self._verifier).__name__} has no batched device path (verify_batch_raw)")
        keys = list(putative_corr_idxs_dict.keys())
        if not keys:
            return {}
        pairs = np.asarray(keys, np.int64).reshape(-1, 2)
        used = [int(g) for g in np.unique(pairs)]
        remap = {g: k for k, g in enumerate(used)}
        local = np.array([[remap[int(i)], remap[int(j)]] for i, j in pairs], np.int32)
        raw, off, ms, ctx = self._verifier.verify_batch_raw([keypoints[g] for g in used], [camera_intrinsics[g] for g in used], local,
                                                            [putative_corr_idxs_dict[k] for k in keys], device=device, pair_ids=pairs)
        P = len(keys)
        # ---- ground truth (two_view_estimator.py:131-153): relative pose and the GT cameras' calibrations, NaN rows where absent
        gt_R = gt_t = gt_cal = None
        if gt_cameras is not None:
            have = [gt_cameras[int(i)] is not None and gt_cameras[int(j)] is not None for i, j in pairs]
            if any(have):
                gt_R = np.full((P, 9), np.nan); gt_t = np.full((P, 3), np.nan); gt_cal = np.ones((P, 10))
                for p, (i, j) in enumerate(pairs):
                    if not have[p]:
                        continue
                    R, t = _gt_relative_pose(gt_cameras[int(i)], gt_cameras[int(j)])
                    gt_R[p], gt_t[p] = R.reshape(9), t
                    gt_cal[p, :5] = cal_to_tuple(gt_cameras[int(i)].calibration())
                    gt_cal[p, 5:] = cal_to_tuple(gt_cameras[int(j)].calibration())
        allm = np.concatenate(ms) if off[-1] else np.zeros((0, 2), np.uint32)
        result = TwoViewBatchResult(pairs.astype(np.int32), off, allm, raw["mask"], raw["R"], raw["t"], raw["n_inliers"], raw["status"])
        reports, keep, _ = two_view_reports(result, ctx, self._corr_metric_dist_threshold, self.processor._min_num_inliers_est_model,
                                            self.processor._min_inlier_ratio_est_model, gt_R, gt_t, gt_cal)
        total = int(off[-1])
        post_reports, post_keep, ba = reports, keep, None
        if self._bundle_adjust_2view:
            # two_view_estimator.py:316-333: BA where enough correspondences were verified, then the post-BA report on its results
            thr = self._ba_reproj_error_thresholds[-1] if self._ba_reproj_error_thresholds else None     # every stage restarts from
            ba = ctx.two_view_ba_batch(P, total, self.processor._min_num_inliers_est_model,              # the initial data
                                       self._bundle_adjust_2view_maxiters,                                # (bundle_adjustment.py:381-390)
                                       float(self._triangulation_options.reproj_error_threshold),
                                       float("inf") if thr is None else float(thr), 1.345,
                                       float(self._triangulation_options.min_triangulation_angle))
            ba_result = TwoViewBatchResult(result.pairs, off, allm, ba["valid_mask"], ba["R"].reshape(P, 9), ba["t"], ba["num_valid"],
                                           raw["status"])
            pre_ratio = np.array([r.inlier_ratio_est_model for r in reports])
            post_reports, post_keep, _ = two_view_reports(ba_result, ctx, self._corr_metric_dist_threshold,
                                                          self.processor._min_num_inliers_est_model,
                                                          self.processor._min_inlier_ratio_est_model, gt_R, gt_t, gt_cal, after_ba=True,
                                                          inlier_ratios=pre_ratio)
        out: Dict[Tuple[int, int], TWO_VIEW_OUTPUT] = {}
        for p, key in enumerate(keys):
            pre_ba_report = reports[p]
            ok = raw["status"][p] == _native.V_OK
            # v_corr_idxs keep the caller's dtype and row order (opencv_verifier_base.py:98-100); failure -> empty uint64
            src = np.asarray(putative_corr_idxs_dict[key]).reshape(-1, 2)
            v = src[raw["mask"][off[p]:off[p + 1]] == 1] if ok else np.array([], dtype=np.uint64)
            pre_ba_report.v_corr_idxs = v
            if ba is not None and ba["status"][p] != 1:
                post_ba_report = post_reports[p]
                post_v = src[ba["valid_mask"][off[p]:off[p + 1]] == 1]
                if len(post_v) == 0:
                    post_v = np.zeros((0, 2), np.int32)                                # two_view_estimator.py:180 / an empty selection
                post_ba_report.v_corr_idxs = post_v
                post_R, post_t = ba["R"][p], ba["t"][p]
            else:
                post_ba_report = dataclasses.replace(pre_ba_report)                  # two_view_estimator.py:338
                post_v, post_R, post_t = v, raw["R"][p], raw["t"][p]
            if post_keep[p]:
                post = (Rot3(np.array(post_R, np.float64).reshape(3, 3)), Unit3(np.array(post_t, np.float64)), post_v,
                        dataclasses.replace(post_ba_report))
            else:                                                                    # inlier_support_processor.py:63-68
                post = (None, None, np.array([], dtype=np.uint64), TwoViewEstimationReport(v_corr_idxs=post_v, num_inliers_est_model=0))
            out[(int(key[0]), int(key[1]))] = (post[0], post[1], post[2], pre_ba_report, post_ba_report, post[3])
        return out


def _apply_two_view_estimator_shard(two_view_estimator, keypoints: Dict[int, Keypoints], putative: Dict[Tuple[int, int], np.ndarray],
                                    intrinsics: Dict[int, Any], priors: Dict[Tuple[int, int], Any], gt_cameras: Dict[int, Any],
                                    gt_scene_mesh: Optional[Any], device: int) -> Dict[Tuple[int, int], TWO_VIEW_OUTPUT]:
    if hasattr(two_view_estimator, "run_2view_batch"):
        return two_view_estimator.run_2view_batch(keypoints, putative, intrinsics, priors, gt_cameras, gt_scene_mesh, device=device)
    # a foreign estimator (only run_2view): the reference's per-pair loop (two_view_estimator.py:567-584) inside the shard
    return {(int(i), int(j)): two_view_estimator.run_2view(
        keypoints[i], keypoints[j], corr, intrinsics[i], intrinsics[j], priors.get((i, j)) if priors else None,
        gt_cameras[i] if gt_cameras is not None else None, gt_cameras[j] if gt_cameras is not None else None, gt_scene_mesh)
        for (i, j), corr in putative.items()}


def run_two_view_estimator_as_futures(client: Any, two_view_estimator: TwoViewEstimator, keypoints_list: List[Keypoints],
                                      putative_corr_idxs_dict: Dict[Tuple[int, int], np.ndarray], camera_intrinsics: List[Any],
                                      relative_pose_priors: Dict[Tuple[int, int], Any], gt_cameras: List[Optional[Any]],
                                      gt_scene_mesh: Optional[Any], num_shards: Optional[int] = None
                                      ) -> Dict[Tuple[int, int], TWO_VIEW_OUTPUT]:
    """Run the two-view estimator for all image pairs (two_view_estimator.py:531-587): one `client.submit` per GPU shard
    (a direct call per shard when `client` is None); the result dictionary has the keys and order of
    putative_corr_idxs_dict."""
    keys = list(putative_corr_idxs_dict.keys())
    if not keys:
        return {}
    pairs = np.asarray(keys, np.int64).reshape(-1, 2)
    n_shards = max(1, min(num_shards or num_gpu_shards(), len(keys)))
    cost = np.array([float(np.asarray(putative_corr_idxs_dict[k]).size // 2 + 1) for k in keys])
    n_dev = num_gpu_devices()
    jobs = []
    for shard, idx in enumerate(shard_pairs(pairs, n_shards, cost)):
        if len(idx) == 0:
            continue
        used = [int(g) for g in np.unique(pairs[idx])]
        sub_keys = [keys[int(k)] for k in idx]
        jobs.append(({g: keypoints_list[g] for g in used}, {k: putative_corr_idxs_dict[k] for k in sub_keys},
                     {g: camera_intrinsics[g] for g in used}, {k: relative_pose_priors.get(k) for k in sub_keys} if relative_pose_priors else {},
                     {g: gt_cameras[g] for g in used} if gt_cameras is not None else None, gt_scene_mesh, shard % n_dev))
    if client is None:
        parts = [_apply_two_view_estimator_shard(two_view_estimator, *job) for job in jobs]
    else:
        est_future = client.scatter(two_view_estimator, broadcast=False)
        parts = client.gather([client.submit(_apply_two_view_estimator_shard, est_future, *job) for job in jobs])
    merged: Dict[Tuple[int, int], TWO_VIEW_OUTPUT] = {}
    for part in parts:
        merged.update(part)
    return {(int(k[0]), int(k[1])): merged[(int(k[0]), int(k[1]))] for k in keys}


def unzip_two_view_results(two_view_results: Dict[Tuple[int, int], TWO_VIEW_OUTPUT]):
    """gtsfm/runner/gtsfm_runner_base.py:415-440: one dictionary per element of TWO_VIEW_OUTPUT, failed pairs dropped."""
    i2Ri1_dict, i2Ui1_dict, v_corr_idxs_dict, two_view_reports_dict = {}, {}, {}, {}
    for key, out in two_view_results.items():
        if out[0] is None or out[1] is None:
            continue
        i2Ri1_dict[key], i2Ui1_dict[key], v_corr_idxs_dict[key], two_view_reports_dict[key] = out[0], out[1], out[2], out[5]
    return i2Ri1_dict, i2Ui1_dict, v_corr_idxs_dict, two_view_reports_dict
