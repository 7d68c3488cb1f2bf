"""TwoViewEstimationReport: same fields as the reference's gtsfm/common/two_view_estimation_report.py:13-57, and
generate_two_view_report as in gtsfm/two_view_estimator.py:354-393."""
from dataclasses import dataclass
from typing import Optional

import numpy as np

from gtsfm_b200.common.types import Rot3, Unit3


@dataclass(frozen=False)
class TwoViewEstimationReport:
    v_corr_idxs: np.ndarray
    num_inliers_est_model: float
    inlier_ratio_est_model: Optional[float] = None
    num_inliers_gt_model: Optional[float] = None
    inlier_ratio_gt_model: Optional[float] = None
    v_corr_idxs_inlier_mask_gt: Optional[np.ndarray] = None
    R_error_deg: Optional[float] = None
    U_error_deg: Optional[float] = None
    i2Ri1: Optional[Rot3] = None
    i2Ui1: Optional[Unit3] = None
    reproj_error_gt_model: Optional[np.ndarray] = None
    inlier_avg_reproj_error_gt_model: Optional[float] = None
    outlier_avg_reproj_error_gt_model: Optional[float] = None


def generate_two_view_report(inlier_ratio_est_model: float, v_corr_idxs: np.ndarray, R_error_deg: Optional[float] = None,
                             U_error_deg: Optional[float] = None, v_corr_idxs_inlier_mask_gt: Optional[np.ndarray] = None,
                             reproj_error_gt_model: Optional[np.ndarray] = None) -> TwoViewEstimationReport:
    """two_view_estimator.py:354-393 (single pair, host arrays)."""
    if v_corr_idxs_inlier_mask_gt is not None and reproj_error_gt_model is not None:
        num_inliers_gt_model = np.count_nonzero(v_corr_idxs_inlier_mask_gt)
        inlier_ratio_gt_model = num_inliers_gt_model / v_corr_idxs.shape[0] if len(v_corr_idxs) > 0 else 0.0
        with np.errstate(all="ignore"):
            inl = reproj_error_gt_model[v_corr_idxs_inlier_mask_gt]
            out = reproj_error_gt_model[np.logical_not(v_corr_idxs_inlier_mask_gt)]
            inlier_avg = float(np.mean(inl)) if inl.size else float("nan")
            outlier_avg = float(np.nanmean(out)) if np.any(~np.isnan(out)) else float("nan")
    else:
        num_inliers_gt_model, inlier_ratio_gt_model, inlier_avg, outlier_avg = 0, float("nan"), float("nan"), float("nan")
    return TwoViewEstimationReport(
        inlier_ratio_est_model=inlier_ratio_est_model, num_inliers_est_model=v_corr_idxs.shape[0],
        num_inliers_gt_model=num_inliers_gt_model, inlier_ratio_gt_model=inlier_ratio_gt_model,
        v_corr_idxs_inlier_mask_gt=v_corr_idxs_inlier_mask_gt, v_corr_idxs=v_corr_idxs, R_error_deg=R_error_deg,
        U_error_deg=U_error_deg, reproj_error_gt_model=reproj_error_gt_model,
        inlier_avg_reproj_error_gt_model=inlier_avg, outlier_avg_reproj_error_gt_model=outlier_avg)
