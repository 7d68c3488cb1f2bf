"""InlierSupportProcessor: drop-in for the reference's gtsfm/frontend/inlier_support_processor.py:19-95 (same constructor
arguments, same decisions), plus the batched form: the decision of every pair of a device batch comes from
gtsfm_two_view_report_batch (`keep`), so no per-pair Python stage is left between verification and the two-view results."""
import dataclasses
from typing import Optional, Tuple

import numpy as np

from gtsfm_b200.common.two_view_estimation_report import TwoViewEstimationReport
from gtsfm_b200.common.types import Rot3, Unit3


class InlierSupportProcessor:
    def __init__(self, min_num_inliers_est_model: int, min_inlier_ratio_est_model: float) -> None:
        self._min_num_inliers_est_model = min_num_inliers_est_model
        self._min_inlier_ratio_est_model = min_inlier_ratio_est_model

    def run_inlier_support(self, i2Ri1: Optional[Rot3], i2Ui1: Optional[Unit3], v_corr_idxs: np.ndarray,
                           two_view_report: TwoViewEstimationReport
                           ) -> Tuple[Optional[Rot3], Optional[Unit3], np.ndarray, TwoViewEstimationReport]:
        """inlier_support_processor.py:39-95: the ratio test first, then the count test (valid models only)."""
        failure_result = (None, None, np.array([], dtype=np.uint64),
                          TwoViewEstimationReport(v_corr_idxs=v_corr_idxs, num_inliers_est_model=0))
        two_view_report_post_isp = dataclasses.replace(two_view_report)
        insufficient_inliers = two_view_report.num_inliers_est_model < self._min_num_inliers_est_model
        valid_model = two_view_report.num_inliers_est_model > 0
        if two_view_report.inlier_ratio_est_model < self._min_inlier_ratio_est_model:
            return failure_result
        if valid_model and insufficient_inliers:
            return failure_result
        return i2Ri1, i2Ui1, v_corr_idxs, two_view_report_post_isp
