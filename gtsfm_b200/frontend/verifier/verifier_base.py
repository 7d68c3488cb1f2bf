"""Verifier plugin interface, same as the reference's gtsfm/frontend/verifier/verifier_base.py:14-82."""
import abc
from typing import Optional, Tuple

import numpy as np

from gtsfm_b200.common.types import Cal3Bundler, GTSFMProcess, Keypoints, Rot3, UiMetadata, Unit3

NUM_MATCHES_REQ_E_MATRIX = 5
NUM_MATCHES_REQ_F_MATRIX = 8


class VerifierBase(GTSFMProcess):
    """Verifiers take putative matches and return the relative pose and the geometrically verified matches."""

    @staticmethod
    def get_ui_metadata() -> UiMetadata:
        return UiMetadata(display_name="Verifier", input_products=("Keypoints", "Putative Correspondences", "Camera Intrinsics"),
                          output_products=("Relative Rotation", "Relative Translation", "Verified Correspondences", "Inlier Ratio"),
                          parent_plate="Two-View Estimator")

    def __init__(self, use_intrinsics_in_verification: bool, estimation_threshold_px: float) -> None:
        self._use_intrinsics_in_verification = use_intrinsics_in_verification
        self._estimation_threshold_px = estimation_threshold_px
        self._min_matches = NUM_MATCHES_REQ_E_MATRIX if use_intrinsics_in_verification else NUM_MATCHES_REQ_F_MATRIX
        # i2Ri1 = None, i2Ui1 = None, no verified correspondences, inlier ratio 0 (verifier_base.py:56)
        self._failure_result = (None, None, np.array([], dtype=np.uint64), 0.0)

    @abc.abstractmethod
    def verify(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, match_indices: np.ndarray,
               camera_intrinsics_i1: Cal3Bundler, camera_intrinsics_i2: Cal3Bundler
               ) -> Tuple[Optional[Rot3], Optional[Unit3], np.ndarray, float]:
        """(i2Ri1, i2Ui1, verified match indices (subset of match_indices, same order), inlier ratio)."""
