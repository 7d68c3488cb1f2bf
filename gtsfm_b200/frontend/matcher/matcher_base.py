"""Matcher plugin interface, same as the reference's gtsfm/frontend/matcher/matcher_base.py:14-63."""
import abc
from typing import Tuple

import numpy as np

from gtsfm_b200.common.types import GTSFMProcess, Keypoints, UiMetadata


class MatcherBase(GTSFMProcess):
    """Matchers work on a pair of descriptor arrays and match them by their distance."""

    @staticmethod
    def get_ui_metadata() -> UiMetadata:
        return UiMetadata(display_name="Matcher", input_products=("Keypoints", "Descriptors", "Image Shapes"),
                          output_products=("Putative Correspondences",), parent_plate="DetDescCorrespondenceGenerator")

    @abc.abstractmethod
    def match(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, descriptors_i1: np.ndarray, descriptors_i2: np.ndarray,
              im_shape_i1: Tuple[int, int, int], im_shape_i2: Tuple[int, int, int]) -> np.ndarray:
        """Match indices (sorted by confidence) as an (N, 2) array, N <= min(N1, N2)."""
