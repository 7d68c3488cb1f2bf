"""TriangulationOptions / TriangulationSamplingMode: the configuration types TwoViewEstimator takes
(gtsfm/data_association/point3d_initializer.py:41-115), so that the reference's YAML (`configs/sift_front_end.yaml:50-55`)
instantiates against this package.  The triangulation itself runs on the device inside gtsfm_two_view_ba_batch; the
multi-view Point3dInitializer (data association, SURVEY.md 8(f): out of scope) is not part of this package.
"""
import sys
from enum import Enum
from typing import NamedTuple

import numpy as np


class TriangulationSamplingMode(str, Enum):
    """point3d_initializer.py:47-60."""
    NO_RANSAC = "NO_RANSAC"
    RANSAC_SAMPLE_UNIFORM = "RANSAC_SAMPLE_UNIFORM"
    RANSAC_SAMPLE_BIASED_BASELINE = "RANSAC_SAMPLE_BIASED_BASELINE"
    RANSAC_TOPK_BASELINES = "RANSAC_TOPK_BASELINES"


class TriangulationOptions(NamedTuple):
    """point3d_initializer.py:63-92 (same fields, same defaults)."""
    mode: TriangulationSamplingMode
    reproj_error_threshold: float = np.inf
    min_triangulation_angle: float = 0.0
    min_inlier_ratio: float = 0.1
    confidence: float = 0.9999
    dyn_num_hypotheses_multiplier: float = 3.0
    min_num_hypotheses: int = 0
    max_num_hypotheses: int = sys.maxsize
