"""B200 TwoWayMatcher: mutual nearest neighbour + optional Lowe ratio test, drop-in for the reference's
gtsfm/frontend/matcher/twoway_matcher.py:24-144 (same constructor kwargs, same return conventions).

The arithmetic runs in libgtsfm_b200.so (tcgen05 screening + exact resolve kernels); nothing is computed on the CPU.
The object holds no CUDA state, so it pickles like the reference's (det_desc_correspondence_generator.py:67 scatters it
to Dask workers); a per-process device context is created lazily on first use.
"""
from enum import Enum
from typing import List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200 import _native
from gtsfm_b200.common.types import Keypoints
from gtsfm_b200.frontend.matcher.matcher_base import MatcherBase


class MatchingDistanceType(Enum):
    """Type of distance metric to use for matching descriptors (twoway_matcher.py:17-21)."""
    HAMMING = 1
    EUCLIDEAN = 2


class TwoWayMatcher(MatcherBase):
    """Two way (mutual nearest neighbour) matcher on the GPU."""

    def __init__(self, distance_type: MatchingDistanceType = MatchingDistanceType.EUCLIDEAN,
                 ratio_test_threshold: Optional[float] = None, device: int = 0):
        super().__init__()
        self._distance_type = distance_type
        self._ratio_test_threshold = ratio_test_threshold
        self._device = device

    def _check_distance_type(self):
        if self._distance_type in (MatchingDistanceType.EUCLIDEAN, MatchingDistanceType.HAMMING):
            return
        raise NotImplementedError("The distance type is not in MatchingDistanceType")     # twoway_matcher.py:100

    @property
    def _hamming(self) -> bool:
        return self._distance_type == MatchingDistanceType.HAMMING

    def match(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, descriptors_i1: np.ndarray, descriptors_i2: np.ndarray,
              im_shape_i1: Tuple[int, int, int] = None, im_shape_i2: Tuple[int, int, int] = None) -> np.ndarray:
        """(K, 2) uint32 match indices ordered by ascending 1->2 distance; np.array([]) when there are none."""
        if descriptors_i1 is None or descriptors_i2 is None or descriptors_i1.size == 0 or descriptors_i2.size == 0:
            return np.array([])                                                            # twoway_matcher.py:71-72
        self._check_distance_type()
        ctx = _native.default_context(self._device)
        if self._hamming:                                                                  # uint8 rows, cv2.NORM_HAMMING (twoway_matcher.py:97-98)
            m = ctx.match_pair_u8(np.asarray(descriptors_i1, np.uint8), np.asarray(descriptors_i2, np.uint8), self._ratio_test_threshold)
        else:
            m = ctx.match_pair(np.asarray(descriptors_i1, np.float32), np.asarray(descriptors_i2, np.float32),
                               self._ratio_test_threshold)
        return m if m.shape[0] else np.array([])                                           # twoway_matcher.py:80-81

    def match_batch(self, descriptors: Sequence[np.ndarray], pairs: np.ndarray, device: Optional[int] = None) -> List[np.ndarray]:
        """All pairs in one device batch (replaces one Dask task per pair,
        det_desc_correspondence_generator.py:71-81).  descriptors: per-image arrays; pairs: (P, 2) image indices; device: the
        GPU of this shard (default: the matcher's own)."""
        self._check_distance_type()
        ctx = _native.default_context(self._device if device is None else device)
        ctx.clear()
        for i, d in enumerate(descriptors):
            if self._hamming:
                ctx.upload_image_u8(i, d)
            else:
                ctx.upload_image(i, d)
        m, off = ctx.match_batch(pairs, self._ratio_test_threshold)
        return [m[off[p]:off[p + 1]] if off[p + 1] > off[p] else np.array([]) for p in range(len(off) - 1)]
