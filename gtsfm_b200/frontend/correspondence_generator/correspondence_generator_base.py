"""Base class for correspondence generators: same interface as the reference's
gtsfm/frontend/correspondence_generator/correspondence_generator_base.py:15-34."""
from abc import abstractmethod
from typing import Any, Dict, List, Tuple

import numpy as np

from gtsfm_b200.common.types import Keypoints


class CorrespondenceGeneratorBase:
    """Base class for correspondence generators."""

    @abstractmethod
    def generate_correspondences(self, client: Any, images: List[Any], image_pairs: List[Tuple[int, int]]
                                 ) -> Tuple[List[Keypoints], Dict[Tuple[int, int], np.ndarray]]:
        """client: Dask client (or None: run in this process); images: all images (futures when a client is given);
        image_pairs: (i1, i2) index pairs.  Returns the keypoints of every image and the putative correspondences of
        every pair as indices of keypoints."""
