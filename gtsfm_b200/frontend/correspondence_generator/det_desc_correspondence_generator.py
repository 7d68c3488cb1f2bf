"""DetDescCorrespondenceGenerator: drop-in for the reference's
gtsfm/frontend/correspondence_generator/det_desc_correspondence_generator.py:17-85 -- same constructor, same
generate_correspondences(client, images, image_pairs), same return structures -- with the matcher submitted as ONE task per
GPU shard of the pair list instead of one task per pair.

Detection / description stays what it is in the reference: one task per image on whatever detector plugin the config
names (out of this path's scope).  The features are then gathered where the shard tasks run; each shard task uploads the
images its pairs name once and matches all of them in one device batch (`match_batch` of the B200 TwoWayMatcher or of a
MatcherCacher around it).  A matcher plugin without `match_batch` (any other MatcherBase) is submitted per pair exactly as
the reference does.
"""
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from gtsfm_b200.common.types import Keypoints
from gtsfm_b200.frontend.correspondence_generator.correspondence_generator_base import CorrespondenceGeneratorBase
from gtsfm_b200.frontend.matcher.matcher_base import MatcherBase
from gtsfm_b200.two_view_batch import num_gpu_devices, num_gpu_shards, shard_pairs


def _apply_det_desc(det_desc, image) -> Tuple[Keypoints, np.ndarray]:
    return det_desc.detect_and_describe(image)


def _get_image_shape(image) -> Tuple[int, int, int]:
    return image.shape


def _apply_matcher(feature_matcher: MatcherBase, features_i1, features_i2, im_shape_i1, im_shape_i2) -> np.ndarray:
    return feature_matcher.match(features_i1[0], features_i2[0], features_i1[1], features_i2[1], im_shape_i1, im_shape_i2)


def _apply_matcher_shard(feature_matcher, features: Dict[int, Tuple[Keypoints, np.ndarray]], im_shapes: Dict[int, Tuple[int, int, int]],
                         pairs: np.ndarray, device: int) -> List[np.ndarray]:
    """All pairs of one shard in one device batch; `features` / `im_shapes` hold only the images the shard names."""
    used = sorted(features)
    remap = {g: k for k, g in enumerate(used)}
    local = np.array([[remap[int(i)], remap[int(j)]] for i, j in pairs], np.int32).reshape(-1, 2)
    descs = [features[g][1] for g in used]
    if hasattr(feature_matcher, "cache_key"):        # MatcherCacher: hits from disk, all misses in one device batch
        return feature_matcher.match_batch([features[g][0] for g in used], descs, [im_shapes[g] for g in used], local, device=device)
    return feature_matcher.match_batch(descs, local, device=device)


class DetDescCorrespondenceGenerator(CorrespondenceGeneratorBase):
    """Pair-wise matching of descriptors, batched per GPU."""

    def __init__(self, matcher: MatcherBase, detector_descriptor: Any, num_shards: Optional[int] = None) -> None:
        self._detector_descriptor = detector_descriptor
        self._matcher = matcher
        self._num_shards = num_shards          # None: one shard per visible GPU

    def __repr__(self) -> str:
        return f"""
        DetDescCorrespondenceGenerator:
           {self._detector_descriptor}
           {self._matcher}
        """

    def generate_correspondences(self, client: Any, images: List[Any], image_pairs: List[Tuple[int, int]]
                                 ) -> Tuple[List[Keypoints], Dict[Tuple[int, int], np.ndarray]]:
        pairs = np.asarray(image_pairs, np.int64).reshape(-1, 2)
        # ---- features and image shapes: one task per image, as in the reference (det_desc_correspondence_generator.py:66-69)
        if client is None:
            features = [_apply_det_desc(self._detector_descriptor, im) for im in images]
            shapes = [_get_image_shape(im) for im in images]
        else:
            det_desc_future = client.scatter(self._detector_descriptor, broadcast=False)
            features = client.gather([client.submit(_apply_det_desc, det_desc_future, im) for im in images])
            shapes = client.gather([client.submit(_get_image_shape, im) for im in images])
        keypoints_list = [f[0] for f in features]

        if not hasattr(self._matcher, "match_batch"):
            # a per-pair plugin: the reference's own submission loop (det_desc_correspondence_generator.py:71-81)
            if client is None:
                out = {(int(i1), int(i2)): _apply_matcher(self._matcher, features[i1], features[i2], shapes[i1], shapes[i2]) for i1, i2 in pairs}
            else:
                mf = client.scatter(self._matcher, broadcast=False)
                out = client.gather({(int(i1), int(i2)): client.submit(_apply_matcher, mf, features[i1], features[i2], shapes[i1], shapes[i2])
                                     for i1, i2 in pairs})
            return keypoints_list, out

        # ---- one matcher task per GPU shard: contiguous blocks of the i1-sorted pair list, balanced by N1 * N2
        n_shards = max(1, min(self._num_shards or num_gpu_shards(), len(pairs)))
        cost = np.array([float(len(features[i][1])) * float(len(features[j][1])) + 1.0 for i, j in pairs])
        shards = shard_pairs(pairs, n_shards, cost)
        jobs = []
        n_dev = num_gpu_devices()
        for shard, idx in enumerate(shards):
            if len(idx) == 0:
                continue
            used = np.unique(pairs[idx])
            jobs.append((idx, {int(g): features[int(g)] for g in used}, {int(g): shapes[int(g)] for g in used}, pairs[idx], shard % n_dev))
        if client is None:
            results = [_apply_matcher_shard(self._matcher, f, s, p, d) for _, f, s, p, d in jobs]
        else:
            mf = client.scatter(self._matcher, broadcast=False)
            results = client.gather([client.submit(_apply_matcher_shard, mf, f, s, p, d) for _, f, s, p, d in jobs])
        per_pair: List[Optional[np.ndarray]] = [None] * len(pairs)
        for (idx, *_), res in zip(jobs, results):
            for k, m in zip(idx, res):
                per_pair[int(k)] = m
        # the reference's dictionary, in the order of image_pairs
        return keypoints_list, {(int(i1), int(i2)): per_pair[k] for k, (i1, i2) in enumerate(pairs)}
