"""MatcherCacher for the batched GPU matcher.  Host only.

Same cache entries as the reference's gtsfm/frontend/cacher/matcher_cacher.py:27-192, so a cache warmed by either
implementation serves the other:
  key    "<type(matcher).__name__>_<sha1>" (matcher_cacher.py:34,80); the sha1 is taken over ONE concatenated array made, per image,
         of the first 10 keypoint coordinates, responses (if any), scales (if any) and the first 10 descriptor rows, followed
         by [h1, w1, c1, h2, w2, c2] (matcher_cacher.py:62-78).  numpy's dtype promotion inside that concatenate is part
         of the key, so the pieces are concatenated here exactly once, with their original dtypes.
  file   <cache root>/matcher/<key>.pbz2 = bz2(pickle(match_indices)) (matcher_cacher.py:42-44, utils/io.py:610-630).
The B200 `TwoWayMatcher` keeps the reference's class name, hence the same namespace.

`match()` is the reference's per-pair call (hit -> return, miss -> compute + store).  `match_batch()` is the batched
form: hits are read from disk, ALL misses go to the device in one `match_batch` call of the wrapped matcher, and their
results are stored one entry per pair.
"""
from pathlib import Path
from typing import List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200.common.types import Keypoints
from gtsfm_b200.frontend.cacher import cache_io
from gtsfm_b200.frontend.matcher.matcher_base import MatcherBase

NUM_KEYPOINTS_TO_SAMPLE_FOR_HASH = 10


def image_key_parts(keypoints: Keypoints, descriptors: np.ndarray) -> List[np.ndarray]:
    """The arrays one image contributes to the key, in the reference's order (matcher_cacher.py:64-71)."""
    k = NUM_KEYPOINTS_TO_SAMPLE_FOR_HASH
    parts = [keypoints.coordinates[:k].flatten()]
    if keypoints.responses is not None:
        parts.append(keypoints.responses[:k].flatten())
    if keypoints.scales is not None:
        parts.append(keypoints.scales[:k].flatten())
    parts.append(descriptors[:k].flatten())
    return parts


def matcher_cache_key(matcher_name: str, parts_i1: List[np.ndarray], parts_i2: List[np.ndarray],
                      im_shape_i1: Tuple[int, int, int], im_shape_i2: Tuple[int, int, int]) -> str:
    h1, w1, c1 = im_shape_i1
    h2, w2, c2 = im_shape_i2
    blob = np.concatenate(parts_i1 + parts_i2 + [np.array([h1, w1, c1, h2, w2, c2])])
    return "{}_{}".format(matcher_name, cache_io.hash_numpy(blob))


class MatcherCacher(MatcherBase):
    """Cacher for matcher output on disk, keyed on the input."""

    def __init__(self, matcher_obj: MatcherBase, cache_root: Optional[Path] = None) -> None:
        super().__init__()
        self._matcher = matcher_obj
        self._matcher_obj_key = type(self._matcher).__name__
        self._cache_root = None if cache_root is None else Path(cache_root)
        self.hits = 0
        self.misses = 0

    def __repr__(self) -> str:
        return f"MatcherCacher({self._matcher!r})"

    def _path(self, key: str) -> Path:
        root = self._cache_root if self._cache_root is not None else cache_io.cache_root()
        return root / "matcher" / "{}.pbz2".format(key)

    def cache_key(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, descriptors_i1: np.ndarray, descriptors_i2: np.ndarray,
                  im_shape_i1: Tuple[int, int, int], im_shape_i2: Tuple[int, int, int]) -> str:
        return matcher_cache_key(self._matcher_obj_key, image_key_parts(keypoints_i1, descriptors_i1),
                                 image_key_parts(keypoints_i2, descriptors_i2), im_shape_i1, im_shape_i2)

    def match(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, descriptors_i1: np.ndarray, descriptors_i2: np.ndarray,
              im_shape_i1: Tuple[int, int, int], im_shape_i2: Tuple[int, int, int]) -> np.ndarray:
        path = self._path(self.cache_key(keypoints_i1, keypoints_i2, descriptors_i1, descriptors_i2, im_shape_i1, im_shape_i2))
        cached = cache_io.read_entry(path)
        if cached is not None:
            self.hits += 1
            return cached
        self.misses += 1
        match_indices = self._matcher.match(keypoints_i1=keypoints_i1, keypoints_i2=keypoints_i2, descriptors_i1=descriptors_i1,
                                            descriptors_i2=descriptors_i2, im_shape_i1=im_shape_i1, im_shape_i2=im_shape_i2)
        cache_io.write_entry(match_indices, path)
        return match_indices

    def match_batch(self, keypoints: Sequence[Keypoints], descriptors: Sequence[np.ndarray], im_shapes: Sequence[Tuple[int, int, int]],
                    pairs: np.ndarray, device: Optional[int] = None) -> List[np.ndarray]:
        """Per-pair results for `pairs` (P, 2) over per-image inputs; one device batch for all cache misses."""
        pairs = np.asarray(pairs).reshape(-1, 2)
        parts = {}
        for i in np.unique(pairs):
            parts[int(i)] = image_key_parts(keypoints[int(i)], descriptors[int(i)])
        paths, results, missing = [], [], []
        for p, (i, j) in enumerate(pairs):
            path = self._path(matcher_cache_key(self._matcher_obj_key, parts[int(i)], parts[int(j)], im_shapes[int(i)], im_shapes[int(j)]))
            paths.append(path)
            cached = cache_io.read_entry(path)
            results.append(cached)
            if cached is None:
                missing.append(p)
        self.hits += len(pairs) - len(missing)
        self.misses += len(missing)
        if missing:
            # the device only sees the images the misses name, renumbered densely
            used = np.unique(pairs[missing])
            remap = {int(g): k for k, g in enumerate(used)}
            sub_pairs = np.array([[remap[int(i)], remap[int(j)]] for i, j in pairs[missing]], np.int32)
            fresh = self._matcher.match_batch([descriptors[int(g)] for g in used], sub_pairs, **({} if device is None else {"device": device}))
            for p, m in zip(missing, fresh):
                cache_io.write_entry(m, paths[p])
                results[p] = m
        return results
