"""Cache of per-pair two-view results, entry-compatible with the reference's gtsfm/two_view_estimator_cacher.py:36-119
(SURVEY 8(f) rank 3).  Host only.

  key    sha1 over concat(kps_i1.coordinates[idx[:10, 0]].flatten(), kps_i2.coordinates[idx[:10, 1]].flatten()), idx = the
         putative correspondences (two_view_estimator_cacher.py:49-63) - no estimator name in it
  file   <cache root>/two_view_estimator/<key>.pbz2 = bz2(pickle(result)) (two_view_estimator_cacher.py:45-47)
The payload is whatever the wrapped estimator's run_2view returns (in the reference the TWO_VIEW_OUTPUT 6-tuple,
two_view_estimator.py:44-51); it is pickled as is, so entries exchange with the reference when both sides hold the
reference's own result types (gtsam importable).

`TwoViewEstimatorCacher.run_2view` is the reference's per-pair call.  `lookup_batch` / `store_batch` serve a batched
estimator (gtsfm_b200.two_view_batch): look every pair up first, send only the misses to the device, store their results.
"""
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200.common.types import Keypoints
from gtsfm_b200.frontend.cacher import cache_io

NUM_CORRESPONDENCES_TO_SAMPLE_FOR_HASH = 10


def two_view_cache_key(keypoints_i1: Keypoints, keypoints_i2: Keypoints, putative_corr_idxs: np.ndarray) -> str:
    sampled = putative_corr_idxs[:NUM_CORRESPONDENCES_TO_SAMPLE_FOR_HASH]
    return cache_io.hash_numpy(np.concatenate([keypoints_i1.coordinates[sampled[:, 0]].flatten(),
                                               keypoints_i2.coordinates[sampled[:, 1]].flatten()]))


class TwoViewEstimatorCacher:
    """Caches two-view relative pose estimation results for an image pair."""

    def __init__(self, two_view_estimator_obj: Any, cache_root: Optional[Path] = None) -> None:
        self._two_view_estimator = two_view_estimator_obj
        self._cache_root = None if cache_root is None else Path(cache_root)
        self.hits = 0
        self.misses = 0

    def __repr__(self) -> str:
        return self._two_view_estimator.__repr__()

    def _path(self, key: str) -> Path:
        root = self._cache_root if self._cache_root is not None else cache_io.cache_root()
        return root / "two_view_estimator" / "{}.pbz2".format(key)

    def run_2view(self, keypoints_i1: Keypoints, keypoints_i2: Keypoints, putative_corr_idxs: np.ndarray, camera_intrinsics_i1,
                  camera_intrinsics_i2, i2Ti1_prior=None, gt_camera_i1=None, gt_camera_i2=None, gt_scene_mesh: Optional[Any] = None):
        path = self._path(two_view_cache_key(keypoints_i1, keypoints_i2, putative_corr_idxs))
        result = cache_io.read_entry(path)
        if result is not None:
            self.hits += 1
            return result
        self.misses += 1
        result = self._two_view_estimator.run_2view(
            keypoints_i1=keypoints_i1, keypoints_i2=keypoints_i2, putative_corr_idxs=putative_corr_idxs,
            camera_intrinsics_i1=camera_intrinsics_i1, camera_intrinsics_i2=camera_intrinsics_i2, i2Ti1_prior=i2Ti1_prior,
            gt_camera_i1=gt_camera_i1, gt_camera_i2=gt_camera_i2, gt_scene_mesh=gt_scene_mesh)
        cache_io.write_entry(result, path)
        return result

    # ---- batched use
    def lookup_batch(self, keypoints: Sequence[Keypoints], pairs: np.ndarray,
                     putative_corr_idxs: Sequence[np.ndarray]) -> Tuple[Dict[int, Any], List[int], List[Path]]:
        """(hits {pair position -> cached result}, positions of the misses, entry path per pair)."""
        pairs = np.asarray(pairs).reshape(-1, 2)
        hits, missing, paths = {}, [], []
        for p, (i, j) in enumerate(pairs):
            path = self._path(two_view_cache_key(keypoints[int(i)], keypoints[int(j)], putative_corr_idxs[p]))
            paths.append(path)
            cached = cache_io.read_entry(path)
            if cached is None:
                missing.append(p)
            else:
                hits[p] = cached
        self.hits += len(hits)
        self.misses += len(missing)
        return hits, missing, paths

    def run_2view_batch(self, keypoints, putative_corr_idxs_dict: Dict[Tuple[int, int], np.ndarray], camera_intrinsics,
                        relative_pose_priors=None, gt_cameras=None, gt_scene_mesh: Optional[Any] = None,
                        device: Optional[int] = None) -> Dict[Tuple[int, int], Any]:
        """TwoViewEstimator.run_2view_batch behind the cache (what run_two_view_estimator_as_futures calls per GPU shard when
        the configuration wraps the estimator, configs/sift_front_end.yaml:36-39): the pairs with an entry come from disk, the
        misses go to the wrapped estimator as ONE batch (per pair when it has no batched path) and are written back.  Keys and
        order of the input dictionary."""
        putative = {(int(k[0]), int(k[1])): v for k, v in putative_corr_idxs_dict.items()}
        keys = list(putative.keys())
        if not keys:
            return {}
        hits, missing, paths = self.lookup_batch(keypoints, np.asarray(keys).reshape(-1, 2), [putative[k] for k in keys])
        fresh: Dict[Tuple[int, int], Any] = {}
        if missing:
            miss_keys = [keys[p] for p in missing]
            sub = {k: putative[k] for k in miss_keys}
            priors = {k: relative_pose_priors.get(k) for k in miss_keys} if relative_pose_priors else None
            if hasattr(self._two_view_estimator, "run_2view_batch"):
                fresh = self._two_view_estimator.run_2view_batch(keypoints, sub, camera_intrinsics, priors, gt_cameras, gt_scene_mesh,
                                                                 device=device)
            else:
                for (i, j) in miss_keys:
                    fresh[(i, j)] = self._two_view_estimator.run_2view(
                        keypoints[i], keypoints[j], sub[(i, j)], camera_intrinsics[i], camera_intrinsics[j],
                        priors[(i, j)] if priors else None, gt_cameras[i] if gt_cameras is not None else None,
                        gt_cameras[j] if gt_cameras is not None else None, gt_scene_mesh)
            self.store_batch(paths, missing, [fresh[k] for k in miss_keys])
        return {k: (hits[p] if p in hits else fresh[k]) for p, k in enumerate(keys)}

    @staticmethod
    def store_batch(paths: Sequence[Path], positions: Sequence[int], results: Sequence[Any]) -> None:
        for p, r in zip(positions, results):
            cache_io.write_entry(r, paths[p])
