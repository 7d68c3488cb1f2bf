"""Cacher compatibility, host only.

Golden keys / files: tests/golden/cacher_keys.npz, recorded from the unmodified reference MatcherCacher and
TwoViewEstimatorCacher by oracle/gen_golden_cacher.py.  Hit / miss behaviour mirrors the reference's
tests/frontend/cacher/test_matcher_cacher.py:33-117 and tests/test_two_view_estimator_cacher.py.
"""
import bz2
import os
import pickle
from unittest.mock import MagicMock

import numpy as np
import pytest

from gtsfm_b200.common.types import Keypoints
from gtsfm_b200.frontend.cacher import cache_io
from gtsfm_b200.frontend.cacher.matcher_cacher import MatcherCacher
from gtsfm_b200.two_view_estimator_cacher import TwoViewEstimatorCacher, two_view_cache_key


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "cacher_keys.npz"))


def matcher_case(g, k):
    p = f"m{k}_"
    extras = bool(g[p + "extras"])
    kp1 = Keypoints(coordinates=g[p + "c1"], scales=g[p + "s1"] if extras else None, responses=g[p + "r1"] if extras else None)
    kp2 = Keypoints(coordinates=g[p + "c2"], scales=g[p + "s2"] if extras else None, responses=g[p + "r2"] if extras else None)
    sh = [int(v) for v in g[p + "sh"]]
    return kp1, kp2, g[p + "d1"], g[p + "d2"], tuple(sh[:3]), tuple(sh[3:]), g[p + "result"], str(g[p + "file"]), g[p + "blob"].tobytes()


class TwoWayMatcher:            # the cache namespace is the wrapped object's class name
    def __init__(self):
        self.match = MagicMock()
        self.match_batch = MagicMock()


@pytest.mark.parametrize("k", [0, 1, 2])
def test_matcher_key_and_entry_equal_the_reference(gold, tmp_path, k):
    kp1, kp2, d1, d2, sh1, sh2, result, fname, blob = matcher_case(gold, k)
    inner = TwoWayMatcher()
    inner.match.return_value = result
    cacher = MatcherCacher(inner, cache_root=tmp_path)
    assert cacher.cache_key(kp1, kp2, d1, d2, sh1, sh2) + ".pbz2" == fname
    # miss: computes through the wrapped matcher (keyword call, as matcher_cacher.py:173-180) and writes the reference's file
    got = cacher.match(keypoints_i1=kp1, keypoints_i2=kp2, descriptors_i1=d1, descriptors_i2=d2, im_shape_i1=sh1, im_shape_i2=sh2)
    np.testing.assert_array_equal(got, result)
    inner.match.assert_called_once()
    assert set(inner.match.call_args.kwargs) == {"keypoints_i1", "keypoints_i2", "descriptors_i1", "descriptors_i2", "im_shape_i1", "im_shape_i2"}
    written = tmp_path / "matcher" / fname
    assert written.exists()
    with bz2.BZ2File(written, "rb") as f:
        back = pickle.load(f)
    np.testing.assert_array_equal(back, result)
    assert back.dtype == result.dtype
    # hit: no second computation
    again = cacher.match(keypoints_i1=kp1, keypoints_i2=kp2, descriptors_i1=d1, descriptors_i2=d2, im_shape_i1=sh1, im_shape_i2=sh2)
    np.testing.assert_array_equal(again, result)
    inner.match.assert_called_once()
    assert (cacher.hits, cacher.misses) == (1, 1)


def test_entry_written_by_the_reference_is_a_hit(gold, tmp_path):
    kp1, kp2, d1, d2, sh1, sh2, result, fname, blob = matcher_case(gold, 0)
    (tmp_path / "matcher").mkdir()
    (tmp_path / "matcher" / fname).write_bytes(blob)            # the exact file the reference cacher wrote
    inner = TwoWayMatcher()
    got = MatcherCacher(inner, cache_root=tmp_path).match(kp1, kp2, d1, d2, sh1, sh2)
    np.testing.assert_array_equal(got, result)
    assert got.dtype == np.uint32
    inner.match.assert_not_called()


def test_corrupt_entry_is_removed_and_recomputed(gold, tmp_path):
    kp1, kp2, d1, d2, sh1, sh2, result, fname, _ = matcher_case(gold, 1)
    (tmp_path / "matcher").mkdir()
    (tmp_path / "matcher" / fname).write_bytes(b"not a bz2 stream")
    inner = TwoWayMatcher()
    inner.match.return_value = result
    got = MatcherCacher(inner, cache_root=tmp_path).match(kp1, kp2, d1, d2, sh1, sh2)
    np.testing.assert_array_equal(got, result)
    inner.match.assert_called_once()
    assert cache_io.read_entry(tmp_path / "matcher" / fname) is not None


def test_match_batch_sends_only_misses_to_the_device(tmp_path):
    rng = np.random.default_rng(0)
    n_img = 5
    kps = [Keypoints(coordinates=rng.uniform(0, 100, (20, 2))) for _ in range(n_img)]
    descs = [rng.uniform(0, 255, (20, 16)).astype(np.float32) for _ in range(n_img)]
    shapes = [(10 + i, 20, 3) for i in range(n_img)]
    pairs = np.array([(0, 1), (0, 3), (1, 3), (2, 4), (3, 4)], np.int32)
    answers = {tuple(p): rng.integers(0, 20, (k + 1, 2)).astype(np.uint32) for k, p in enumerate(pairs)}
    answers[(2, 4)] = np.array([])                              # an empty result is cached too, as the reference does

    inner = TwoWayMatcher()
    cacher = MatcherCacher(inner, cache_root=tmp_path)
    # warm two entries through the per-pair path
    for p in ((0, 3), (2, 4)):
        inner.match.return_value = answers[p]
        cacher.match(kps[p[0]], kps[p[1]], descs[p[0]], descs[p[1]], shapes[p[0]], shapes[p[1]])

    def fake_batch(sub_descs, sub_pairs):
        # misses are (0,1), (1,3), (3,4): images {0,1,3,4} renumbered 0..3
        assert len(sub_descs) == 4
        used = [0, 1, 3, 4]
        for a, b in zip(sub_descs, (descs[g] for g in used)):
            assert a is b
        return [answers[(used[i], used[j])] for i, j in sub_pairs]

    inner.match_batch.side_effect = fake_batch
    out = cacher.match_batch(kps, descs, shapes, pairs)
    inner.match_batch.assert_called_once()
    np.testing.assert_array_equal(inner.match_batch.call_args.args[1], [[0, 1], [1, 2], [2, 3]])
    for p, m in zip(pairs, out):
        np.testing.assert_array_equal(m, answers[tuple(p)])
    # everything is warm now: a second batch never reaches the matcher
    out2 = cacher.match_batch(kps, descs, shapes, pairs)
    inner.match_batch.assert_called_once()
    for a, b in zip(out, out2):
        np.testing.assert_array_equal(a, b)
    # and the per-pair reference-style call hits the entries the batch wrote
    inner.match.reset_mock()
    got = cacher.match(kps[1], kps[3], descs[1], descs[3], shapes[1], shapes[3])
    np.testing.assert_array_equal(got, answers[(1, 3)])
    inner.match.assert_not_called()


@pytest.mark.parametrize("k", [0, 1, 2])
def test_two_view_key_equals_the_reference(gold, k):
    p = f"t{k}_"
    key = two_view_cache_key(Keypoints(coordinates=gold[p + "c1"]), Keypoints(coordinates=gold[p + "c2"]), gold[p + "idx"])
    assert key == str(gold[p + "key"])


def test_two_view_cacher_hit_miss_and_batch(gold, tmp_path):
    kp1, kp2, idx = Keypoints(coordinates=gold["t0_c1"]), Keypoints(coordinates=gold["t0_c2"]), gold["t0_idx"]
    result = ("R", "U", np.arange(6).reshape(3, 2), {"pre": 1}, None, None)          # any picklable 6-tuple
    inner = MagicMock()
    inner.run_2view.return_value = result
    cacher = TwoViewEstimatorCacher(inner, cache_root=tmp_path)
    got = cacher.run_2view(kp1, kp2, idx, None, None, None, None, None)
    assert got is result
    assert (tmp_path / "two_view_estimator" / (str(gold["t0_key"]) + ".pbz2")).exists()
    got2 = cacher.run_2view(kp1, kp2, idx, None, None, None, None, None)
    inner.run_2view.assert_called_once()
    assert got2[0] == "R" and np.array_equal(got2[2], result[2])
    # batched lookup: pair 0 is warm, pair 1 is not; storing it makes the next lookup a full hit
    kps = [kp1, kp2, Keypoints(coordinates=gold["t1_c1"]), Keypoints(coordinates=gold["t1_c2"])]
    pairs = np.array([(0, 1), (2, 3)])
    putative = [idx, gold["t1_idx"]]
    hits, missing, paths = cacher.lookup_batch(kps, pairs, putative)
    assert list(hits) == [0] and missing == [1]
    assert paths[1].name == str(gold["t1_key"]) + ".pbz2"
    cacher.store_batch(paths, missing, [("R2", "U2", np.zeros((0, 2)), None, None, None)])
    hits, missing, _ = cacher.lookup_batch(kps, pairs, putative)
    assert sorted(hits) == [0, 1] and not missing and hits[1][0] == "R2"


def test_two_view_cacher_run_2view_batch_sends_only_misses(gold, tmp_path):
    """The default configuration wraps the estimator in the cacher (configs/sift_front_end.yaml:36-39), so the per-shard call of
    run_two_view_estimator_as_futures lands here: hits from disk, the misses as ONE batch, everything in the input order."""
    from gtsfm_b200.two_view_estimator import run_two_view_estimator_as_futures
    kps = [Keypoints(coordinates=gold["t0_c1"]), Keypoints(coordinates=gold["t0_c2"]), Keypoints(coordinates=gold["t1_c1"]),
           Keypoints(coordinates=gold["t1_c2"])]
    putative = {(0, 1): gold["t0_idx"], (2, 3): gold["t1_idx"], (1, 0): np.ascontiguousarray(gold["t0_idx"][:5, ::-1])}

    class Inner:
        def __init__(self):
            self.batches = []

        def run_2view_batch(self, keypoints, corr, intrinsics, priors=None, gt_cameras=None, gt_scene_mesh=None, device=None):
            self.batches.append(list(corr.keys()))
            return {k: ("R%d%d" % k, "U", np.asarray(v)[:2], None, None, None) for k, v in corr.items()}

    inner = Inner()
    cacher = TwoViewEstimatorCacher(inner, cache_root=tmp_path)
    cacher.store_batch([cacher._path(two_view_cache_key(kps[2], kps[3], putative[(2, 3)]))], [0], [("warm", "U", np.zeros((0, 2)), None, None, None)])
    out = cacher.run_2view_batch(kps, putative, [None] * 4)
    assert list(out.keys()) == [(0, 1), (2, 3), (1, 0)] and inner.batches == [[(0, 1), (1, 0)]]
    assert out[(2, 3)][0] == "warm" and out[(0, 1)][0] == "R01" and out[(1, 0)][0] == "R10"
    again = cacher.run_2view_batch(kps, putative, [None] * 4)                      # everything is warm now
    assert inner.batches == [[(0, 1), (1, 0)]] and again[(1, 0)][0] == "R10" and cacher.hits == 1 + 3
    # through the batched caller (one shard, no client), and with a per-pair-only estimator behind the cache
    out2 = run_two_view_estimator_as_futures(None, cacher, kps, putative, [None] * 4, {}, None, None, num_shards=1)
    assert list(out2.keys()) == list(putative.keys()) and out2[(0, 1)][0] == "R01"
    per_pair = MagicMock(spec=["run_2view"])
    per_pair.run_2view.return_value = ("P", "U", np.zeros((0, 2)), None, None, None)
    cold = TwoViewEstimatorCacher(per_pair, cache_root=tmp_path / "cold")
    out3 = run_two_view_estimator_as_futures(None, cold, kps, putative, [None] * 4, {}, None, None, num_shards=1)
    assert per_pair.run_2view.call_count == 3 and out3[(2, 3)][0] == "P"
    # a foreign estimator without the cacher: the reference's per-pair loop inside the shard
    out4 = run_two_view_estimator_as_futures(None, per_pair, kps, putative, [None] * 4, {}, None, None, num_shards=1)
    assert per_pair.run_2view.call_count == 6 and list(out4.keys()) == list(putative.keys())
