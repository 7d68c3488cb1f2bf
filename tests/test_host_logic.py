"""Host-side logic that needs no GPU: pair sharding, synthetic scenes, the plugin classes' early-return / error
conventions (mirroring the reference's own tests, tests/frontend/matcher/test_matcher_base.py and
tests/frontend/verifier/test_verifier_base.py), result dictionaries and the record packing used for the multi-GPU gather."""
import pickle

import numpy as np
import pytest

from gtsfm_b200 import synthetic
from gtsfm_b200.common.types import Cal3Bundler, Keypoints
from gtsfm_b200.frontend.matcher.twoway_matcher import MatchingDistanceType, TwoWayMatcher
from gtsfm_b200.frontend.verifier.ransac import Ransac
from gtsfm_b200.two_view_batch import (HEADER_BYTES, ResultExchange, TwoViewBatchResult, block_bytes, block_layout, pack_block, shard_pairs,
                                       unpack_blocks)


# ---------------------------------------------------------------------------------------------------- sharding
@pytest.mark.parametrize("n_shards", [1, 2, 4, 8])
def test_shard_pairs_is_a_contiguous_partition(n_shards):
    pairs = synthetic.make_pairs(40, "exhaustive")
    shards = shard_pairs(pairs, n_shards)
    assert len(shards) == n_shards
    allidx = np.concatenate(shards)
    assert sorted(allidx.tolist()) == list(range(len(pairs)))
    sizes = [len(s) for s in shards]
    assert max(sizes) - min(sizes) <= 1
    # contiguous in (i1, i2) order: every shard touches a compact range of first images
    firsts = [pairs[s][:, 0] for s in shards if len(s)]
    for a, b in zip(firsts[:-1], firsts[1:]):
        assert a.max() <= b.min()


def test_shard_pairs_balances_cost_and_handles_degenerate_inputs():
    pairs = synthetic.make_pairs(30, "exhaustive")
    cost = np.ones(len(pairs)); cost[:10] = 50.0
    shards = shard_pairs(pairs, 4, cost)
    loads = [cost[s].sum() for s in shards]
    assert max(loads) <= sum(loads) / 4 + 50.0
    assert [len(s) for s in shard_pairs(np.zeros((0, 2), np.int32), 3)] == [0, 0, 0]
    one = shard_pairs(np.array([[0, 1]], np.int32), 4)
    assert sum(len(s) for s in one) == 1


# ---------------------------------------------------------------------------------------------------- synthetic scenes
def test_scene_is_deterministic_and_sift_like():
    a = synthetic.make_scene(n_images=4, n_kps=512, seed=11)
    b = synthetic.make_scene(n_images=4, n_kps=512, seed=11)
    c = synthetic.make_scene(n_images=4, n_kps=512, seed=12)
    for x, y in zip(a.descriptors, b.descriptors):
        assert (x == y).all()
    assert any((x != y).any() for x, y in zip(a.descriptors, c.descriptors))
    d = a.descriptors[0]
    assert d.dtype == np.float32 and d.shape == (512, 128)
    assert (d == np.rint(d)).all() and d.min() >= 0 and d.max() <= 255          # integer-valued like cv2 SIFT output
    n = np.linalg.norm(d, axis=1)
    assert abs(np.median(n) - 512) < 8
    assert a.keypoints[0].dtype == np.float64 and a.keypoints[0].shape == (512, 2)
    assert len(a.pairs) == 6 and (a.pairs[:, 0] < a.pairs[:, 1]).all()
    R, t = a.relative_pose(0, 1)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and abs(np.linalg.norm(t) - 1) < 1e-12


def test_superpoint_scene_is_unit_norm_float():
    s = synthetic.make_scene(n_images=2, n_kps=256, dim=256, kind="superpoint", seed=3, ratio=None)
    d = s.descriptors[0]
    assert d.shape == (256, 256) and np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-5)
    assert (d != np.rint(d)).any()


def test_config_table_matches_baseline_configs():
    c = synthetic.CONFIGS
    assert (c["C2"]["n_images"], c["C2"]["n_kps"], c["C2"]["dim"]) == (100, 8192, 128)
    assert len(synthetic.make_pairs(100, "exhaustive")) == 4950
    assert c["C3"]["n_images"] == 1000 and c["C3"]["top_k"] == 50
    assert c["C4"]["dim"] == 256 and c["C4"]["ratio"] is None
    assert c["C5"]["n_hyp"] == 10000 and c["C5"]["max_pairs"] == 20000
    topk = synthetic.make_pairs(200, "topk", top_k=50, angles=np.linspace(0, 2 * np.pi, 200, endpoint=False))
    assert (topk[:, 0] < topk[:, 1]).all() and len(np.unique(topk, axis=0)) == len(topk)


# ---------------------------------------------------------------------------------------------------- plugin conventions
def test_matcher_empty_inputs_return_empty_array_without_touching_the_gpu():
    m = TwoWayMatcher(ratio_test_threshold=0.8)
    kp = Keypoints(coordinates=np.zeros((0, 2)))
    d0 = np.zeros((0, 128), np.float32)
    d1 = np.ones((5, 128), np.float32)
    for a, b in ((d0, d1), (d1, d0), (d0, d0)):
        out = m.match(kp, kp, a, b, (10, 10, 3), (10, 10, 3))
        assert isinstance(out, np.ndarray) and out.size == 0 and out.ndim == 1       # twoway_matcher.py:72: np.array([])


def test_matcher_rejects_unknown_distance_types_loudly():
    m = TwoWayMatcher(distance_type="cosine")
    d = np.ones((5, 32), np.float32)
    with pytest.raises(NotImplementedError):                                           # twoway_matcher.py:100
        m.match(None, None, d, d)
    assert TwoWayMatcher(distance_type=MatchingDistanceType.HAMMING)._hamming


def test_verifier_gates_and_failure_tuple():
    v = Ransac(use_intrinsics_in_verification=True, estimation_threshold_px=4)
    kp = Keypoints(coordinates=np.random.default_rng(0).uniform(0, 100, (10, 2)))
    cal = Cal3Bundler(1000.0, 0.0, 0.0, 640.0, 480.0)
    few = np.stack([np.arange(5), np.arange(5)], 1).astype(np.uint32)
    R, t, idx, ratio = v.verify(kp, kp, few, cal, cal)                               # < 6 matches: failure tuple, no GPU
    assert R is None and t is None and idx.size == 0 and ratio == 0.0
    R, t, idx, ratio = v.verify(kp, kp, np.array([]), cal, cal)
    assert R is None and idx.size == 0
    # fundamental-matrix mode: fewer than 8 matches -> failure tuple (opencv_verifier_base.py:38-40,69), no GPU needed
    vf = Ransac(use_intrinsics_in_verification=False, estimation_threshold_px=4)
    seven = np.stack([np.arange(7), np.arange(7)], 1).astype(np.uint32)
    R, t, idx, ratio = vf.verify(kp, kp, seven, cal, cal)
    assert R is None and t is None and idx.size == 0 and ratio == 0.0
    assert vf._params().estimate_fundamental == 1 and v._params().estimate_fundamental == 0


def test_plugins_are_picklable_like_the_reference_requires():
    m = pickle.loads(pickle.dumps(TwoWayMatcher(ratio_test_threshold=0.7)))
    assert m._ratio_test_threshold == 0.7
    v = pickle.loads(pickle.dumps(Ransac(True, 3.0, num_hypotheses=77, seed=5)))
    assert v._num_hypotheses == 77 and v._seed == 5 and v._estimation_threshold_px == 3.0
    assert TwoWayMatcher.get_ui_metadata().display_name and Ransac.get_ui_metadata().display_name


# ---------------------------------------------------------------------------------------------------- result containers
def _fake_result(pairs, seed):
    rng = np.random.default_rng(seed)
    P = len(pairs)
    counts = rng.integers(0, 20, P)
    off = np.zeros(P + 1, np.int64); np.cumsum(counts, out=off[1:])
    matches = rng.integers(0, 100, (off[-1], 2)).astype(np.uint32)
    mask = rng.integers(0, 2, off[-1]).astype(np.uint8)
    status = (counts < 6).astype(np.int32)
    R = np.tile(np.eye(3), (P, 1, 1)) * rng.uniform(0.5, 1.5, (P, 1, 1))
    t = rng.normal(size=(P, 3))
    ninl = np.array([mask[off[p]:off[p + 1]].sum() for p in range(P)], np.int32)
    return TwoViewBatchResult(np.asarray(pairs, np.int32), off, matches, mask, R, t, ninl, status)


def test_two_view_dicts_follow_the_reference_conventions():
    pairs = synthetic.make_pairs(6, "exhaustive")
    res = _fake_result(pairs, 1)
    Rd, td, vd, rd = res.two_view_dicts()
    put = res.putative_corr_idxs_dict()
    assert set(Rd) == {(int(a), int(b)) for a, b in pairs} == set(put)
    for p, (a, b) in enumerate(pairs):
        key = (int(a), int(b))
        sl = slice(res.match_offsets[p], res.match_offsets[p + 1])
        assert (put[key] == res.matches[sl]).all()
        if res.status[p] != 0:
            assert Rd[key] is None and td[key] is None and vd[key].size == 0 and rd[key] == 0.0
        else:
            assert (vd[key] == res.matches[sl][res.inlier_mask[sl] == 1]).all()
            assert np.allclose(Rd[key].matrix(), res.R[p])


def test_result_blocks_round_trip_without_a_process_group():
    """[headers | payload] blocks: pack on the host mirror, pad like an uneven all-gather, unpack."""
    pairs = synthetic.make_pairs(7, "exhaustive")
    res = _fake_result(pairs, 2)
    hdr, payload = pack_block(res, np.arange(len(pairs)))
    assert hdr.dtype.itemsize == HEADER_BYTES == 128 and payload.dtype == np.uint32
    ok = res.status == 0
    assert (hdr["n_matches"] == np.diff(res.match_offsets)).all() and (hdr["status"] == res.status).all()
    assert (hdr["n_inliers"][ok] == res.n_inliers[ok]).all() and (hdr["n_inliers"][~ok] == 0).all()
    assert (hdr["R"][ok] == res.R.reshape(-1, 9)[ok]).all() and (hdr["t"][ok] == res.t[ok]).all()
    assert len(payload) == hdr["n_inliers"].sum()
    # two "ranks" with different sizes: the second block is padded to the first one's layout
    half = len(pairs) // 3
    parts = [_fake_result(pairs[:half], 5), _fake_result(pairs[half:], 6)]
    blocks = [pack_block(parts[0], np.arange(half)), pack_block(parts[1], np.arange(half, len(pairs)))]
    counts = np.array([[len(h), len(p)] for h, p in blocks])
    off, size = block_layout(counts[:, 0].max(), counts[:, 1].max())
    buf = np.stack([block_bytes(h, p, counts[:, 0].max(), counts[:, 1].max()) for h, p in blocks])
    assert buf.shape == (2, size) and off == 128 * counts[:, 0].max()
    table = unpack_blocks(buf, counts, len(pairs))
    assert (table.headers["pair_id"] == np.arange(len(pairs))).all()
    for r, (h, p) in enumerate(blocks):
        o = np.concatenate([[0], np.cumsum(h["n_inliers"])])
        for k, g in enumerate(h["pair_id"]):
            assert (table.v_corr_idxs(int(g)) == p[o[k]:o[k + 1]]).all()
    # the single-process exchange is the identity
    ex = ResultExchange(None)
    t1 = ex.finish(ex.start_host(res, np.arange(len(pairs))), len(pairs))
    assert (t1.headers == hdr).all()


def test_inlier_support_processor_decisions():
    """inlier_support_processor.py:39-95: ratio test first, then the count test for valid models."""
    from gtsfm_b200.common.two_view_estimation_report import TwoViewEstimationReport, generate_two_view_report
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    isp = InlierSupportProcessor(min_num_inliers_est_model=15, min_inlier_ratio_est_model=0.1)
    v = np.zeros((20, 2), np.uint32)
    ok = isp.run_inlier_support("R", "U", v, TwoViewEstimationReport(v_corr_idxs=v, num_inliers_est_model=20, inlier_ratio_est_model=0.5))
    assert ok[0] == "R" and ok[2] is v
    low_ratio = isp.run_inlier_support("R", "U", v, TwoViewEstimationReport(v_corr_idxs=v, num_inliers_est_model=20, inlier_ratio_est_model=0.05))
    assert low_ratio[0] is None and low_ratio[2].size == 0 and low_ratio[3].num_inliers_est_model == 0
    few = isp.run_inlier_support("R", "U", v[:10], TwoViewEstimationReport(v_corr_idxs=v[:10], num_inliers_est_model=10, inlier_ratio_est_model=0.5))
    assert few[0] is None
    rep = generate_two_view_report(0.5, v, 1.0, 2.0, np.array([True] * 15 + [False] * 5), np.r_[np.full(15, 0.25), np.full(4, 9.0), np.nan])
    assert rep.num_inliers_gt_model == 15 and rep.inlier_ratio_gt_model == 0.75
    assert rep.inlier_avg_reproj_error_gt_model == 0.25 and rep.outlier_avg_reproj_error_gt_model == 9.0


def test_report_oracle_matches_the_pinned_sampson_restatement():
    """The numpy restatement of compute_epipolar_distances_sq_sampson used by the report oracle equals the C oracle's
    sampson_sq (itself pinned against the reference's known answers, tests/test_oracle_verify.py)."""
    from oracle import oracle as O
    rng = np.random.default_rng(2)
    F = rng.standard_normal((3, 3))
    c1, c2 = rng.uniform(0, 1000, (50, 2)), rng.uniform(0, 1000, (50, 2))
    np.testing.assert_allclose(O.sampson_sq_numpy(c1, c2, F), O.sampson_sq(F, c1, c2), rtol=1e-10)


def test_plugin_contexts_cap_their_workspace(monkeypatch):
    """default_context() (one per worker thread) sets a 4 GiB work-buffer budget instead of the library's 12 GiB;
    GTSFM_B200_WORKSPACE_GIB overrides it.  No GPU: the Context class is replaced by a recorder."""
    from gtsfm_b200 import _native
    seen = []

    class FakeCtx:
        def __init__(self, device):
            self.device = device

        def set_workspace(self, nbytes):
            seen.append(nbytes)

    monkeypatch.setattr(_native, "Context", FakeCtx)
    monkeypatch.setattr(_native, "_default_tls", __import__("threading").local())
    _native.default_context(0)
    monkeypatch.setenv("GTSFM_B200_WORKSPACE_GIB", "1.5")
    _native.default_context(1)
    assert seen == [4 << 30, int(1.5 * (1 << 30))]
    assert _native.default_context(0) is _native.default_context(0) and len(seen) == 2
