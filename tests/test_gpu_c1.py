"""BASELINE config C1 on the device: the 12 Lund-door images (reference SIFT features, tests/golden/c1_lund_door_12.npz,
made by oracle/gen_golden.py c1) through match + verify for all 66 pairs of the ExhaustiveRetriever.

Matches: bit-exact against the matches recorded from the unmodified reference TwoWayMatcher(ratio 0.8).
Verifier: against the CPU oracle on the identical sample sets (mask exact outside the 1e-6 band, R / t within 0.1 deg),
and against the data set's ground-truth poses (the reference's own tolerance style: a few degrees).
"""
import hashlib
import os

import numpy as np
import pytest

from tests.test_gpu_verify import check_against_oracle, dir_angle_deg, params, rot_angle_deg

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
H = 1024        # 1,024 hypotheses for C1-C4


@pytest.fixture(scope="module")
def c1():
    import gtsfm_b200
    g = np.load(os.path.join(GOLDEN, "c1_lund_door_12.npz"))
    ctx = gtsfm_b200.Context(0)
    off = g["img_off"]
    for i in range(12):
        ctx.upload_image(i, g["desc_u8"][off[i]:off[i + 1]].astype(np.float32), g["kps"][off[i]:off[i + 1]], tuple(g["cal"][i]))
    yield ctx, g
    ctx.close()


def test_c1_matches_equal_reference_for_all_66_pairs(c1):
    ctx, g = c1
    pairs = np.ascontiguousarray(g["pairs"])
    m, off = ctx.match_batch(pairs, 0.8)
    assert m.dtype == np.uint32 and len(off) == 67
    np.testing.assert_array_equal(off, g["m_off"])
    np.testing.assert_array_equal(m, g["m_u16"].astype(np.uint32))
    for k in range(66):
        sha = hashlib.sha1(np.ascontiguousarray(m[off[k]:off[k + 1]]).tobytes()).hexdigest()
        assert sha == str(g["m_sha1"][k])


def test_c1_single_pair_plugin_equals_batch(c1):
    """TwoWayMatcher.match on one pair (the reference's call shape) returns the same rows as the 66-pair batch."""
    from gtsfm_b200.common.types import Keypoints
    from gtsfm_b200.frontend.matcher.twoway_matcher import TwoWayMatcher
    _, g = c1
    off, mo = g["img_off"], g["m_off"]
    for k in (0, 37, 65):
        i, j = g["pairs"][k]
        d1 = g["desc_u8"][off[i]:off[i + 1]].astype(np.float32)
        d2 = g["desc_u8"][off[j]:off[j + 1]].astype(np.float32)
        kp = Keypoints(coordinates=np.zeros((0, 2)))
        got = TwoWayMatcher(ratio_test_threshold=0.8).match(kp, kp, d1, d2, (760, 1135, 3), (760, 1135, 3))
        np.testing.assert_array_equal(got, g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32))


def test_c1_verifier_vs_oracle_and_ground_truth(c1):
    import gtsfm_b200
    ctx, g = c1
    pairs = np.ascontiguousarray(g["pairs"])
    m, off = ctx.match_batch(pairs, 0.8)
    keys = np.arange(66, dtype=np.uint64)
    out = ctx.verify_batch(pairs, params(4.0, H, seed=20260001, keys=keys), total=len(m))
    assert (out["status"] == 0).all()
    io = g["img_off"]
    # oracle on a spread of pairs (adjacent, mid-baseline, the widest)
    for k in (0, 1, 12, 30, 47, 65):
        i, j = pairs[k]
        mk = m[off[k]:off[k + 1]]
        samples = gtsfm_b200.ransac_samples(20260001, int(keys[k]), len(mk), H)
        got = dict(status=out["status"][k], mask=out["mask"][off[k]:off[k + 1]], R=out["R"][k], t=out["t"][k], n_inliers=out["n_inliers"][k])
        check_against_oracle(got, g["kps"][io[i]:io[i + 1]].astype(np.float64), g["kps"][io[j]:io[j + 1]].astype(np.float64), mk,
                             tuple(g["cal"][i]), tuple(g["cal"][j]), 4.0, samples)
    # ground truth from data.mat: i2Ri1 = wRi2^T wRi1, i2ti1 = wRi2^T (wti1 - wti2)
    r_err, t_err = [], []
    for k, (i, j) in enumerate(pairs):
        R_gt = g["wRc"][j].T @ g["wRc"][i]
        t_gt = g["wRc"][j].T @ (g["wtc"][i] - g["wtc"][j])
        r_err.append(rot_angle_deg(out["R"][k], R_gt))
        t_err.append(dir_angle_deg(out["t"][k], t_gt))
    r_err, t_err = np.array(r_err), np.array(t_err)
    ratio = out["n_inliers"] / np.diff(off)
    assert np.median(r_err) < 1.0 and np.median(t_err) < 3.0, (np.median(r_err), np.median(t_err))
    assert (r_err < 5.0).mean() > 0.9, np.sort(r_err)[-8:]
    assert np.median(ratio) > 0.5


def test_c1_matcher_cacher_batch(c1, tmp_path):
    """MatcherCacher over the device matcher: one batch for the misses, reference-compatible entries, second pass all hits."""
    from gtsfm_b200.common.types import Keypoints
    from gtsfm_b200.frontend.cacher.matcher_cacher import MatcherCacher
    from gtsfm_b200.frontend.matcher.twoway_matcher import TwoWayMatcher
    _, g = c1
    off, mo = g["img_off"], g["m_off"]
    kps = [Keypoints(coordinates=g["kps"][off[i]:off[i + 1]]) for i in range(12)]
    descs = [g["desc_u8"][off[i]:off[i + 1]].astype(np.float32) for i in range(12)]
    shapes = [(1135, 760, 3)] * 12
    sel = [0, 5, 11, 40, 65]
    pairs = g["pairs"][sel]
    cacher = MatcherCacher(TwoWayMatcher(ratio_test_threshold=0.8), cache_root=tmp_path)
    out = cacher.match_batch(kps, descs, shapes, pairs)
    assert (cacher.hits, cacher.misses) == (0, 5)
    for k, m in zip(sel, out):
        np.testing.assert_array_equal(m, g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32))
    assert len(os.listdir(tmp_path / "matcher")) == 5
    assert all(f.startswith("TwoWayMatcher_") and f.endswith(".pbz2") for f in os.listdir(tmp_path / "matcher"))
    out2 = cacher.match_batch(kps, descs, shapes, pairs)
    assert (cacher.hits, cacher.misses) == (5, 5)
    for a, b in zip(out, out2):
        np.testing.assert_array_equal(a, b)
    i, j = pairs[2]
    single = cacher.match(kps[i], kps[j], descs[i], descs[j], shapes[i], shapes[j])
    np.testing.assert_array_equal(single, out[2])
    assert cacher.hits == 6
