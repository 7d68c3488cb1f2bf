"""Two-view triangulation + bundle adjustment on the device (gtsfm_two_view_ba_batch) against the
numpy restatement of the reference's formulation (oracle/two_view_ba.py, PARITY UNPINNED: gtsam is not available here).

The two sides optimise the same robust reprojection cost with the same LM schedule on different parametrisations (kernel:
gauge-fixed 5-DOF relative pose, analytic Jacobians, Schur complement; oracle: two 6-DOF poses held by the reference's soft
priors, finite differences), so the comparison is at convergence: relative rotation and translation direction within 0.05
degrees, valid-track masks equal outside a band around the 0.5 px filter threshold.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_DEG = 0.05


def _angle_deg(A, B):
    c = np.clip(0.5 * (np.trace(A.T @ B) - 1.0), -1, 1)
    return float(np.degrees(np.arccos(c)))


def _dir_deg(a, b):
    return float(np.degrees(np.arccos(np.clip(a @ b / (np.linalg.norm(a) * np.linalg.norm(b)), -1, 1))))


@pytest.fixture(scope="module")
def c1():
    from gtsfm_b200.common.types import Cal3Bundler, Keypoints
    g = np.load(os.path.join(GOLDEN, "c1_lund_door_12.npz"))
    off = g["img_off"]
    kps = [Keypoints(coordinates=g["kps"][off[i]:off[i + 1]].astype(np.float64)) for i in range(12)]
    cals = [Cal3Bundler(*[float(v) for v in g["cal"][i]]) for i in range(12)]
    pairs = [(int(a), int(b)) for a, b in g["pairs"]]
    mo = g["m_off"]
    putative = {key: g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32) for k, key in enumerate(pairs)}
    return g, kps, cals, pairs, putative


def _verify(c1, sel):
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    g, kps, cals, pairs, putative = c1
    keys = [pairs[k] for k in sel]
    ver = Ransac(True, 4.0, num_hypotheses=512, seed=3)
    # every 3rd putative match: a few hundred tracks per pair keep the dense numpy restatement at seconds per pair
    raw, off, ms, ctx = ver.verify_batch_raw(kps, cals, np.array(keys, np.int32), [putative[k][::3] for k in keys])
    return keys, raw, off, ms, ctx


def test_ba_batch_matches_the_restatement_at_convergence(c1):
    from oracle import two_view_ba as OB
    g, kps, cals, pairs, putative = c1
    sel = [0, 5, 15, 30, 44]
    keys, raw, off, ms, ctx = _verify(c1, sel)
    P = len(keys)
    ba = ctx.two_view_ba_batch(P, int(off[-1]), 15, 100, 100.0, 0.5, 1.345)
    assert (ba["status"] == 0).all() and (ba["iterations"] > 0).all()
    for p, (i, j) in enumerate(keys):
        mask = raw["mask"][off[p]:off[p + 1]] == 1
        m = ms[p][mask]
        uv1, uv2 = kps[i].coordinates[m[:, 0]], kps[j].coordinates[m[:, 1]]
        ref = OB.bundle_adjust(tuple(g["cal"][i]), tuple(g["cal"][j]), raw["R"][p], raw["t"][p], uv1, uv2)
        assert ref["status"] == 0
        assert abs(np.linalg.norm(ba["t"][p]) - 1.0) < 1e-12
        np.testing.assert_allclose(ba["R"][p] @ ba["R"][p].T, np.eye(3), atol=1e-12)
        assert _angle_deg(ba["R"][p], ref["R"]) < TOL_DEG, (keys[p], _angle_deg(ba["R"][p], ref["R"]))
        assert _dir_deg(ba["t"][p], ref["t"]) < TOL_DEG, (keys[p], _dir_deg(ba["t"][p], ref["t"]))
        got = ba["valid_mask"][off[p]:off[p + 1]]
        assert not got[~mask].any()                                   # only verified correspondences can stay
        assert int(got.sum()) == int(ba["num_valid"][p])
        agree = (got[mask] == 1) == ref["valid"]
        assert agree.mean() > 0.97, (keys[p], agree.mean())           # rows within a few 1e-3 px of the 0.5 px threshold may flip
        # BA does not make the pose worse against the ground truth than the verifier's
        Rg = g["wRc"][j].T @ g["wRc"][i]
        assert _angle_deg(ba["R"][p], Rg) < _angle_deg(raw["R"][p].reshape(3, 3), Rg) + 0.05


def test_ba_skips_pairs_below_the_inlier_floor_and_failed_pairs(c1):
    g, kps, cals, pairs, putative = c1
    keys, raw, off, ms, ctx = _verify(c1, [0, 5])
    big = int(raw["n_inliers"].max()) + 1
    ba = ctx.two_view_ba_batch(2, int(off[-1]), big, 100, 100.0, 0.5, 1.345)      # floor above every pair: nothing runs
    assert (ba["status"] == 1).all() and (ba["iterations"] == 0).all()
    np.testing.assert_array_equal(ba["R"].reshape(2, 9), raw["R"].reshape(2, 9))
    np.testing.assert_array_equal(ba["t"], raw["t"])
    np.testing.assert_array_equal(ba["valid_mask"], raw["mask"][: int(off[-1])])
    np.testing.assert_array_equal(ba["num_valid"], raw["n_inliers"])


def test_ba_with_an_unreachable_filter_threshold_returns_the_verifier_pose(c1):
    g, kps, cals, pairs, putative = c1
    keys, raw, off, ms, ctx = _verify(c1, [0])
    ba = ctx.two_view_ba_batch(1, int(off[-1]), 15, 100, 100.0, 1e-9, 1.345)      # no track can pass a 1e-9 px filter
    assert ba["status"][0] == 3 and ba["num_valid"][0] == 0 and not ba["valid_mask"].any()
    np.testing.assert_array_equal(ba["R"].reshape(9), raw["R"].reshape(9))
    ba = ctx.two_view_ba_batch(1, int(off[-1]), 15, 100, 1e-9, 0.5, 1.345)        # nothing triangulates under a 1e-9 px threshold
    assert ba["status"][0] == 2 and ba["num_valid"][0] == 0 and ba["iterations"][0] == 0


def test_ba_needs_the_last_verify_batch_and_sane_parameters(c1):
    from gtsfm_b200 import _native
    keys, raw, off, ms, ctx = _verify(c1, [0, 5])
    with pytest.raises(_native.GtsfmB200Error):
        ctx.two_view_ba_batch(3, int(off[-1]), 15)
    with pytest.raises(_native.GtsfmB200Error):
        ctx.two_view_ba_batch(2, int(off[-1]), 15, ba_reproj_error_thresh=0.0)
    with pytest.raises(_native.GtsfmB200Error):                                         # post-BA report without a BA batch
        ctx.two_view_report_batch(2, int(off[-1]), 4.0, 15, 0.1, after_ba=True)


def test_two_view_estimator_with_bundle_adjustment_builds_the_reference_tuples(c1):
    """TwoViewEstimator(bundle_adjust_2view=True): the post-BA report is taken on BA's outputs with the pre-BA ratio
    (two_view_estimator.py:316-333), the InlierSupportProcessor runs on it, skipped pairs copy the pre-BA report."""
    from tests.test_batched_callers import GtCamera
    from gtsfm_b200.data_association.point3d_initializer import TriangulationOptions, TriangulationSamplingMode
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    from gtsfm_b200.two_view_estimator import TwoViewEstimator
    g, kps, cals, pairs, putative = c1
    cams = [GtCamera(g["wRc"][i], g["wtc"][i], cals[i]) for i in range(12)]
    opts = TriangulationOptions(mode=TriangulationSamplingMode.NO_RANSAC, reproj_error_threshold=100.0)
    mk = lambda ba: TwoViewEstimator(Ransac(True, 4.0, num_hypotheses=512, seed=3), InlierSupportProcessor(15, 0.1), ba, 4.0,   # noqa: E731
                                     triangulation_options=opts)
    sub = {k: putative[k] for k in pairs[:24]}
    sub[pairs[3]] = sub[pairs[3]][:4]                                   # verification fails -> failure tuple, BA skipped
    with_ba = mk(True).run_2view_batch(kps, sub, cals, None, cams)
    without = mk(False).run_2view_batch(kps, sub, cals, None, cams)
    better = 0
    for key in sub:
        R, U, v, pre, post_ba, post_isp = with_ba[key]
        R0, U0, v0, pre0, _, _ = without[key]
        np.testing.assert_array_equal(pre.v_corr_idxs, pre0.v_corr_idxs)              # BA does not touch the pre-BA report
        if key == pairs[3]:
            assert R is None and v.size == 0 and post_ba.num_inliers_est_model == 0
            continue
        if pre.num_inliers_est_model < 15:
            assert post_ba.num_inliers_est_model == pre.num_inliers_est_model
            continue
        assert post_ba.inlier_ratio_est_model == pre.inlier_ratio_est_model           # the reference's "hack"
        assert post_ba.num_inliers_est_model == len(post_ba.v_corr_idxs) <= pre.num_inliers_est_model
        # every post-BA correspondence is one of the verified ones, in order
        pre_rows = {tuple(r) for r in pre.v_corr_idxs.tolist()}
        assert all(tuple(r) in pre_rows for r in post_ba.v_corr_idxs.tolist())
        assert post_ba.R_error_deg is not None and post_ba.v_corr_idxs_inlier_mask_gt.shape[0] == len(post_ba.v_corr_idxs)
        if R is not None:
            assert v.shape == post_ba.v_corr_idxs.shape and post_isp.num_inliers_est_model == len(v)
            assert abs(np.linalg.norm(U.point3()) - 1.0) < 1e-12
        better += post_ba.R_error_deg <= pre.R_error_deg + 1e-3
    assert better >= 0.7 * (len(sub) - 1)                                # refinement helps on this scene (GT poses are SfM output)
    # run_2view: a batch of one through the same code
    key = pairs[5]
    one = mk(True).run_2view(kps[key[0]], kps[key[1]], putative[key], cals[key[0]], cals[key[1]], None, cams[key[0]], cams[key[1]])
    assert one[0] is not None and one[4].num_inliers_est_model == len(one[4].v_corr_idxs)


def test_ba_with_radial_distortion_and_different_calibrations():
    """Cal3Bundler k1, k2 != 0 and K1 != K2 through triangulation (fixed-point undistortion), projection and its Jacobian."""
    import gtsfm_b200
    from oracle import two_view_ba as OB
    from tests.test_two_view_ba_restatement import _scene
    cal1, cal2, R, t, X, u1, u2 = _scene(seed=3, n=260, noise=0.3, outliers=12)
    rng = np.random.default_rng(1)
    perm1, perm2 = rng.permutation(len(u1)), rng.permutation(len(u2))          # keypoint order differs from match order
    k1, k2 = np.empty_like(u1), np.empty_like(u2)
    k1[perm1], k2[perm2] = u1, u2
    matches = np.stack([perm1, perm2], 1).astype(np.uint32)
    ctx = gtsfm_b200.Context(0)
    ctx.upload_image(0, None, k1, cal1)
    ctx.upload_image(1, None, k2, cal2)
    prm = gtsfm_b200.VerifyParams()
    prm.threshold_px, prm.n_hypotheses, prm.lo_iters, prm.seed = 4.0, 512, 20, 7
    v = ctx.verify_batch(np.array([[0, 1]], np.int32), prm, matches, np.array([0, len(matches)], np.int64))
    assert v["status"][0] == 0 and v["n_inliers"][0] >= 200
    ba = ctx.two_view_ba_batch(1, len(matches), 15, 100, 100.0, 0.5, 1.345)
    mask = v["mask"] == 1
    ref = OB.bundle_adjust(cal1, cal2, v["R"][0], v["t"][0], u1[mask], u2[mask])
    assert ba["status"][0] == ref["status"] == 0
    assert _angle_deg(ba["R"][0], ref["R"]) < TOL_DEG and _dir_deg(ba["t"][0], ref["t"]) < TOL_DEG
    assert ((ba["valid_mask"][mask] == 1) == ref["valid"]).mean() > 0.97
    assert _angle_deg(ba["R"][0], R) < 0.1 and _dir_deg(ba["t"][0], t) < 0.5      # and it lands on the scene's pose
    # a minimum triangulation angle above what the scene offers rejects every track (point3d_initializer.py:287-291)
    far = ctx.two_view_ba_batch(1, len(matches), 15, 100, 100.0, 0.5, 1.345, min_triangulation_angle_deg=60.0)
    assert far["status"][0] == 2 and far["num_valid"][0] == 0
    ctx.close()
