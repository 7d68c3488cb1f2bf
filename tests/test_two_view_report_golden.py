"""The post-verification report stage pinned to the UNMODIFIED reference:
tests/golden/two_view_report.npz holds inputs and outputs of the reference's compute_correspondence_metrics /
epipolar_inlier_correspondences (utils/metrics.py:38-128), compute_relative_pose_metrics + generate_two_view_report
(two_view_estimator.py:354-422) and InlierSupportProcessor.run_inlier_support (inlier_support_processor.py:39-95), recorded by
oracle/gen_golden_report.py.  CPU: the oracle's numpy restatement and the host mirror classes reproduce them.  GPU: the
device report (gtsfm_two_view_report_batch through TwoViewEstimator) equals the restatement, incl. GT cameras that carry
their own calibration.
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "two_view_report.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLDEN)


def _case(g, c):
    p = f"c{c}_"
    return {k[len(p):]: g[k] for k in g.files if k.startswith(p)}


def _nan_equal(a, b, rtol=1e-12):
    a, b = float(a), float(b)
    return (np.isnan(a) and np.isnan(b)) or abs(a - b) <= rtol * max(1.0, abs(b))


@pytest.mark.parametrize("c", range(8))
def test_oracle_report_restatement_reproduces_the_reference(gold, c):
    from oracle import oracle as O
    k = _case(gold, c)
    verified = bool(k["verified"])
    v = k["v"] if verified else np.zeros((0, 2), np.uint32)
    min_inl, min_ratio = int(k["isp"][0]), float(k["isp"][1])
    rep = O.two_view_report(k["k1"], k["k2"], v, int(k["n_put"]), k["R_est"] if verified else None, k["t_est"] if verified else None,
                            k["gcal1"], k["gcal2"], k["gt_R"], k["gt_t"], 4.0, min_inl, min_ratio)
    assert rep["keep"] == bool(k["isp_keep"])
    if not verified:
        assert rep["R_error_deg"] is None and np.isnan(k["R_err"]) and not bool(k["has_mask"])
        return
    # arccos near 1 amplifies the last bit of the dot product to ~1e-6 degree (case 1: identical directions)
    assert abs(rep["R_error_deg"] - float(k["R_err"])) < 1e-9 and abs(rep["U_error_deg"] - float(k["U_err"])) < 1e-5
    np.testing.assert_allclose(rep["dist_sq"], k["err_gt"], rtol=1e-12)
    np.testing.assert_array_equal(rep["gt_mask"], k["mask_gt"])
    assert rep["num_inliers_gt"] == int(k["rep_num_inliers_gt"])
    assert _nan_equal(rep["inlier_avg"], k["rep_inlier_avg"]) and _nan_equal(rep["outlier_avg"], k["rep_outlier_avg"])


@pytest.mark.parametrize("c", range(8))
def test_host_mirror_classes_reproduce_the_reference(gold, c):
    """gtsfm_b200's generate_two_view_report and InlierSupportProcessor on the reference's own intermediate values."""
    from gtsfm_b200.common.two_view_estimation_report import generate_two_view_report
    from gtsfm_b200.common.types import Rot3, Unit3
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    k = _case(gold, c)
    verified = bool(k["verified"])
    v = k["v"] if verified else np.array([], dtype=np.uint64)
    ratio = len(k["v"]) / int(k["n_put"]) if verified else 0.0
    has = bool(k["has_mask"])
    rep = generate_two_view_report(ratio, v, R_error_deg=None if np.isnan(k["R_err"]) else float(k["R_err"]),
                                   U_error_deg=None if np.isnan(k["U_err"]) else float(k["U_err"]),
                                   v_corr_idxs_inlier_mask_gt=k["mask_gt"] if has else None, reproj_error_gt_model=k["err_gt"] if has else None)
    assert rep.num_inliers_est_model == int(k["rep_num_inliers_est"]) and _nan_equal(rep.inlier_ratio_est_model, k["rep_ratio_est"])
    assert rep.num_inliers_gt_model == int(k["rep_num_inliers_gt"]) and _nan_equal(rep.inlier_ratio_gt_model, k["rep_ratio_gt"])
    assert _nan_equal(rep.inlier_avg_reproj_error_gt_model, k["rep_inlier_avg"])
    assert _nan_equal(rep.outlier_avg_reproj_error_gt_model, k["rep_outlier_avg"])
    proc = InlierSupportProcessor(int(k["isp"][0]), float(k["isp"][1]))
    R, U, pv, prep = proc.run_inlier_support(Rot3(k["R_est"]) if verified else None, Unit3(k["t_est"]) if verified else None, v, rep)
    assert (R is not None) == bool(k["isp_keep"]) and len(pv) == int(k["isp_v_len"])
    assert prep.num_inliers_est_model == int(k["isp_rep_num_inliers_est"])


@pytest.mark.gpu
def test_device_report_with_gt_cameras_of_their_own_calibration():
    """The device report through TwoViewEstimator.run_2view, GT cameras whose calibration differs from camera_intrinsics
    (utils/metrics.py:90-97 uses gt_camera.calibration()), against the restatement pinned above."""
    from gtsfm_b200.common.types import Cal3Bundler, Keypoints
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    from gtsfm_b200.two_view_estimator import TwoViewEstimator
    from oracle import oracle as O
    from tests.test_batched_callers import GtCamera
    from tests.test_gpu_verify import make_pixel_scene
    rng = np.random.default_rng(99)
    n = 900
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=0.35)
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    gcal1, gcal2 = (1008.0, 0.0, 0.0, 642.0, 478.0), (995.0, 0.0, 0.0, 639.0, 483.0)
    # camera 1 at the origin, camera 2 = (R, t): wTi2 = (R^T, -R^T t)
    cam1 = GtCamera(np.eye(3), np.zeros(3), Cal3Bundler(*gcal1))
    cam2 = GtCamera(R.T, -R.T @ t, Cal3Bundler(*gcal2))
    est = TwoViewEstimator(Ransac(True, 4.0, num_hypotheses=512, seed=2), InlierSupportProcessor(15, 0.1), False, 3.0)
    Rr, Ur, v, pre, post_ba, post_isp = est.run_2view(Keypoints(kp1), Keypoints(kp2), mi, Cal3Bundler(*cal), Cal3Bundler(*cal), None, cam1, cam2)
    assert Rr is not None and len(v) > 400
    ref = O.two_view_report(kp1, kp2, v, n, Rr.matrix(), Ur.point3(), gcal1, gcal2, R, t, 3.0, 15, 0.1)
    np.testing.assert_allclose(pre.reproj_error_gt_model, ref["dist_sq"], rtol=1e-9, atol=1e-12)   # px^2; cancellation in near-zero numerators
    band = np.abs(ref["dist_sq"] / 9.0 - 1.0) < 1e-9
    assert ((pre.v_corr_idxs_inlier_mask_gt == ref["gt_mask"]) | band).all()
    assert abs(pre.R_error_deg - ref["R_error_deg"]) < 1e-6 and abs(pre.U_error_deg - ref["U_error_deg"]) < 1e-6
    # and it is NOT what the images' own calibration would give
    other = O.two_view_report(kp1, kp2, v, n, Rr.matrix(), Ur.point3(), cal, cal, R, t, 3.0, 15, 0.1)
    assert np.abs(other["dist_sq"] - ref["dist_sq"]).max() > 1e-3
    assert post_isp.num_inliers_est_model == len(v) and post_ba.num_inliers_est_model == len(v)
