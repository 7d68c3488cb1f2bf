"""Golden vectors for the post-verification report stage, recorded from the UNMODIFIED reference
functions imported from /root/reference.  ORACLE-side test infrastructure; run in the build container:

    python oracle/gen_golden_report.py        -> tests/golden/two_view_report.npz

Reference code exercised, as is:
  gtsfm.utils.metrics.compute_correspondence_metrics -> epipolar_inlier_correspondences        utils/metrics.py:38-128
  gtsfm.utils.verification.essential_to_fundamental_matrix / compute_epipolar_distances_sq_sampson   utils/verification.py:113-214
  gtsfm.two_view_estimator.compute_relative_pose_metrics / generate_two_view_report             two_view_estimator.py:354-422
  gtsfm.utils.geometry_comparisons.compute_relative_rotation_angle / ..._unit_translation_angle utils/geometry_comparisons.py:266-309
  gtsfm.frontend.inlier_support_processor.InlierSupportProcessor.run_inlier_support             frontend/inlier_support_processor.py:39-95

Modules these files import that are absent from this image and that the exercised functions never touch (dask, matplotlib,
seaborn, trimesh, ...) are MagicMock placeholders - in this generator only.  `gtsam` (un-vendored, gtsam==4.2) is
oracle/gtsam_stub: Pose3.between, Rot3.between, Unit3, EssentialMatrix.matrix = [t]x R and Cal3Bundler.K restated in numpy.
"""
import importlib
import os
import re
import sys
from unittest.mock import MagicMock

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(HERE, "gtsam_stub"))


def import_with_placeholders(name):
    stubbed = []
    for _ in range(400):
        try:
            return importlib.import_module(name), stubbed
        except ModuleNotFoundError as e:
            sys.modules[e.name] = MagicMock()
            stubbed.append(e.name)
        except ImportError as e:
            # a gtsam name the stub does not define (imported by the reference module, never used by the exercised functions)
            m = re.search(r"cannot import name '(\w+)' from '([\w.]+)'", str(e))
            if not m or m.group(2) != "gtsam":
                raise
            setattr(sys.modules["gtsam"], m.group(1), MagicMock())
            stubbed.append("gtsam." + m.group(1))
    raise RuntimeError("too many missing modules")


def rodrigues(w):
    th = np.linalg.norm(w)
    if th == 0:
        return np.eye(3)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K


def main():
    import gtsam
    missing_gtsam = []

    def _gtsam_placeholder(name):
        # names the reference modules import / annotate with but the exercised functions never use (PEP 562 module __getattr__)
        if name.startswith("__"):
            raise AttributeError(name)
        missing_gtsam.append(name)
        return MagicMock()
    gtsam.__getattr__ = _gtsam_placeholder
    metrics, s1 = import_with_placeholders("gtsfm.utils.metrics")
    tve, s2 = import_with_placeholders("gtsfm.two_view_estimator")
    isp_mod, s3 = import_with_placeholders("gtsfm.frontend.inlier_support_processor")
    from gtsfm.common.keypoints import Keypoints
    print("placeholders:", sorted(set(s1 + s2 + s3)), "gtsam names:", sorted(set(missing_gtsam)))
    rng = np.random.default_rng(20260707)
    out = {}
    # (n keypoints, fraction of verified rows that are wrong w.r.t. GT, pose perturbation deg, verified?, putative count factor,
    #  ISP thresholds, different GT calibration?)
    cases = [
        (400, 0.10, 0.3, True, 1.6, (15, 0.1), False),
        (400, 0.00, 0.0, True, 1.2, (15, 0.1), False),          # every verified row correct: outlier average is NaN
        (300, 1.00, 5.0, True, 1.5, (15, 0.1), False),          # every verified row wrong: inlier average is NaN
        (12, 0.20, 0.5, True, 1.5, (15, 0.1), False),           # fewer inliers than min_num_inliers -> dropped
        (200, 0.10, 0.5, True, 25.0, (15, 0.1), False),         # inlier ratio below the minimum -> dropped
        (100, 0.10, 0.5, False, 2.0, (15, 0.1), False),         # verification failed: (None, None, [], 0.0)
        (500, 0.30, 1.0, True, 1.3, (50, 0.5), True),           # GT cameras with their own calibration
        (60, 0.50, 2.0, True, 1.1, (0, 0.0), False),
    ]
    for c, (n, wrong, pert, verified, putf, (min_inl, min_ratio), own_cal) in enumerate(cases):
        cal1 = gtsam.Cal3Bundler(1000.0 + 10 * c, 0.0, 0.0, 640.0, 480.0)
        cal2 = gtsam.Cal3Bundler(900.0 - 5 * c, 0.0, 0.0, 620.0, 470.0)
        gcal1 = gtsam.Cal3Bundler(1010.0, 0.0, 0.0, 642.0, 481.0) if own_cal else cal1
        gcal2 = gtsam.Cal3Bundler(905.0, 0.0, 0.0, 618.0, 473.0) if own_cal else cal2
        wR1, wt1 = rodrigues(rng.normal(0, 0.3, 3)), rng.normal(0, 1, 3)
        wR2, wt2 = rodrigues(rng.normal(0, 0.3, 3)), rng.normal(0, 1, 3) + [2.0, 0, 0]
        cam1 = gtsam.PinholeCameraCal3Bundler(gtsam.Pose3(gtsam.Rot3(wR1), wt1), gcal1)
        cam2 = gtsam.PinholeCameraCal3Bundler(gtsam.Pose3(gtsam.Rot3(wR2), wt2), gcal2)
        X = rng.uniform(-2, 2, (n, 3)) + [0, 0, 8]
        uv1 = np.array([cam1.project(x) for x in X]) + rng.normal(0, 0.7, (n, 2))
        uv2 = np.array([cam2.project(x) for x in X]) + rng.normal(0, 0.7, (n, 2))
        nw = int(round(wrong * n))
        uv2[:nw] = rng.uniform(0, 1, (nw, 2)) * [1280, 960]
        # keypoint arrays are larger than the verified set and shuffled; v_corr_idxs index into them
        n1, n2 = n + 17, n + 5
        p1, p2 = rng.permutation(n1)[:n], rng.permutation(n2)[:n]
        k1 = rng.uniform(0, 1000, (n1, 2)); k1[p1] = uv1
        k2 = rng.uniform(0, 1000, (n2, 2)); k2[p2] = uv2
        v = np.stack([p1, p2], 1).astype(np.uint32)
        n_put = int(round(putf * n))
        i2Ti1 = cam2.pose().between(cam1.pose())
        R_est = rodrigues(np.radians(pert) * np.array([0.6, -0.5, 0.62])) @ i2Ti1.rotation().matrix()
        t_est = rodrigues(np.radians(2 * pert) * np.array([0.1, 0.9, -0.4])) @ (i2Ti1.translation() / np.linalg.norm(i2Ti1.translation()))
        if verified:
            i2Ri1, i2Ui1, v_in, ratio = gtsam.Rot3(R_est), gtsam.Unit3(t_est), v, len(v) / n_put
        else:
            i2Ri1, i2Ui1, v_in, ratio = None, None, np.array([], dtype=np.uint64), 0.0
        # ---- the reference's sequence (two_view_estimator.py:131-160 then :340-347)
        R_err, U_err = tve.compute_relative_pose_metrics(i2Ri1, i2Ui1, cam1.pose(), cam2.pose())
        mask_gt, err_gt = metrics.compute_correspondence_metrics(keypoints_i1=Keypoints(k1), keypoints_i2=Keypoints(k2), corr_idxs_i1i2=v_in,
                                                                 dist_threshold=4.0, gt_camera_i1=cam1, gt_camera_i2=cam2, gt_scene_mesh=None)
        with np.errstate(all="ignore"):
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                rep = tve.generate_two_view_report(ratio, v_in, R_error_deg=R_err, U_error_deg=U_err, v_corr_idxs_inlier_mask_gt=mask_gt,
                                                   reproj_error_gt_model=err_gt)
        proc = isp_mod.InlierSupportProcessor(min_num_inliers_est_model=min_inl, min_inlier_ratio_est_model=min_ratio)
        pR, pU, pv, prep = proc.run_inlier_support(i2Ri1, i2Ui1, v_in, rep)
        p = f"c{c}_"
        out[p + "k1"], out[p + "k2"], out[p + "v"], out[p + "n_put"] = k1, k2, v, n_put
        out[p + "verified"] = verified
        out[p + "R_est"], out[p + "t_est"] = R_est, t_est
        out[p + "cal1"] = np.array([cal1.fx(), 0, 0, cal1.px(), cal1.py()]); out[p + "cal2"] = np.array([cal2.fx(), 0, 0, cal2.px(), cal2.py()])
        out[p + "gcal1"] = np.array([gcal1.fx(), 0, 0, gcal1.px(), gcal1.py()]); out[p + "gcal2"] = np.array([gcal2.fx(), 0, 0, gcal2.px(), gcal2.py()])
        out[p + "wR1"], out[p + "wt1"], out[p + "wR2"], out[p + "wt2"] = wR1, wt1, wR2, wt2
        out[p + "gt_R"], out[p + "gt_t"] = i2Ti1.rotation().matrix(), i2Ti1.translation()
        out[p + "isp"] = np.array([min_inl, min_ratio])
        # outputs
        out[p + "R_err"] = np.nan if R_err is None else R_err
        out[p + "U_err"] = np.nan if U_err is None else U_err
        out[p + "has_mask"] = mask_gt is not None
        out[p + "mask_gt"] = np.zeros(0, bool) if mask_gt is None else mask_gt
        out[p + "err_gt"] = np.zeros(0) if err_gt is None else err_gt
        out[p + "rep_num_inliers_est"] = rep.num_inliers_est_model
        out[p + "rep_ratio_est"] = rep.inlier_ratio_est_model
        out[p + "rep_num_inliers_gt"] = rep.num_inliers_gt_model
        out[p + "rep_ratio_gt"] = rep.inlier_ratio_gt_model
        out[p + "rep_inlier_avg"] = rep.inlier_avg_reproj_error_gt_model
        out[p + "rep_outlier_avg"] = rep.outlier_avg_reproj_error_gt_model
        out[p + "isp_keep"] = pR is not None
        out[p + "isp_v_len"] = len(pv)
        out[p + "isp_rep_num_inliers_est"] = prep.num_inliers_est_model
        print(c, "R_err", R_err, "U_err", U_err, "gt inliers", rep.num_inliers_gt_model, "/", rep.num_inliers_est_model, "keep", pR is not None,
              "avg", rep.inlier_avg_reproj_error_gt_model, rep.outlier_avg_reproj_error_gt_model)
    out["n_cases"] = len(cases)
    np.savez_compressed(os.path.join(OUT, "two_view_report.npz"), **out)
    print("wrote", os.path.join(OUT, "two_view_report.npz"))


if __name__ == "__main__":
    main()
