"""CPU checks of the numpy restatement of two-view triangulation + BA (oracle/two_view_ba.py, PARITY UNPINNED: gtsam is absent)
and of the configuration mirror the host side takes (TriangulationOptions, point3d_initializer.py:41-92)."""
import os
import sys

import numpy as np
import pytest

from oracle import two_view_ba as OB


def _scene(seed=0, n=120, noise=0.4, outliers=8):
    rng = np.random.default_rng(seed)
    cal1 = (1000.0, 0.01, 0.001, 640.0, 480.0)
    cal2 = (900.0, -0.02, 0.0, 600.0, 400.0)
    R = OB.expmap(np.array([0.05, -0.2, 0.03]))
    t = np.array([1.0, 0.1, 0.2]); t /= np.linalg.norm(t)
    X = np.c_[rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(4, 9, n)]
    u1, _ = OB.project(cal1, X)
    u2, _ = OB.project(cal2, X @ R.T + t)
    u1 += rng.normal(0, noise, u1.shape); u2 += rng.normal(0, noise, u2.shape)
    u2[:outliers, 1] += 25.0                 # across the (nearly horizontal) epipolar lines: a free point cannot absorb it
    return cal1, cal2, R, t, X, u1, u2


def _rot_deg(A, B):
    return float(np.degrees(np.linalg.norm(OB.logmap(A.T @ B))))


def test_triangulation_recovers_noise_free_points_and_applies_the_gates():
    cal1, cal2, R, t, X, _, _ = _scene(noise=0.0, outliers=0)
    u1, _ = OB.project(cal1, X); u2, _ = OB.project(cal2, X @ R.T + t)
    P, ok = OB.triangulate(cal1, cal2, R, t, u1, u2, 100.0)
    assert ok.all()
    np.testing.assert_allclose(P, X, rtol=1e-7, atol=1e-7)
    u2b = u2.copy(); u2b[0] += 400.0                                  # a gross mismatch exceeds the reprojection gate
    _, ok = OB.triangulate(cal1, cal2, R, t, u1, u2b, 100.0)
    assert not ok[0] and ok[1:].all()
    _, ok = OB.triangulate(cal1, cal2, R, t, u1, u2, 100.0, min_angle_deg=60.0)   # no track subtends 60 degrees
    assert not ok.any()


def test_bundle_adjustment_improves_a_perturbed_pose_and_filters_outliers():
    cal1, cal2, R, t, X, u1, u2 = _scene()
    R0 = R @ OB.expmap(np.array([0.004, -0.003, 0.002]))
    t0 = t + np.array([0.01, -0.02, 0.015]); t0 /= np.linalg.norm(t0)
    out = OB.bundle_adjust(cal1, cal2, R0, t0, u1, u2)
    assert out["status"] == 0 and out["iterations"] > 0 and out["triangulated"].all()
    assert _rot_deg(out["R"], R) < 0.5 * _rot_deg(R0, R)
    assert np.degrees(np.arccos(np.clip(out["t"] @ t, -1, 1))) < 0.5 * np.degrees(np.arccos(t0 @ t))
    assert abs(np.linalg.norm(out["t"]) - 1.0) < 1e-12
    assert not out["valid"][:8].any() and out["valid"][8:].mean() > 0.5      # 25-px outliers fail the 0.5 px filter


def test_bundle_adjustment_exits():
    cal1, cal2, R, t, X, u1, u2 = _scene(n=40)
    out = OB.bundle_adjust(cal1, cal2, R, t, u1, u2, tri_thr=1e-9)
    assert out["status"] == 2 and not out["valid"].any() and out["iterations"] == 0
    np.testing.assert_array_equal(out["R"], R)
    out = OB.bundle_adjust(cal1, cal2, R, t, u1, u2, ba_thr=1e-9)
    assert out["status"] == 3 and not out["valid"].any()
    np.testing.assert_array_equal(out["R"], R)                        # the initial pose comes back (two_view_estimator.py:199-202)
    out = OB.bundle_adjust(cal1, cal2, R, t, u1, u2, ba_thr=None)
    assert out["valid"].all()                                         # a None threshold keeps every track (bundle_adjustment.py:350-352)


def test_triangulation_options_mirror_the_reference_fields():
    from gtsfm_b200.data_association.point3d_initializer import TriangulationOptions, TriangulationSamplingMode
    o = TriangulationOptions(mode=TriangulationSamplingMode.NO_RANSAC, reproj_error_threshold=100.0)
    assert o._fields == ("mode", "reproj_error_threshold", "min_triangulation_angle", "min_inlier_ratio", "confidence",
                         "dyn_num_hypotheses_multiplier", "min_num_hypotheses", "max_num_hypotheses")
    assert o.min_triangulation_angle == 0.0 and o.confidence == 0.9999 and o.max_num_hypotheses == sys.maxsize
    assert TriangulationOptions(TriangulationSamplingMode("NO_RANSAC")).reproj_error_threshold == np.inf
    assert [m.value for m in TriangulationSamplingMode] == ["NO_RANSAC", "RANSAC_SAMPLE_UNIFORM", "RANSAC_SAMPLE_BIASED_BASELINE",
                                                            "RANSAC_TOPK_BASELINES"]
