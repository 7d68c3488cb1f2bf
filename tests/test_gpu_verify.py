"""GPU parity of the batched Ransac verifier against the CPU oracle, driven by the identical (H, 5) sample sets.

Bars (BASELINE.json north_star): inlier masks bit-exact except for correspondences whose squared Sampson distance is
within 1e-6 (relative) of the threshold; rotation and translation direction within 0.1 degree of the oracle.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BAND = 1e-6
ANGLE_TOL_DEG = 0.1


def rot_angle_deg(R1, R2):
    return np.degrees(np.arccos(np.clip((np.trace(R1.T @ R2) - 1) / 2, -1, 1)))


def dir_angle_deg(a, b):
    return np.degrees(np.arccos(np.clip(a @ b / np.linalg.norm(a) / np.linalg.norm(b), -1, 1)))


@pytest.fixture(scope="module")
def ctx():
    import gtsfm_b200
    c = gtsfm_b200.Context(0)
    yield c
    c.close()


def make_pixel_scene(rng, n, noise_px=0.5, out_frac=0.4, f=1000.0, u0=640.0, v0=480.0):
    import cv2
    X = rng.uniform(-2, 2, (n, 3)) + [0, 0, 6]
    R, _ = cv2.Rodrigues(rng.normal(0, 0.2, 3))
    t = rng.normal(0, 1, 3)
    t /= np.linalg.norm(t)
    x1 = X[:, :2] / X[:, 2:]
    Xc = X @ R.T + t
    x2 = Xc[:, :2] / Xc[:, 2:]
    p1 = x1 * f + [u0, v0] + rng.normal(0, noise_px, x1.shape)
    p2 = x2 * f + [u0, v0] + rng.normal(0, noise_px, x2.shape)
    no = int(out_frac * n)
    p2[:no] = rng.uniform(0, 1, (no, 2)) * [2 * u0, 2 * v0]
    perm = rng.permutation(n)
    return p1[perm], p2[perm], R, t, (f, 0.0, 0.0, u0, v0)


def params(thr, H, lo=20, seed=7, samples=None, keys=None):
    import gtsfm_b200
    p = gtsfm_b200.VerifyParams()
    p.threshold_px = thr
    p.n_hypotheses = H
    p.lo_iters = lo
    p.seed = seed
    p.pair_keys = None if keys is None else keys.ctypes.data
    p.samples = None if samples is None else samples.ctypes.data
    return p


def check_against_oracle(got, kp1, kp2, mi, cal1, cal2, thr, samples, lo=20):
    from oracle import oracle as O
    ref = O.verify_pair(kp1, kp2, mi, cal1, cal2, thr, samples, lo_iters=lo)
    assert got["status"] == ref["status"]
    if ref["status"] != 0:
        return ref
    n1 = O.calibrate(kp1[mi[:, 0]], cal1)
    n2 = O.calibrate(kp2[mi[:, 1]], cal2)
    thr2 = (thr / max(cal1[0], cal2[0])) ** 2
    s = O.sampson_sq(ref["E"], n1, n2)
    differ = got["mask"] != ref["mask"]
    in_band = np.abs(s / thr2 - 1.0) < BAND
    assert not (differ & ~in_band).any(), f"{(differ & ~in_band).sum()} mask entries differ outside the 1e-6 band"
    assert abs(int(got["n_inliers"]) - int(ref["ninl"])) <= int(in_band.sum())
    assert rot_angle_deg(got["R"], ref["R"]) < ANGLE_TOL_DEG
    assert dir_angle_deg(got["t"], ref["t"]) < ANGLE_TOL_DEG
    return ref


@pytest.mark.parametrize("n,out_frac,H", [(500, 0.4, 256), (3000, 0.5, 512), (64, 0.2, 128), (6, 0.0, 32)])
def test_verify_pair_vs_oracle_same_samples(ctx, n, out_frac, H):
    import gtsfm_b200
    rng = np.random.default_rng(n)
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=out_frac)
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    samples = gtsfm_b200.ransac_samples(7, 0, n, H)
    # explicit samples and the device sampler must give the same result
    g1 = ctx.verify_pair(kp1, kp2, mi, cal, cal, params(4.0, H, samples=samples))
    g2 = ctx.verify_pair(kp1, kp2, mi, cal, cal, params(4.0, H))
    np.testing.assert_array_equal(g1["mask"], g2["mask"])
    np.testing.assert_array_equal(g1["R"], g2["R"])
    ref = check_against_oracle(g1, kp1, kp2, mi, cal, cal, 4.0, samples)
    if n >= 64:
        assert rot_angle_deg(g1["R"], R) < 1.0 and dir_angle_deg(g1["t"], t) < 3.0


def test_verify_without_local_optimisation(ctx):
    import gtsfm_b200
    rng = np.random.default_rng(3)
    kp1, kp2, R, t, cal = make_pixel_scene(rng, 800)
    mi = np.c_[np.arange(800), np.arange(800)].astype(np.uint32)
    samples = gtsfm_b200.ransac_samples(1, 0, 800, 200)
    g = ctx.verify_pair(kp1, kp2, mi, cal, cal, params(4.0, 200, lo=0, seed=1))
    check_against_oracle(g, kp1, kp2, mi, cal, cal, 4.0, samples, lo=0)


def test_reference_two_plane_scene(ctx):
    # tests/frontend/verifier/test_verifier_base.py:81-100 -- all 8 matches verified, R/t within 2 degrees of GT
    import gtsfm_b200
    g = np.load(os.path.join(GOLDEN, "verifier_two_planes.npz"))
    mi = np.c_[np.arange(8), np.arange(8)].astype(np.uint32)
    cal = (1.0, 0.0, 0.0, 0.0, 0.0)
    samples = gtsfm_b200.ransac_samples(7, 0, 8, 200)
    got = ctx.verify_pair(g["kp1"], g["kp2"], mi, cal, cal, params(0.5, 200))
    assert got["status"] == 0 and got["mask"].all()
    assert rot_angle_deg(got["R"], g["R_gt"]) < 2 and dir_angle_deg(got["t"], g["t_gt"]) < 2
    check_against_oracle(got, g["kp1"], g["kp2"], mi, cal, cal, 0.5, samples)


def test_reference_argoverse_pair(ctx):
    # tests/frontend/verifier/test_verifier_argoverse.py:73-119 (euler within 1 degree, i1ti2 within 0.01).
    # The 20 annotated points admit two optima with 11 inliers each (MSAC cost 1.314e-6 and 1.337e-6); cv2's
    # USAC_ACCURATE returns the second one.  With sampler seed 1 the best minimal model lies in that basin and the
    # refined pose reproduces the recorded cv2 output (ref_R / ref_t) itself; other seeds reach the lower-cost optimum,
    # which is 0.012 from the annotated translation.  Oracle parity is asserted for both.
    import gtsfm_b200
    from scipy.spatial.transform import Rotation
    g = np.load(os.path.join(GOLDEN, "verifier_argoverse.npz"))
    n = g["kp1"].shape[0]
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    kp1, kp2, cal = g["kp1"].astype(np.float64), g["kp2"].astype(np.float64), tuple(g["cal"])
    got = ctx.verify_pair(kp1, kp2, mi, cal, cal, params(0.5, 1024, seed=1))
    assert got["status"] == 0
    eul = Rotation.from_matrix(got["R"].T).as_euler("zyx", degrees=True)
    assert np.allclose(eul, g["gt_euler_zyx_deg"], atol=1.0), eul
    assert np.allclose(-got["R"].T @ got["t"], g["gt_i1ti2"], atol=0.01)
    assert rot_angle_deg(got["R"], g["ref_R"]) < 1e-3 and dir_angle_deg(got["t"], g["ref_t"]) < 1e-3
    np.testing.assert_array_equal(mi[got["mask"] == 1], g["ref_idx"])
    check_against_oracle(got, kp1, kp2, mi, cal, cal, 0.5, gtsfm_b200.ransac_samples(1, 0, n, 1024))
    got = ctx.verify_pair(kp1, kp2, mi, cal, cal, params(0.5, 1024, seed=0))
    check_against_oracle(got, kp1, kp2, mi, cal, cal, 0.5, gtsfm_b200.ransac_samples(0, 0, n, 1024))
    eul = Rotation.from_matrix(got["R"].T).as_euler("zyx", degrees=True)
    assert np.allclose(eul, g["gt_euler_zyx_deg"], atol=1.0), eul


def test_gates_too_few_matches(ctx):
    # opencv_verifier_base.py:69-78 and tests/frontend/verifier/test_verifier_base.py:120-138
    kp = np.random.default_rng(0).uniform(0, 100, (10, 2))
    cal = (1.0, 0.0, 0.0, 0.0, 0.0)
    for M in (0, 4, 5):
        mi = np.c_[np.arange(M), np.arange(M)].astype(np.uint32).reshape(-1, 2)
        got = ctx.verify_pair(kp, kp, mi, cal, cal, params(0.5, 16))
        assert got["status"] == 1 and got["n_inliers"] == 0 and not got["mask"].any()


def test_batch_of_ragged_pairs(ctx):
    """Several pairs of different sizes in one call, small workspace -> several chunks; equals per-pair results."""
    import gtsfm_b200
    rng = np.random.default_rng(5)
    sizes = [300, 5, 1200, 40, 0, 700]
    scenes = [make_pixel_scene(rng, max(n, 1)) for n in sizes]
    ctx.clear()
    for k, (kp1, kp2, _, _, cal) in enumerate(scenes):
        ctx.upload_image(2 * k, None, kp1, cal)
        ctx.upload_image(2 * k + 1, None, kp2, cal)
    pairs = np.array([(2 * k, 2 * k + 1) for k in range(len(sizes))], np.int32)
    mi = np.concatenate([np.c_[np.arange(n), np.arange(n)] for n in sizes]).astype(np.uint32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    keys = np.arange(100, 100 + len(sizes), dtype=np.uint64)
    ctx.set_workspace(64 << 20)
    out = ctx.verify_batch(pairs, params(4.0, 256, seed=9, keys=keys), mi, off)
    ctx.set_workspace(4 << 30)
    for k, n in enumerate(sizes):
        kp1, kp2, _, _, cal = scenes[k]
        got = dict(status=out["status"][k], mask=out["mask"][off[k]:off[k + 1]], R=out["R"][k], t=out["t"][k], n_inliers=out["n_inliers"][k])
        if n < 6:
            assert got["status"] == 1
            continue
        samples = gtsfm_b200.ransac_samples(9, int(keys[k]), n, 256)
        check_against_oracle(got, kp1, kp2, mi[off[k]:off[k + 1]], cal, cal, 4.0, samples)


def test_match_then_verify_on_device(ctx):
    """The fused path: matches stay on the device between gtsfm_match_batch and gtsfm_verify_batch."""
    import gtsfm_b200
    from tests.test_gpu_match import sift_like
    rng = np.random.default_rng(8)
    n = 1500
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=0.0)
    protos = np.abs(rng.standard_normal((n, 128)))
    d1 = sift_like(rng, n, protos=protos)
    d2 = sift_like(rng, n, protos=protos)
    d2[: n // 3] = sift_like(rng, n // 3)            # a third of image 2 is clutter
    perm = rng.permutation(n)
    d2, kp2 = d2[perm], kp2[perm]
    ctx.clear()
    ctx.upload_image(0, d1, kp1, cal)
    ctx.upload_image(1, d2, kp2, cal)
    pairs = np.array([[0, 1]], np.int32)
    m, off = ctx.match_batch(pairs, 0.8)
    out = ctx.verify_batch(pairs, params(4.0, 512, seed=3), total=len(m))
    assert out["status"][0] == 0
    samples = gtsfm_b200.ransac_samples(3, 0, len(m), 512)
    got = dict(status=out["status"][0], mask=out["mask"], R=out["R"][0], t=out["t"][0], n_inliers=out["n_inliers"][0])
    check_against_oracle(got, kp1, kp2, m, cal, cal, 4.0, samples)
    assert rot_angle_deg(out["R"][0], R) < 1.0
    assert out["n_inliers"][0] > 0.5 * len(m)


def test_async_fetch_and_pinned_outputs_equal_the_plain_path(ctx):
    """gtsfm_match_fetch_async + pinned result buffers (the end-to-end leg of bench.py) return the same arrays as the
    synchronous, caller-owned path, twice in a row (the arena is reused)."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_scene(n_images=4, n_kps=1024, seed=77)
    ctx.clear()
    for i in range(4):
        ctx.upload_image(i, scene.descriptors[i], scene.keypoints[i], scene.cal)
    prm = params(4.0, 128, seed=9)
    m0, off0 = ctx.match_batch(scene.pairs, 0.8)
    out0 = ctx.verify_batch(scene.pairs, prm, total=len(m0))
    ref = {k: np.array(v, copy=True) for k, v in out0.items()}
    for _ in range(2):
        m1, off1 = ctx.match_batch(scene.pairs, 0.8, fetch="async")
        out1 = ctx.verify_batch(scene.pairs, prm, total=int(off1[-1]), reuse_outputs=True)
        ctx.sync()
        np.testing.assert_array_equal(off1, off0)
        np.testing.assert_array_equal(m1, m0)
        for k in ("R", "t", "E", "n_inliers", "status", "mask"):
            np.testing.assert_array_equal(out1[k], ref[k])


def test_low_inlier_rate_many_hypotheses(ctx):
    """C5-shaped pair: 25 % inliers and several thousand hypotheses; same-sample parity with the oracle."""
    import gtsfm_b200
    rng = np.random.default_rng(55)
    n, H = 1200, 4096
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=0.75)
    mi = np.stack([np.arange(n), np.arange(n)], 1).astype(np.uint32)
    samples = gtsfm_b200.ransac_samples(21, 7, n, H)
    prm = params(4.0, H, seed=21)
    keys = np.array([7], np.uint64)
    prm.pair_keys = keys.ctypes.data
    ctx.clear()
    ctx.upload_image(0, None, kp1, cal)
    ctx.upload_image(1, None, kp2, cal)
    out = ctx.verify_batch(np.array([[0, 1]], np.int32), prm, mi, np.array([0, n], np.int64))
    got = dict(status=out["status"][0], mask=out["mask"], R=out["R"][0], t=out["t"][0], n_inliers=out["n_inliers"][0])
    check_against_oracle(got, kp1, kp2, mi, cal, cal, 4.0, samples)
    assert rot_angle_deg(out["R"][0], R) < 1.0
    cand, hyps, evals, executed = ctx.verify_stats(reset=True)
    assert hyps >= H and 0 < executed <= evals


def test_full_size_properties_c2_pair(ctx):
    """Size-independent properties at the BASELINE size (8192 keypoints per image, 1024 hypotheses; the oracle comparison at
    this size is tests/test_gpu_baseline_sizes.py): the inlier mask IS sampson^2 < thr^2 of the returned E (fp64, strict), E = [t]x R
    up to scale, R is a rotation, t a unit vector, the pose is the scene's, and matching is symmetric under swapping the
    images (mutual nearest neighbours with the ratio test off)."""
    from gtsfm_b200 import synthetic
    from oracle import oracle as O
    scene = synthetic.make_scene(n_images=2, n_kps=8192, seed=20260002)
    ctx.clear()
    for i in range(2):
        ctx.upload_image(i, scene.descriptors[i], scene.keypoints[i], scene.cal)
    pairs = np.array([[0, 1]], np.int32)
    m, off = ctx.match_batch(pairs, 0.8)
    out = ctx.verify_batch(pairs, params(4.0, 1024, seed=1), total=len(m))
    assert out["status"][0] == 0 and len(m) > 4000
    R, t, E, mask = out["R"][0], out["t"][0], out["E"][0], out["mask"]
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-9) and abs(np.linalg.det(R) - 1) < 1e-9 and abs(np.linalg.norm(t) - 1) < 1e-9
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    En, Ep = E / np.linalg.norm(E), (tx @ R) / np.linalg.norm(tx @ R)
    assert min(np.abs(En - Ep).max(), np.abs(En + Ep).max()) < 1e-6
    n1 = O.calibrate(scene.keypoints[0][m[:, 0]], scene.cal)
    n2 = O.calibrate(scene.keypoints[1][m[:, 1]], scene.cal)
    s = O.sampson_sq(E, n1, n2) / (4.0 / scene.cal[0]) ** 2
    differ = (s < 1.0) != (mask == 1)
    assert not (differ & (np.abs(s - 1.0) > BAND)).any()
    assert out["n_inliers"][0] == int(mask.sum())
    Rgt, tgt = scene.relative_pose(0, 1)
    assert rot_angle_deg(R, Rgt) < 0.5 and np.degrees(np.arccos(np.clip(t @ tgt, -1, 1))) < 2.0
    # symmetry of mutual nearest neighbours
    m01, _ = ctx.match_batch(pairs, None)
    m10, _ = ctx.match_batch(np.array([[1, 0]], np.int32), None)
    assert {(int(a), int(b)) for a, b in m01} == {(int(b), int(a)) for a, b in m10}
    # every ratio-test match is also a mutual nearest neighbour
    assert {(int(a), int(b)) for a, b in m} <= {(int(a), int(b)) for a, b in m01}


# ------------------------------------------------------------------------------------------ fundamental-matrix mode
def params_f(thr, H, seed=7, samples=None, keys=None):
    p = params(thr, H, lo=0, seed=seed, samples=samples, keys=keys)
    p.estimate_fundamental = 1
    return p


def check_against_oracle_F(got, kp1, kp2, mi, cal1, cal2, thr, samples):
    """Ransac(use_intrinsics_in_verification=False): the inlier rule is cv2's float32 error <= thr^2 on integer-counted
    inliers, so the bar is exact equality of the masks and of the winning model (E to 1e-9 relative); R, t within 0.1 deg."""
    from oracle import oracle as O
    ref = O.verify_pair_F(kp1, kp2, mi, cal1, cal2, thr, samples)
    assert got["status"] == ref["status"]
    if ref["status"] != 0:
        return ref
    np.testing.assert_array_equal(got["mask"], ref["mask"])
    assert int(got["n_inliers"]) == ref["ninl"]
    scale = np.abs(ref["E"]).max()
    assert np.abs(got["E"] - ref["E"]).max() < 1e-9 * scale
    assert rot_angle_deg(got["R"], ref["R"]) < ANGLE_TOL_DEG
    assert dir_angle_deg(got["t"], ref["t"]) < ANGLE_TOL_DEG
    return ref


@pytest.mark.parametrize("n,out_frac,H", [(500, 0.4, 256), (3000, 0.5, 512), (64, 0.2, 128), (8, 0.0, 32), (5000, 0.3, 300)])
def test_fundamental_mode_vs_oracle_same_samples(ctx, n, out_frac, H):
    import gtsfm_b200
    rng = np.random.default_rng(1000 + n)
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=out_frac)
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    samples = gtsfm_b200.ransac_samples(7, 0, n, H, 7)
    assert samples.shape == (H, 7) and all(len(set(r)) == 7 for r in samples.tolist())
    g1 = ctx.verify_pair(kp1, kp2, mi, cal, cal, params_f(4.0, H, samples=samples))
    g2 = ctx.verify_pair(kp1, kp2, mi, cal, cal, params_f(4.0, H))
    np.testing.assert_array_equal(g1["mask"], g2["mask"])
    np.testing.assert_array_equal(g1["R"], g2["R"])
    ref = check_against_oracle_F(g1, kp1, kp2, mi, cal, cal, 4.0, samples)
    if n >= 500:
        assert ref["status"] == 0 and rot_angle_deg(g1["R"], R) < 3.0


def test_fundamental_mode_golden_fixture(ctx):
    """tests/golden/fmat_synth1500.npz (recorded output of the unmodified reference Ransac(False, 4).verify and of cv2,
    oracle/gen_golden.py): oracle parity on the same samples, inlier-set IoU and pose agreement with the reference."""
    import gtsfm_b200
    g = np.load(os.path.join(GOLDEN, "fmat_synth1500.npz"))
    kp1, kp2, mi, cal = g["kp1"].astype(np.float64), g["kp2"].astype(np.float64), g["mi"].astype(np.uint32), tuple(g["cal"])
    M = mi.shape[0]
    samples = gtsfm_b200.ransac_samples(11, 0, M, 1024, 7)
    got = ctx.verify_pair(kp1, kp2, mi, cal, cal, params_f(4.0, 1024, seed=11))
    check_against_oracle_F(got, kp1, kp2, mi, cal, cal, 4.0, samples)
    mine = set(map(tuple, mi[got["mask"] == 1]))
    ref = set(map(tuple, g["ref_idx"]))
    assert len(mine & ref) / len(mine | ref) > 0.95
    # no refit after the best minimal model (as cv2 FM_RANSAC): the translation direction is only good to a few degrees
    assert rot_angle_deg(got["R"], g["R_gt"]) < 1.0 and dir_angle_deg(got["t"], g["t_gt"]) < 5.0


def test_fundamental_mode_gates_and_batch(ctx):
    """< 8 matches -> TOO_FEW (opencv_verifier_base.py:38-40,69); ragged batch over several chunks equals the oracle."""
    import gtsfm_b200
    rng = np.random.default_rng(6)
    sizes = [300, 7, 1200, 40, 0, 9000, 8]
    scenes = [make_pixel_scene(rng, max(n, 1)) for n in sizes]
    ctx.clear()
    for k, (kp1, kp2, _, _, cal) in enumerate(scenes):
        ctx.upload_image(2 * k, None, kp1, cal)
        ctx.upload_image(2 * k + 1, None, kp2, cal)
    pairs = np.array([(2 * k, 2 * k + 1) for k in range(len(sizes))], np.int32)
    mi = np.concatenate([np.c_[np.arange(n), np.arange(n)] for n in sizes]).astype(np.uint32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    keys = np.arange(200, 200 + len(sizes), dtype=np.uint64)
    ctx.set_workspace(64 << 20)
    out = ctx.verify_batch(pairs, params_f(3.0, 200, seed=9, keys=keys), mi, off)
    ctx.set_workspace(4 << 30)
    for k, n in enumerate(sizes):
        kp1, kp2, _, _, cal = scenes[k]
        got = dict(status=out["status"][k], mask=out["mask"][off[k]:off[k + 1]], R=out["R"][k], t=out["t"][k],
                   n_inliers=out["n_inliers"][k], E=out["E"][k])
        if n < 8:
            assert got["status"] == 1 and not got["mask"].any()
            continue
        samples = gtsfm_b200.ransac_samples(9, int(keys[k]), n, 200, 7)
        check_against_oracle_F(got, kp1, kp2, mi[off[k]:off[k + 1]], cal, cal, 3.0, samples)
    cand, hyps, evals, executed = ctx.verify_stats(reset=True)
    assert hyps > 0 and 0 < executed <= evals and hyps <= cand <= 3 * hyps


def test_fundamental_mode_plugin(ctx):
    """Ransac(use_intrinsics_in_verification=False).verify through the plugin class (configs/verifier/ransac_8pt.yaml)."""
    from gtsfm_b200.common.types import Cal3Bundler, Keypoints
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    rng = np.random.default_rng(12)
    n = 1500
    kp1, kp2, R, t, cal = make_pixel_scene(rng, n, out_frac=0.3)
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    v = Ransac(use_intrinsics_in_verification=False, estimation_threshold_px=2.0, num_hypotheses=512, seed=4)
    c = Cal3Bundler(*cal)
    i2Ri1, i2Ui1, v_corr, ratio = v.verify(Keypoints(kp1), Keypoints(kp2), mi, c, c)
    assert i2Ri1 is not None and v_corr.dtype == mi.dtype and 0.5 < ratio <= 0.75
    assert rot_angle_deg(i2Ri1.matrix(), R) < 2.0


def test_fundamental_mode_adaptive_budget_low_inlier_pairs(ctx):
    """cv2.findFundamentalMat(FM_RANSAC, confidence 0.999999, maxIters 1e6) adapts its iteration count to the inlier ratio
    (ransac.py:22-23,103-110).  Pairs at 30 % and 70 % inliers in one batch: with adaptive_max_hypotheses the low-inlier pair
    is re-run with the budget the stopping rule asks for (tens of thousands of 7-point samples), the other one is not;
    each equals the oracle on the sample set of the budget it ended with, and the low-inlier pair recovers as many inliers
    as cv2 itself finds, which 1,024 fixed hypotheses do not."""
    import cv2
    import gtsfm_b200
    rng = np.random.default_rng(41)
    n = 1200
    scenes = [make_pixel_scene(rng, n, out_frac=0.70), make_pixel_scene(rng, n, out_frac=0.30)]
    ctx.clear()
    for k, (kp1, kp2, _, _, cal) in enumerate(scenes):
        ctx.upload_image(2 * k, None, kp1, cal)
        ctx.upload_image(2 * k + 1, None, kp2, cal)
    pairs = np.array([[0, 1], [2, 3]], np.int32)
    mi = np.concatenate([np.c_[np.arange(n), np.arange(n)]] * 2).astype(np.uint32)
    off = np.array([0, n, 2 * n], np.int64)
    keys = np.array([5, 6], np.uint64)
    fixed = ctx.verify_batch(pairs, params_f(3.0, 1024, seed=13, keys=keys), mi, off)
    assert (ctx.verify_hypotheses_used(2) == 1024).all()
    prm = params_f(3.0, 1024, seed=13, keys=keys)
    prm.adaptive_max_hypotheses = 1 << 20
    prm.confidence = 0.999999
    out = ctx.verify_batch(pairs, prm, mi, off)
    used = ctx.verify_hypotheses_used(2)
    assert used[1] == 1024 and 16384 <= used[0] <= (1 << 18), used
    for k in range(2):
        kp1, kp2, R, t, cal = scenes[k]
        got = dict(status=out["status"][k], mask=out["mask"][off[k]:off[k + 1]], R=out["R"][k], t=out["t"][k],
                   n_inliers=out["n_inliers"][k], E=out["E"][k])
        samples = gtsfm_b200.ransac_samples(13, int(keys[k]), n, int(used[k]), 7)
        check_against_oracle_F(got, kp1, kp2, mi[off[k]:off[k + 1]], cal, cal, 3.0, samples)
        # the stopping rule holds for the returned model
        w = got["n_inliers"] / n
        assert np.log(1 - 0.999999) / np.log(1 - w ** 7) <= used[k]
    assert out["n_inliers"][0] >= fixed["n_inliers"][0]
    kp1, kp2, R, t, cal = scenes[0]
    _, cmask = cv2.findFundamentalMat(kp1, kp2, method=cv2.FM_RANSAC, ransacReprojThreshold=3.0, confidence=0.999999, maxIters=1000000)
    assert out["n_inliers"][0] >= 0.95 * int(cmask.sum())
    assert rot_angle_deg(out["R"][0], R) < 2.0


# ------------------------------------------------------------------------------ post-verification report (8(f) rank 2)
def test_two_view_report_batch_vs_oracle(ctx):
    """GT correspondence metrics, relative-pose errors, report aggregates and the InlierSupportProcessor decision of every
    pair of a ragged batch (one pair fails verification, one has few inliers) against the numpy restatement."""
    from oracle import oracle as O
    rng = np.random.default_rng(31)
    sizes = [400, 5, 1500, 30, 900]
    out_fracs = [0.3, 0.0, 0.6, 0.0, 0.97]
    scenes = [make_pixel_scene(rng, max(n, 1), out_frac=f) for n, f in zip(sizes, out_fracs)]
    ctx.clear()
    for k, (kp1, kp2, _, _, cal) in enumerate(scenes):
        ctx.upload_image(2 * k, None, kp1, cal)
        ctx.upload_image(2 * k + 1, None, kp2, cal)
    P = len(sizes)
    pairs = np.array([(2 * k, 2 * k + 1) for k in range(P)], np.int32)
    mi = np.concatenate([np.c_[np.arange(n), np.arange(n)] for n in sizes]).astype(np.uint32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    out = ctx.verify_batch(pairs, params(4.0, 256, seed=2), mi, off)
    gt_R = np.stack([s[2] for s in scenes])
    gt_t = np.stack([s[3] * 2.5 for s in scenes])             # any scale: Unit3 normalises
    thr, min_inl, min_ratio = 1.5, 15, 0.1
    rep = ctx.two_view_report_batch(P, int(off[-1]), thr, min_inl, min_ratio, gt_R, gt_t)
    for k, n in enumerate(sizes):
        kp1, kp2, _, _, cal = scenes[k]
        ok = out["status"][k] == 0
        mask = out["mask"][off[k]:off[k + 1]] == 1
        v = mi[off[k]:off[k + 1]][mask] if ok else np.zeros((0, 2), np.uint32)
        ref = O.two_view_report(kp1, kp2, v, n, out["R"][k] if ok else None, out["t"][k] if ok else None, cal, cal,
                                gt_R[k], gt_t[k], thr, min_inl, min_ratio)
        assert bool(rep["keep"][k]) == ref["keep"], k
        d = rep["dist_sq"][off[k]:off[k + 1]]
        g = rep["gt_mask"][off[k]:off[k + 1]]
        if not ok or len(v) == 0:
            assert np.isnan(d).all() and not g.any() and rep["num_inliers_gt"][k] == 0
            continue
        assert np.isnan(d[~mask]).all() and not g[~mask].any()
        np.testing.assert_allclose(d[mask], ref["dist_sq"], rtol=1e-9, atol=1e-18)
        band = np.abs(ref["dist_sq"] / thr ** 2 - 1.0) < 1e-9
        assert ((g[mask] == 1) == ref["gt_mask"])[~band].all()
        assert abs(int(rep["num_inliers_gt"][k]) - ref["num_inliers_gt"]) <= int(band.sum())
        for a, b in ((rep["inlier_avg_err"][k], ref["inlier_avg"]), (rep["outlier_avg_err"][k], ref["outlier_avg"])):
            assert (np.isnan(a) and np.isnan(b)) or abs(a - b) <= 1e-9 * max(abs(b), 1e-30)
        assert abs(rep["R_error_deg"][k] - ref["R_error_deg"]) < 1e-7
        assert abs(rep["U_error_deg"][k] - ref["U_error_deg"]) < 1e-6
    assert not rep["keep"][1] and rep["keep"][0] and rep["keep"][2]          # too few matches fails, good pairs pass
    # without ground truth only the InlierSupportProcessor decision is defined
    rep2 = ctx.two_view_report_batch(P, int(off[-1]), thr, min_inl, min_ratio)
    np.testing.assert_array_equal(rep2["keep"], rep["keep"])
    assert np.isnan(rep2["dist_sq"]).all() and np.isnan(rep2["R_error_deg"]).all()


def test_two_view_reports_python_layer(ctx):
    """two_view_reports(): TwoViewEstimationReport objects + keep mask; agree with the per-pair reference-style classes."""
    from gtsfm_b200 import synthetic
    from gtsfm_b200.common.two_view_estimation_report import generate_two_view_report
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    from gtsfm_b200.two_view_batch import run_two_view_batch, two_view_reports
    scene = synthetic.make_scene(n_images=4, n_kps=1024, seed=5)
    res = run_two_view_batch(scene.descriptors, scene.keypoints, [scene.cal] * 4, scene.pairs, ratio_test_threshold=0.8,
                             estimation_threshold_px=4.0, num_hypotheses=256, seed=1, ctx=ctx)
    gt = [scene.relative_pose(int(a), int(b)) for a, b in scene.pairs]
    reports, keep, _ = two_view_reports(res, ctx, 4.0, 15, 0.1, np.stack([g[0] for g in gt]), np.stack([g[1] for g in gt]))
    isp = InlierSupportProcessor(15, 0.1)
    Rd, td, vd, rd = res.two_view_dicts()
    for p, (a, b) in enumerate(scene.pairs):
        r = reports[p]
        key = (int(a), int(b))
        ref = generate_two_view_report(rd[key], vd[key], r.R_error_deg, r.U_error_deg, r.v_corr_idxs_inlier_mask_gt, r.reproj_error_gt_model)
        assert ref.num_inliers_gt_model == r.num_inliers_gt_model and ref.num_inliers_est_model == r.num_inliers_est_model
        assert abs(ref.inlier_avg_reproj_error_gt_model - r.inlier_avg_reproj_error_gt_model) < 1e-9
        assert r.R_error_deg < 1.0 and r.num_inliers_gt_model > 0.9 * r.num_inliers_est_model
        kept = isp.run_inlier_support(Rd[key], td[key], vd[key], r)[0] is not None
        assert kept == bool(keep[p])


def test_run_two_view_batch_pinned_outputs_equal_owned_outputs(ctx):
    """pinned_outputs only changes where the results live (the context's page-locked arena), never what they are."""
    from gtsfm_b200 import synthetic
    from gtsfm_b200.two_view_batch import run_two_view_batch
    scene = synthetic.make_scene(n_images=5, n_kps=1024, seed=11)
    kw = dict(ratio_test_threshold=0.8, estimation_threshold_px=4.0, num_hypotheses=128, seed=2, ctx=ctx)
    a = run_two_view_batch(scene.descriptors, scene.keypoints, [scene.cal] * 5, scene.pairs, **kw)
    b = run_two_view_batch(scene.descriptors, scene.keypoints, [scene.cal] * 5, scene.pairs, pinned_outputs=True, **kw)
    for name in ("match_offsets", "matches", "inlier_mask", "R", "t", "n_inliers", "status"):
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name))
    assert not b.matches.flags.owndata and a.matches.flags.owndata


def test_verify_batch_rejects_out_of_range_match_rows(ctx):
    """Caller-supplied match rows are checked against each image's keypoint count (the reference raises IndexError)."""
    import gtsfm_b200
    rng = np.random.default_rng(2)
    kp1, kp2, _, _, cal = make_pixel_scene(rng, 50)
    ctx.clear()
    ctx.upload_image(0, None, kp1, cal)
    ctx.upload_image(1, None, kp2[:40], cal)
    mi = np.c_[np.arange(50), np.arange(50)].astype(np.uint32)
    with pytest.raises(gtsfm_b200.GtsfmB200Error, match="does not have"):
        ctx.verify_batch(np.array([[0, 1]], np.int32), params(4.0, 32), mi, np.array([0, 50], np.int64))
    out = ctx.verify_batch(np.array([[0, 1]], np.int32), params(4.0, 32), mi[:40], np.array([0, 40], np.int64))
    assert out["status"][0] == 0
