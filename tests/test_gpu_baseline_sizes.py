"""Oracle parity at the BASELINE.json sizes (C2 8192 x 8192 x 128, C3-shaped ragged chunked batches,
C4 2048 x 2048 x 256 floats with the ratio test off, C5 20,480 keypoints / 25 % inliers / 10,000 hypotheses) and the radial
distortion branch of normalize_coordinates.  Everything goes through the C ABI (gtsfm_b200.Context) and is compared with the
CPU oracle (oracle/liboracle.so, threads over independent query rows): matches bit-exact including order, inlier masks
identical outside the 1e-6 band around the threshold, R and t within 0.1 degree on identical sample sets.
"""
import numpy as np
import pytest

from tests.test_gpu_verify import check_against_oracle, dir_angle_deg, params, rot_angle_deg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import gtsfm_b200
    c = gtsfm_b200.Context(0)
    yield c
    c.close()


def _match_and_verify_vs_oracle(ctx, descs, kps, cal, pairs, ratio, H, seed, verify=True, lo=20):
    """Upload, match the pair list in one batch, verify on the device-resident matches; compare every pair with the oracle."""
    import gtsfm_b200
    from oracle import oracle as O
    ctx.clear()
    for i, (d, k) in enumerate(zip(descs, kps)):
        ctx.upload_image(i, d, k, cal)
    pairs = np.ascontiguousarray(pairs, np.int32)
    m, off = ctx.match_batch(pairs, ratio)
    out = ctx.verify_batch(pairs, params(4.0, H, lo=lo, seed=seed), total=len(m)) if verify else None
    n_checked = 0
    for p, (i, j) in enumerate(pairs):
        want = O.twoway_match(descs[i], descs[j], ratio).reshape(-1, 2)
        got = m[off[p]:off[p + 1]]
        assert got.shape == want.shape, f"pair {p} ({i},{j}): {len(got)} matches, oracle {len(want)}"
        assert (got == want).all(), f"pair {p} ({i},{j}): match rows differ from the oracle"
        if not verify:
            continue
        g = dict(status=out["status"][p], mask=out["mask"][off[p]:off[p + 1]], R=out["R"][p], t=out["t"][p], n_inliers=out["n_inliers"][p])
        if len(want) < 6:
            assert g["status"] == 1
            continue
        samples = gtsfm_b200.ransac_samples(seed, p, len(want), H)
        check_against_oracle(g, kps[i], kps[j], want, cal, cal, 4.0, samples, lo=lo)
        n_checked += 1
    return m, off, out, n_checked


def test_c2_pairs_8192x8192x128_vs_oracle(ctx):
    """BASELINE config C2 (seed 20260002): a neighbouring and a wide-baseline pair of the 100-image scene, ratio 0.8 and ratio
    off, then the verifier on the ~6.6k matches with H = 1,024."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_config("C2")
    pick = [0, 1, 57]
    descs = [scene.descriptors[i] for i in pick]
    kps = [scene.keypoints[i] for i in pick]
    pairs = [(0, 1), (0, 2), (1, 2)]
    m, off, out, n = _match_and_verify_vs_oracle(ctx, descs, kps, scene.cal, pairs, 0.8, 1024, seed=20260002)
    assert n == 3 and (np.diff(off) > 3000).all()
    Rgt, tgt = scene.relative_pose(0, 1)
    assert rot_angle_deg(out["R"][0], Rgt) < 0.5 and dir_angle_deg(out["t"][0], tgt) < 2.0
    _match_and_verify_vs_oracle(ctx, descs, kps, scene.cal, pairs[:2], None, 1024, seed=3, verify=False)


def test_c4_pairs_2048x2048x256_float_screened_vs_oracle(ctx):
    """BASELINE config C4's shape (SuperPoint-like unit-norm 256-d floats, 2,048 keypoints, ratio test off): the fp16
    tensor-core screening + exact resolve route against the ORACLE (cv2's float32 arithmetic), every pair of 6 images."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_config("C4", n_images=6)
    assert scene.descriptors[0].shape == (2048, 256) and scene.ratio is None
    ctx.set_float_screening(True)
    pairs = [(i, j) for i in range(6) for j in range(i + 1, 6)]
    m, off, out, n = _match_and_verify_vs_oracle(ctx, scene.descriptors, scene.keypoints, scene.cal, pairs, None, 1024, seed=20260004)
    assert n == len(pairs) and (np.diff(off) > 500).all()


def test_c5_pair_20480_keypoints_10000_hypotheses_vs_oracle(ctx):
    """BASELINE config C5's shape: 20,480 keypoints per image (160 lane-class rows, 80 train blocks), 25 % inliers, 10,000
    hypotheses."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_config("C5", n_images=2)
    assert scene.descriptors[0].shape == (20480, 128) and scene.n_hyp == 10000
    m, off, out, n = _match_and_verify_vs_oracle(ctx, scene.descriptors, scene.keypoints, scene.cal, [(0, 1)], 0.8, 10000, seed=20260005)
    assert n == 1 and len(m) > 10000
    frac = out["n_inliers"][0] / len(m)
    assert 0.15 < frac < 0.40, frac
    Rgt, tgt = scene.relative_pose(0, 1)
    assert rot_angle_deg(out["R"][0], Rgt) < 0.5


def test_c3_shaped_ragged_chunked_batch_vs_oracle(ctx):
    """A C3-shaped job in small: 18 images of 3,000-8,192 keypoints (ragged lane classes / train blocks inside one batch), top-3
    pair list, 160 MiB of workspace so the batch runs as several chunks; every pair against the per-pair oracle."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_scene(n_images=18, n_kps=8192, pairs="topk", top_k=3, seed=20260003, visible=0.6)
    rng = np.random.default_rng(33)
    sizes = rng.integers(3000, 8193, 18)
    sizes[3], sizes[11] = 8192, 3001
    descs = [d[:n] for d, n in zip(scene.descriptors, sizes)]
    kps = [k[:n] for k, n in zip(scene.keypoints, sizes)]
    ctx.set_workspace(160 << 20)
    try:
        m, off, out, n = _match_and_verify_vs_oracle(ctx, descs, kps, scene.cal, scene.pairs, 0.8, 512, seed=20260003)
    finally:
        ctx.set_workspace(12 << 30)
    assert n == len(scene.pairs) >= 45


@pytest.mark.parametrize("k1,k2", [(-0.12, 0.03), (0.05, -0.01)])
def test_radial_distortion_branch_of_normalize_coordinates(ctx, k1, k2):
    """Cal3Bundler with k1, k2 != 0 (gtsfm/utils/features.py:40-50 -> gtsam Cal3Bundler.calibrate: fixed-point undistortion)
    through gtsfm_verify_pair against the oracle's calibrate; different intrinsics per image."""
    import gtsfm_b200
    from oracle import oracle as O
    rng = np.random.default_rng(17)
    n, H = 2000, 512
    cal1, cal2 = (900.0, k1, k2, 650.0, 470.0), (1100.0, 0.5 * k1, 2.0 * k2, 630.0, 490.0)
    X = rng.uniform(-2, 2, (n, 3)) + [0, 0, 6]
    w = rng.normal(0, 0.15, 3)
    th = np.linalg.norm(w)
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]]) / th
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    t = rng.normal(0, 1, 3)
    t /= np.linalg.norm(t)

    def distort(xn, cal):
        r2 = (xn ** 2).sum(1, keepdims=True)
        return cal[0] * (1 + (cal[1] + cal[2] * r2) * r2) * xn + [cal[3], cal[4]]

    x1 = X[:, :2] / X[:, 2:]
    Xc = X @ R.T + t
    x2 = Xc[:, :2] / Xc[:, 2:]
    p1 = distort(x1, cal1) + rng.normal(0, 0.5, x1.shape)
    p2 = distort(x2, cal2) + rng.normal(0, 0.5, x2.shape)
    p2[: n // 3] = rng.uniform(0, 1, (n // 3, 2)) * [1280, 960]
    mi = np.c_[np.arange(n), np.arange(n)].astype(np.uint32)
    # the oracle's calibrate inverts the distortion model to gtsam's tolerance (1e-5 px^2)
    back = distort(O.calibrate(p1, cal1), cal1)
    assert ((back - p1) ** 2).sum(1).max() <= 1e-5
    assert np.abs(O.calibrate(p1, cal1) - (p1 - [cal1[3], cal1[4]]) / cal1[0]).max() > 1e-3     # the branch does something
    samples = gtsfm_b200.ransac_samples(7, 0, n, H)
    got = ctx.verify_pair(p1, p2, mi, cal1, cal2, params(4.0, H))
    check_against_oracle(got, p1, p2, mi, cal1, cal2, 4.0, samples)
    assert got["status"] == 0 and got["n_inliers"] > 0.55 * n
    assert rot_angle_deg(got["R"], R) < 0.5 and dir_angle_deg(got["t"], t) < 2.0
