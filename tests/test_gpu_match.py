"""GPU parity of the batched TwoWayMatcher against the CPU oracle and the reference's golden vectors (bit-exact)."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAGS = (("none", None), ("r08", 0.8), ("r095", 0.95))


@pytest.fixture(scope="module")
def ctx():
    import gtsfm_b200
    c = gtsfm_b200.Context(0)
    yield c
    c.close()


def sift_like(rng, n, dim=128, protos=None, noise=0.03):
    """integer-valued float32 descriptors in [0,255] with L2 norm ~512."""
    p = np.abs(rng.standard_normal((n, dim))) if protos is None else protos
    p = p / np.linalg.norm(p, axis=1, keepdims=True)
    p = np.minimum(p, 0.2)
    p = p / np.linalg.norm(p, axis=1, keepdims=True)
    p = p + rng.normal(0, noise, p.shape)
    p = np.abs(p) / np.linalg.norm(p, axis=1, keepdims=True)
    return np.clip(np.rint(p * 512), 0, 255).astype(np.float32)


def check(got, want):
    if want.shape[0] == 0:
        assert got.shape[0] == 0
    else:
        assert got.dtype == np.uint32
        np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "matcher_*.npz"))), ids=os.path.basename)
def test_golden_vectors(ctx, path):
    g = np.load(path)
    for tag, ratio in TAGS:
        if "raises_" + tag in g:
            with pytest.raises(ValueError):
                ctx.match_pair(g["d1"], g["d2"], ratio)
            continue
        check(ctx.match_pair(g["d1"], g["d2"], ratio), g["m_" + tag])


def test_lund_door_pair(ctx):
    g = np.load(os.path.join(GOLDEN, "lund_door_pair.npz"))
    d0, d1 = g["d0_u8"].astype(np.float32), g["d1_u8"].astype(np.float32)
    check(ctx.match_pair(d0, d1, 0.8), g["m_r08"])
    check(ctx.match_pair(d0, d1, None), g["m_none"])


@pytest.mark.parametrize("n1,n2", [(1000, 900), (256, 256), (300, 2049), (2500, 130), (5, 3)])
def test_integer_path_vs_oracle(ctx, n1, n2):
    from oracle import oracle as O
    rng = np.random.default_rng(n1 * 7 + n2)
    shared = min(n1, n2) // 2
    protos = np.abs(rng.standard_normal((shared, 128)))
    d1 = np.concatenate([sift_like(rng, shared, protos=protos), sift_like(rng, n1 - shared)])[rng.permutation(n1)]
    d2 = np.concatenate([sift_like(rng, shared, protos=protos), sift_like(rng, n2 - shared)])[rng.permutation(n2)]
    for ratio in (None, 0.8):
        want = O.twoway_match(d1, d2, ratio)
        check(ctx.match_pair(d1, d2, ratio), want.reshape(-1, 2))


def test_integer_ties_and_duplicates(ctx):
    """Duplicate descriptors force equal distances: lowest-index tie-breaks and the stable output order must hold."""
    from oracle import oracle as O
    rng = np.random.default_rng(5)
    base = sift_like(rng, 200)
    d1 = np.concatenate([base, base[:50], sift_like(rng, 100)])
    d2 = np.concatenate([base[::-1], base[:30], sift_like(rng, 77)])
    for ratio in (None, 0.8, 1.0):
        check(ctx.match_pair(d1, d2, ratio), O.twoway_match(d1, d2, ratio).reshape(-1, 2))


def test_integer_small_values_dim64(ctx):
    from oracle import oracle as O
    rng = np.random.default_rng(9)
    d1 = rng.integers(0, 4, (700, 64)).astype(np.float32)
    d2 = rng.integers(0, 4, (650, 64)).astype(np.float32)
    for ratio in (None, 0.9):
        check(ctx.match_pair(d1, d2, ratio), O.twoway_match(d1, d2, ratio).reshape(-1, 2))


def test_float_path_vs_oracle(ctx):
    from oracle import oracle as O
    rng = np.random.default_rng(3)
    for dim in (1, 37, 128, 256):
        d1 = rng.standard_normal((400, dim)).astype(np.float32)
        d2 = (d1[rng.permutation(400)[:350]] + 0.1 * rng.standard_normal((350, dim))).astype(np.float32)
        for ratio in (None, 0.8):
            check(ctx.match_pair(d1, d2, ratio), O.twoway_match(d1, d2, ratio).reshape(-1, 2))


def test_nan_rows_integer_path(ctx):
    from oracle import oracle as O
    rng = np.random.default_rng(4)
    d1, d2 = sift_like(rng, 600), sift_like(rng, 500)
    d2[:200] = d1[100:300]
    d1[[3, 150, 599]] = np.nan
    d2[[0, 120]] = np.nan
    for ratio in (None, 0.8):
        check(ctx.match_pair(d1, d2, ratio), O.twoway_match(d1, d2, ratio).reshape(-1, 2))


def test_empty_inputs(ctx):
    e = np.zeros((0, 128), np.float32)
    d = np.ones((4, 128), np.float32)
    for a, b in ((e, d), (d, e), (e, e)):
        assert ctx.match_pair(a, b, None).shape[0] == 0


def test_batch_of_pairs_matches_single_pair_calls(ctx):
    """Ragged batch, tiny workspace so several chunks are used; results must equal one-pair-at-a-time results."""
    from oracle import oracle as O
    rng = np.random.default_rng(11)
    sizes = [700, 1300, 256, 90, 2100, 513]
    protos = np.abs(rng.standard_normal((2100, 128)))
    imgs = [sift_like(rng, n, protos=protos[rng.permutation(2100)[:n]]) for n in sizes]
    pairs = np.array([(i, j) for i in range(len(sizes)) for j in range(i + 1, len(sizes))], np.int32)
    ctx.set_workspace(64 << 20)
    ctx.clear()
    for k, d in enumerate(imgs):
        ctx.upload_image(k, d)
    m, off = ctx.match_batch(pairs, 0.8)
    ctx.set_workspace(4 << 30)
    for p, (i, j) in enumerate(pairs):
        check(m[off[p]:off[p + 1]], O.twoway_match(imgs[i], imgs[j], 0.8).reshape(-1, 2))


def test_superpoint_like_batch_ratio_off(ctx):
    """C4-shaped inputs: unit-norm float 256-d descriptors, ratio test off, several pairs in one batch."""
    from gtsfm_b200 import synthetic
    from oracle import oracle as O
    scene = synthetic.make_scene(n_images=4, n_kps=700, dim=256, kind="superpoint", seed=20260004, ratio=None)
    ctx.clear()
    for i, d in enumerate(scene.descriptors):
        ctx.upload_image(i, d)
    m, off = ctx.match_batch(scene.pairs, None)
    for p, (i, j) in enumerate(scene.pairs):
        want = O.twoway_match(scene.descriptors[i], scene.descriptors[j], None).reshape(-1, 2)
        check(m[off[p]:off[p + 1]], want)
    assert off[-1] > 0


# ------------------------------------------------------------------ float descriptors: fp16 tensor-core screening path
def unit_float_descs(rng, n, dim, protos=None, noise=0.1):
    p = rng.standard_normal((n, dim)) if protos is None else protos + noise * rng.standard_normal(protos.shape)
    return (p / np.linalg.norm(p, axis=1, keepdims=True)).astype(np.float32)


@pytest.mark.parametrize("n1,n2,dim", [(900, 1000, 256), (256, 256, 128), (300, 2049, 64), (2500, 130, 256), (5, 3, 128)])
def test_float_screened_path_vs_oracle(ctx, n1, n2, dim):
    """Unit-norm float descriptors, ratio test off: fp16 screening + exact recompute must reproduce cv2's float32 matches
    bit for bit (indices and order), and equal the exhaustive kernel's result."""
    from oracle import oracle as O
    rng = np.random.default_rng(n1 * 3 + n2 + dim)
    shared = min(n1, n2) // 2
    protos = rng.standard_normal((shared, dim))
    d1 = np.concatenate([unit_float_descs(rng, shared, dim, protos), unit_float_descs(rng, n1 - shared, dim)])[rng.permutation(n1)]
    d2 = np.concatenate([unit_float_descs(rng, shared, dim, protos), unit_float_descs(rng, n2 - shared, dim)])[rng.permutation(n2)]
    want = O.twoway_match(d1, d2, None).reshape(-1, 2)
    got = ctx.match_pair(d1, d2, None)
    check(got, want)
    ctx.set_float_screening(False)
    try:
        check(ctx.match_pair(d1, d2, None), want)
    finally:
        ctx.set_float_screening(True)
    # ratio test on: the screened route certifies d1 <= ratio * d2 from the screening bounds or recomputes d2 exactly
    for ratio in (0.9, 0.8, 0.6, 1.0):
        want_r = O.twoway_match(d1, d2, ratio).reshape(-1, 2)
        check(ctx.match_pair(d1, d2, ratio), want_r)
    ctx.set_float_screening(False)
    try:
        check(ctx.match_pair(d1, d2, 0.8), O.twoway_match(d1, d2, 0.8).reshape(-1, 2))
    finally:
        ctx.set_float_screening(True)


def test_float_screened_ties_nan_and_scales(ctx):
    """Duplicate rows (exact ties -> lowest index), NaN rows, rows of different norms inside the eligible range, and
    descriptors outside it (|x|^2 > 1024 -> exhaustive kernel)."""
    from oracle import oracle as O
    rng = np.random.default_rng(77)
    base = unit_float_descs(rng, 300, 128)
    d1 = np.concatenate([base, base[:60], unit_float_descs(rng, 150, 128)])
    d2 = np.concatenate([base[::-1], base[:40], unit_float_descs(rng, 90, 128)])
    check(ctx.match_pair(d1, d2, None), O.twoway_match(d1, d2, None).reshape(-1, 2))
    d1n, d2n = d1.copy(), d2.copy()
    d1n[[3, 150, 400]] = np.nan
    d2n[[0, 120]] = np.nan
    check(ctx.match_pair(d1n, d2n, None), O.twoway_match(d1n, d2n, None).reshape(-1, 2))
    s1 = (d1 * rng.uniform(0.05, 20.0, (d1.shape[0], 1))).astype(np.float32)
    s2 = (d2 * rng.uniform(0.05, 20.0, (d2.shape[0], 1))).astype(np.float32)
    check(ctx.match_pair(s1, s2, None), O.twoway_match(s1, s2, None).reshape(-1, 2))
    b1, b2 = (d1 * 512).astype(np.float32), (d2 * 512).astype(np.float32)
    check(ctx.match_pair(b1, b2, None), O.twoway_match(b1, b2, None).reshape(-1, 2))
    z = np.zeros((40, 128), np.float32)                       # all-zero rows: every distance ties
    check(ctx.match_pair(z, d2, None), O.twoway_match(z, d2, None).reshape(-1, 2))
    # the same inputs with the ratio test on: exact duplicates make d1 == d2 (kept only when ratio >= 1), NaN rows, scales
    for ratio in (0.8, 1.0, 0.99):
        check(ctx.match_pair(d1, d2, ratio), O.twoway_match(d1, d2, ratio).reshape(-1, 2))
        check(ctx.match_pair(d1n, d2n, ratio), O.twoway_match(d1n, d2n, ratio).reshape(-1, 2))
        check(ctx.match_pair(s1, s2, ratio), O.twoway_match(s1, s2, ratio).reshape(-1, 2))
        check(ctx.match_pair(z, d2, ratio), O.twoway_match(z, d2, ratio).reshape(-1, 2))


def test_float_screened_ratio_on_c4_size_vs_oracle(ctx):
    """TwoWayMatcher(ratio_test_threshold=0.8) on float descriptors (the reference's kaze.yaml / d2net.yaml correspondence
    configs) at C4's pair size, 2048 x 2048 x 256, batched: the tensor-core route against the ORACLE for every pair, with
    descriptors whose nearest / second-nearest ratios straddle the threshold (noise levels 0.1 .. 1.4 of the prototype amplitude)."""
    from oracle import oracle as O
    rng = np.random.default_rng(4242)
    n, dim = 2048, 256
    protos = rng.standard_normal((n, dim))
    imgs = []
    for noise in (0.1, 0.6, 1.0, 1.4):
        d = unit_float_descs(rng, n, dim, protos, noise=noise)
        d[: n // 4] = unit_float_descs(rng, n // 4, dim)
        imgs.append(d[rng.permutation(n)])
    pairs = np.array([(i, j) for i in range(4) for j in range(i + 1, 4)], np.int32)
    ctx.clear()
    for i, d in enumerate(imgs):
        ctx.upload_image(i, d)
    for ratio in (0.8, 0.95):
        m, off = ctx.match_batch(pairs, ratio)
        n_flag = int(ctx.debug_fetch(10, 1, np.int32)[0])
        kept = []
        for p, (i, j) in enumerate(pairs):
            want = O.twoway_match(imgs[i], imgs[j], ratio).reshape(-1, 2)
            check(m[off[p]:off[p + 1]], want)
            kept.append(len(want))
        assert min(kept) > 50 and n_flag < 0.25 * 2 * n * len(pairs), (kept, n_flag)


def test_float_screened_c4_size_flag_rate(ctx):
    """C4-sized pair (2048 x 2048 x 256): screened == exhaustive, and only a small fraction of the rows and
    columns needs the exhaustive re-scan."""
    from gtsfm_b200 import synthetic
    scene = synthetic.make_scene(n_images=3, n_kps=2048, dim=256, kind="superpoint", seed=20260004, ratio=None)
    ctx.clear()
    for i, d in enumerate(scene.descriptors):
        ctx.upload_image(i, d)
    m, off = ctx.match_batch(scene.pairs, None)
    n_flag = int(ctx.debug_fetch(10, 1, np.int32)[0])
    ctx.set_float_screening(False)
    try:
        m2, off2 = ctx.match_batch(scene.pairs, None)
    finally:
        ctx.set_float_screening(True)
    np.testing.assert_array_equal(off, off2)
    np.testing.assert_array_equal(m, m2)
    assert off[-1] > 1000
    total = sum(len(scene.descriptors[i]) + len(scene.descriptors[j]) for i, j in scene.pairs)
    assert n_flag < 0.25 * total, (n_flag, total)
    print("flagged rows+columns:", n_flag, "of", total)


# ------------------------------------------------------------------------------------------ HAMMING (binary descriptors)
def test_hamming_golden_and_oracle(ctx):
    """TwoWayMatcher(distance_type=HAMMING): the reference's recorded matches on ORB-like 256-bit descriptors, then random
    32- and 61-byte descriptors (heavy ties: integer distances) against the oracle, ratio test on and off."""
    from oracle import oracle as O
    g = np.load(os.path.join(GOLDEN, "hamming_orb256.npz"))
    for tag, ratio in TAGS:
        check(ctx.match_pair_u8(g["d1"], g["d2"], ratio), g["m_" + tag])
    rng = np.random.default_rng(8)
    for n1, n2, nb in ((700, 900, 32), (300, 257, 61), (2, 5, 32), (1100, 64, 64)):
        base = rng.integers(0, 256, (min(n1, n2), nb), dtype=np.uint8)
        d1 = np.concatenate([base, rng.integers(0, 256, (n1 - len(base), nb), dtype=np.uint8)])[rng.permutation(n1)]
        flip = (rng.random((len(base), nb)) < 0.05).astype(np.uint8) * rng.integers(1, 256, (len(base), nb), dtype=np.uint8)
        d2 = np.concatenate([base ^ flip, rng.integers(0, 256, (n2 - len(base), nb), dtype=np.uint8)])[rng.permutation(n2)]
        for ratio in (None, 0.8):
            check(ctx.match_pair_u8(d1, d2, ratio), O.twoway_match_hamming(d1, d2, ratio).reshape(-1, 2))


def test_hamming_plugin_and_batch(ctx):
    from gtsfm_b200.frontend.matcher.twoway_matcher import MatchingDistanceType, TwoWayMatcher
    from oracle import oracle as O
    rng = np.random.default_rng(9)
    imgs = [rng.integers(0, 256, (n, 32), dtype=np.uint8) for n in (500, 260, 900)]
    imgs[1][:200] = imgs[0][100:300]
    imgs[2][:300] = imgs[0][:300] ^ (rng.random((300, 32)) < 0.03).astype(np.uint8)
    m = TwoWayMatcher(distance_type=MatchingDistanceType.HAMMING, ratio_test_threshold=0.8)
    pairs = np.array([[0, 1], [0, 2], [1, 2]], np.int32)
    out = m.match_batch(imgs, pairs)
    for (i, j), got in zip(pairs, out):
        want = O.twoway_match_hamming(imgs[i], imgs[j], 0.8).reshape(-1, 2)
        check(np.asarray(got).reshape(-1, 2), want)
        single = m.match(None, None, imgs[i], imgs[j])
        check(np.asarray(single).reshape(-1, 2), want)
    with pytest.raises(Exception):                       # one store cannot mix float32 and binary descriptors
        ctx.clear(); ctx.upload_image_u8(0, imgs[0]); ctx.upload_image(1, np.ones((4, 32), np.float32))
    ctx.clear()


def test_hamming_multi_chunk_batch_last_chunk_single_pair(ctx):
    """A binary-descriptor batch that runs as several chunks (64 MiB of workspace: 8,192-row pairs take 80 B per row and
    column, so 102 pairs fit and the 103rd opens a chunk of its own), plus an image without any features.  Every pair equals
    the single-pair entry point; the last pair and two others also equal the oracle."""
    from gtsfm_b200.frontend.matcher.twoway_matcher import MatchingDistanceType, TwoWayMatcher
    from oracle import oracle as O
    rng = np.random.default_rng(91)
    n_img, n = 15, 8192
    base = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    imgs = []
    for i in range(n_img):
        d = base ^ (rng.random((n, 32)) < 0.02).astype(np.uint8) * rng.integers(1, 256, (n, 32), dtype=np.uint8)
        d[: n // 4] = rng.integers(0, 256, (n // 4, 32), dtype=np.uint8)
        imgs.append(d[rng.permutation(n)])
    imgs.append(np.array([]))                                       # image 15: no features (twoway_matcher.py:71-72)
    all_pairs = [(i, j) for i in range(n_img) for j in range(i + 1, n_img)]
    pairs = np.array(all_pairs[:102] + [(3, 15)] + all_pairs[102:103], np.int32)          # 104 pairs, 103 of them non-empty
    m = TwoWayMatcher(distance_type=MatchingDistanceType.HAMMING, ratio_test_threshold=0.8)
    from gtsfm_b200 import _native
    dc = _native.default_context(0)                                 # the context the plugin's match_batch runs on
    dc.set_workspace(64 << 20)
    try:
        out = m.match_batch(imgs, pairs)
    finally:
        dc.set_workspace(12 << 30)
    assert len(out) == len(pairs) and np.asarray(out[102]).size == 0
    for p, (i, j) in enumerate(pairs):
        if j == 15:
            continue
        single = ctx.match_pair_u8(imgs[i], imgs[j], 0.8)
        check(np.asarray(out[p]).reshape(-1, 2), single)
        assert len(single) > 1000
    for p in (0, 57, 103):
        i, j = pairs[p]
        check(np.asarray(out[p]).reshape(-1, 2), O.twoway_match_hamming(imgs[i], imgs[j], 0.8).reshape(-1, 2))
