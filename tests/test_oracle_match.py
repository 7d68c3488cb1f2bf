"""The CPU oracle for TwoWayMatcher against (a) the golden vectors produced by the unmodified reference
(oracle/gen_golden.py) and (b) cv2 itself, the library the reference calls (twoway_matcher.py:102,136,139)."""
import glob
import os

import cv2
import numpy as np
import pytest

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAGS = (("none", None), ("r08", 0.8), ("r095", 0.95))


def test_reference_known_answer_vectors():
    # tests/frontend/matcher/test_twoway_withoutratiotest_matcher.py:25-48 / test_twoway_withratiotest_matcher.py:25-49
    g = np.load(os.path.join(GOLDEN, "matcher_dummy1d.npz"))
    np.testing.assert_array_equal(O.twoway_match(g["d1"], g["d2"], None), [[9, 5], [2, 4], [3, 2], [1, 7], [8, 6], [0, 3]])
    np.testing.assert_array_equal(O.twoway_match(g["d1"], g["d2"], 0.8), [[9, 5], [2, 4], [3, 2], [0, 3]])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "matcher_*.npz"))), ids=os.path.basename)
def test_oracle_matches_golden(path):
    g = np.load(path)
    for tag, ratio in TAGS:
        if "raises_" + tag in g:
            with pytest.raises(ValueError):
                O.twoway_match(g["d1"], g["d2"], ratio)
            continue
        got = O.twoway_match(g["d1"], g["d2"], ratio)
        want = g["m_" + tag]
        if want.shape[0] == 0:
            assert got.size == 0 and got.shape == (0,)
        else:
            assert got.dtype == np.uint32
            np.testing.assert_array_equal(got, want)


def test_oracle_matches_golden_lund_door():
    g = np.load(os.path.join(GOLDEN, "lund_door_pair.npz"))
    d0, d1 = g["d0_u8"].astype(np.float32), g["d1_u8"].astype(np.float32)
    np.testing.assert_array_equal(O.twoway_match(d0, d1, 0.8), g["m_r08"])
    np.testing.assert_array_equal(O.twoway_match(d0, d1, None), g["m_none"])


@pytest.mark.parametrize("dim", [1, 3, 16, 37, 128, 256])
def test_l2_bitexact_vs_cv2(dim):
    rng = np.random.default_rng(dim)
    a = rng.standard_normal((40, dim)).astype(np.float32)
    b = rng.standard_normal((55, dim)).astype(np.float32)
    idx, d1, d2 = O.knn2(a, b)
    knn = cv2.BFMatcher(cv2.NORM_L2).knnMatch(a, b, k=2)
    for i, (m1, m2) in enumerate(knn):
        assert m1.trainIdx == idx[i]
        assert np.float32(m1.distance).tobytes() == d1[i].tobytes()
        assert np.float32(m2.distance).tobytes() == d2[i].tobytes()


def test_tie_break_lowest_train_index():
    a = np.zeros((3, 4), np.float32)
    b = np.zeros((5, 4), np.float32)
    idx, d1, d2 = O.knn2(a, b)
    assert (idx == 0).all() and (d1 == 0).all() and (d2 == 0).all()
    m = cv2.BFMatcher(cv2.NORM_L2).match(a, b)
    assert all(x.trainIdx == 0 for x in m)


def test_empty_inputs():
    e = np.array([])
    d = np.ones((4, 8), np.float32)
    for a, b in ((e, d), (d, e), (e, e)):
        r = O.twoway_match(a, b, None)
        assert r.size == 0 and r.shape == (0,)


def test_hamming_oracle_matches_reference_golden():
    """TwoWayMatcher(distance_type=HAMMING) of the unmodified reference on ORB-like 256-bit descriptors (gen_golden.py hamming)."""
    g = np.load(os.path.join(GOLDEN, "hamming_orb256.npz"))
    for tag, ratio in (("none", None), ("r08", 0.8), ("r095", 0.95)):
        got = O.twoway_match_hamming(g["d1"], g["d2"], ratio).reshape(-1, 2)
        np.testing.assert_array_equal(got, g["m_" + tag])


def test_oracle_reproduces_reference_matches_on_c1_pairs(golden_dir):
    """BASELINE config C1 (12 Lund-door images, reference SIFT + TwoWayMatcher(0.8), oracle/gen_golden.py c1): the
    oracle's matches equal the recorded reference output on the first and the last of the 66 pairs."""
    import os
    from oracle import oracle as O
    g = np.load(os.path.join(golden_dir, "c1_lund_door_12.npz"))
    off, mo = g["img_off"], g["m_off"]
    for k in (0, 65):
        i, j = g["pairs"][k]
        m = O.twoway_match(g["desc_u8"][off[i]:off[i + 1]].astype(np.float32), g["desc_u8"][off[j]:off[j + 1]].astype(np.float32), 0.8)
        np.testing.assert_array_equal(m, g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32))
