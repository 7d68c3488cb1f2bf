"""The cv2 restatement used as bench.py's CPU baseline reproduces the golden vectors of the unmodified reference."""
import glob
import os

import numpy as np
import pytest

from oracle import ref_cv2

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "matcher_*.npz"))), ids=os.path.basename)
def test_match_golden(path):
    g = np.load(path)
    for tag, ratio in (("none", None), ("r08", 0.8), ("r095", 0.95)):
        if "raises_" + tag in g:
            with pytest.raises(ValueError):
                ref_cv2.match_pair(g["d1"], g["d2"], ratio)
            continue
        got, want = ref_cv2.match_pair(g["d1"], g["d2"], ratio), g["m_" + tag]
        if want.shape[0] == 0:
            assert got.size == 0
        else:
            np.testing.assert_array_equal(got, want)


def test_verify_golden_synth2000():
    g = np.load(os.path.join(GOLDEN, "verifier_synth2000.npz"))
    R, t, v, ratio = ref_cv2.verify_pair(g["kp1"], g["kp2"], g["mi"], tuple(g["cal"]), tuple(g["cal"]), 4.0)
    iou = len(set(map(tuple, v)) & set(map(tuple, g["ref_idx"]))) / len(set(map(tuple, v)) | set(map(tuple, g["ref_idx"])))
    assert iou > 0.98
    assert np.degrees(np.arccos(np.clip((np.trace(R.T @ g["ref_R"]) - 1) / 2, -1, 1))) < 0.2
