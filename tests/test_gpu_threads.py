"""Re-entrancy of the per-pair plugins (SURVEY 8(b) "Threading"): a Dask worker may run match() / verify() in several
threads of one process (--threads_per_worker > 1, gtsfm_runner_base.py:58-69).  Every thread gets its own device context
(gtsfm_b200._native.default_context), so concurrent calls must return what the same calls return one after another.
"""
import threading

import numpy as np
import pytest

from tests.test_gpu_match import sift_like
from tests.test_gpu_verify import make_pixel_scene

pytestmark = pytest.mark.gpu


def one_job(seed):
    from gtsfm_b200.common.types import Cal3Bundler, Keypoints
    from gtsfm_b200.frontend.matcher.twoway_matcher import TwoWayMatcher
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    rng = np.random.default_rng(seed)
    n = 700 + 100 * (seed % 4)
    kp1, kp2, _, _, cal = make_pixel_scene(rng, n, out_frac=0.2)
    protos = np.abs(rng.standard_normal((n, 128)))
    d1, d2 = sift_like(rng, n, protos=protos), sift_like(rng, n, protos=protos)
    k1, k2 = Keypoints(coordinates=kp1), Keypoints(coordinates=kp2)
    m = TwoWayMatcher(ratio_test_threshold=0.8).match(k1, k2, d1, d2, (960, 1280, 3), (960, 1280, 3))
    c = Cal3Bundler(*cal)
    R, U, v, ratio = Ransac(True, 4.0, num_hypotheses=256, seed=seed).verify(k1, k2, m, c, c)
    return m, R.matrix(), U.point3(), v, ratio


def test_concurrent_match_and_verify_equal_sequential():
    from gtsfm_b200 import _native
    seeds = list(range(8))
    want = [one_job(s) for s in seeds]
    got = [None] * len(seeds)
    ctxs = []                                       # kept alive so that ids stay distinct
    errors = []

    def worker(k):
        try:
            for _ in range(3):
                got[k] = one_job(seeds[k])
            ctxs.append(_native.default_context(0))
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(seeds))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert len({id(c) for c in ctxs}) == len(seeds)  # one context per thread
    for w, g in zip(want, got):
        np.testing.assert_array_equal(w[0], g[0])
        np.testing.assert_array_equal(w[1], g[1])
        np.testing.assert_array_equal(w[2], g[2])
        np.testing.assert_array_equal(w[3], g[3])
        assert w[4] == g[4]
