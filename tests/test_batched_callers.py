"""the batched callers with the reference's signatures --
DetDescCorrespondenceGenerator.generate_correspondences(client, images, image_pairs)
(gtsfm/frontend/correspondence_generator/det_desc_correspondence_generator.py:31-85) and
run_two_view_estimator_as_futures(client, two_view_estimator, keypoints_list, putative_corr_idxs_dict, camera_intrinsics,
relative_pose_priors, gt_cameras, gt_scene_mesh) (gtsfm/two_view_estimator.py:531-587).

CPU part: host logic with a minimal fake Dask client and stub plugins (task counts, dictionary keys / order, shard contents).
GPU part: the C1 fixture (12 Lund-door images, 66 pairs) -- the returned structures equal the per-pair plugin path, the
recorded reference matches, and the oracle's report restatement.
"""
import os

import numpy as np
import pytest

from tests.fake_dask import FakeClient

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class FakeImage:
    def __init__(self, idx, shape=(760, 1135, 3)):
        self.idx, self.shape = idx, shape


class FixtureDetDesc:
    """Detector-descriptor stub: returns precomputed (Keypoints, descriptors) for image idx."""

    def __init__(self, keypoints, descriptors):
        self._k, self._d = keypoints, descriptors

    def detect_and_describe(self, image):
        return self._k[image.idx], self._d[image.idx]


class PerPairMatcher:
    """A foreign MatcherBase-style plugin: only match()."""

    def __init__(self):
        self.calls = 0

    def match(self, k1, k2, d1, d2, s1, s2):
        self.calls += 1
        return np.array([[len(d1), len(d2)]], np.uint32)


class BatchStubMatcher(PerPairMatcher):
    def __init__(self):
        super().__init__()
        self.batches = []

    def match_batch(self, descriptors, pairs, device=None):
        self.batches.append((len(descriptors), np.array(pairs), device))
        return [np.array([[len(descriptors[i]), len(descriptors[j])]], np.uint32) for i, j in pairs]


def _stub_features(n=7):
    from gtsfm_b200.common.types import Keypoints
    kps = [Keypoints(coordinates=np.full((10 + i, 2), float(i))) for i in range(n)]
    descs = [np.zeros((10 + i, 8), np.float32) for i in range(n)]
    return kps, descs


@pytest.mark.parametrize("use_client", [True, False])
def test_generator_per_pair_plugin_follows_the_reference_loop(use_client):
    from gtsfm_b200.frontend.correspondence_generator.det_desc_correspondence_generator import DetDescCorrespondenceGenerator
    kps, descs = _stub_features()
    m = PerPairMatcher()
    gen = DetDescCorrespondenceGenerator(m, FixtureDetDesc(kps, descs))
    pairs = [(0, 1), (2, 5), (1, 6)]
    client = FakeClient() if use_client else None
    images = [FakeImage(i) for i in range(7)]
    keypoints_list, corr = gen.generate_correspondences(client, images, pairs)
    assert [k.coordinates.shape[0] for k in keypoints_list] == [10 + i for i in range(7)]
    assert list(corr.keys()) == pairs and m.calls == 3
    assert (corr[(2, 5)] == [[12, 15]]).all()
    if use_client:
        assert client.count("_apply_matcher") == 3 and client.count("_apply_det_desc") == 7


@pytest.mark.parametrize("num_shards", [1, 3])
def test_generator_batched_plugin_submits_one_task_per_shard(num_shards):
    from gtsfm_b200.frontend.correspondence_generator.det_desc_correspondence_generator import DetDescCorrespondenceGenerator
    kps, descs = _stub_features()
    m = BatchStubMatcher()
    gen = DetDescCorrespondenceGenerator(m, FixtureDetDesc(kps, descs), num_shards=num_shards)
    pairs = [(3, 4), (0, 1), (0, 2), (5, 6), (1, 2), (2, 6)]              # not sorted: the result keeps THIS order
    client = FakeClient()
    keypoints_list, corr = gen.generate_correspondences(client, [FakeImage(i) for i in range(7)], pairs)
    assert list(corr.keys()) == pairs
    for (i, j), v in corr.items():
        assert (v == [[10 + i, 10 + j]]).all()
    assert client.count("_apply_matcher_shard") == num_shards == len(m.batches)
    assert client.count("_apply_matcher") == 0 and m.calls == 0
    assert sum(len(b[1]) for b in m.batches) == len(pairs)
    for n_img, local, device in m.batches:                               # a shard only receives the images its pairs name
        assert n_img == len(np.unique(local)) and local.max() == n_img - 1


class StubEstimator:
    def __init__(self):
        self.batches = []

    def run_2view_batch(self, keypoints, putative, intrinsics, priors, gt_cameras, gt_scene_mesh, device=None):
        self.batches.append((sorted(keypoints), list(putative), device))
        assert set(np.unique(np.array(list(putative)))) == set(keypoints) == set(intrinsics)
        return {k: (None, None, np.array([]), k, priors.get(k), gt_cameras[k[0]] if gt_cameras else None) for k in putative}


@pytest.mark.parametrize("use_client,num_shards", [(True, 1), (True, 4), (False, 2)])
def test_run_two_view_estimator_as_futures_shards_and_keeps_the_dictionary_order(use_client, num_shards):
    from gtsfm_b200.two_view_estimator import run_two_view_estimator_as_futures
    kps, _ = _stub_features(9)
    keys = [(4, 7), (0, 1), (0, 3), (2, 8), (1, 2), (5, 6), (3, 4)]
    putative = {k: np.zeros((5 + k[0], 2), np.uint32) for k in keys}
    est = StubEstimator()
    client = FakeClient() if use_client else None
    priors = {(0, 3): "prior03"}
    gt = [f"cam{i}" for i in range(9)]
    out = run_two_view_estimator_as_futures(client, est, kps, putative, list(range(9)), priors, gt, None, num_shards=num_shards)
    assert list(out.keys()) == keys
    assert out[(0, 3)][4] == "prior03" and out[(4, 7)][4] is None and out[(2, 8)][5] == "cam2"
    assert len(est.batches) == num_shards
    if use_client:
        assert client.count("_apply_two_view_estimator_shard") == num_shards and client.scattered == [est]
    assert sorted(k for b in est.batches for k in b[1]) == sorted(keys)


def test_unsupported_stages_fail_loudly():
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    from gtsfm_b200.two_view_estimator import TwoViewEstimator
    kps, _ = _stub_features(2)
    put = {(0, 1): np.zeros((8, 2), np.uint32)}
    isp = InlierSupportProcessor(15, 0.1)
    with pytest.raises(NotImplementedError):
        TwoViewEstimator(Ransac(True, 4.0), isp, True, 4.0).run_2view_batch(kps, put, [None, None])
    from gtsfm_b200.data_association.point3d_initializer import TriangulationOptions, TriangulationSamplingMode
    with pytest.raises(NotImplementedError):                 # BA triangulates with NO_RANSAC only
        TwoViewEstimator(Ransac(True, 4.0), isp, True, 4.0, TriangulationOptions(TriangulationSamplingMode.RANSAC_SAMPLE_UNIFORM)) \
            .run_2view_batch(kps, put, [None, None])
    with pytest.raises(NotImplementedError):                 # BA with a relative pose prior
        TwoViewEstimator(Ransac(True, 4.0), isp, True, 4.0, TriangulationOptions(TriangulationSamplingMode.NO_RANSAC)) \
            .run_2view_batch(kps, put, [None, None], {(0, 1): object()})
    with pytest.raises(NotImplementedError):
        TwoViewEstimator(Ransac(True, 4.0), isp, False, 4.0).run_2view_batch(kps, put, [None, None], gt_scene_mesh=object())
    with pytest.raises(NotImplementedError):
        TwoViewEstimator(PerPairMatcher(), isp, False, 4.0).run_2view_batch(kps, put, [None, None])


# ------------------------------------------------------------------------------------------------------------ GPU: C1
class Pose:
    def __init__(self, R, t):
        self._R, self._t = R, t

    def rotation(self):
        return self

    def matrix(self):
        return self._R

    def translation(self):
        return self._t


class GtCamera:
    def __init__(self, wRc, wtc, cal):
        self._pose, self._cal = Pose(wRc, wtc), cal

    def pose(self):
        return self._pose

    def calibration(self):
        return self._cal


@pytest.fixture(scope="module")
def c1_data():
    from gtsfm_b200.common.types import Cal3Bundler, Keypoints
    g = np.load(os.path.join(GOLDEN, "c1_lund_door_12.npz"))
    off = g["img_off"]
    kps = [Keypoints(coordinates=g["kps"][off[i]:off[i + 1]].astype(np.float64)) for i in range(12)]
    descs = [g["desc_u8"][off[i]:off[i + 1]].astype(np.float32) for i in range(12)]
    cals = [Cal3Bundler(*[float(v) for v in g["cal"][i]]) for i in range(12)]
    cams = [GtCamera(g["wRc"][i], g["wtc"][i], cals[i]) for i in range(12)]
    return g, kps, descs, cals, cams


@pytest.mark.gpu
def test_c1_generate_correspondences_equals_the_reference_matches(c1_data):
    from gtsfm_b200.frontend.correspondence_generator.det_desc_correspondence_generator import DetDescCorrespondenceGenerator
    from gtsfm_b200.frontend.matcher.twoway_matcher import TwoWayMatcher
    g, kps, descs, cals, cams = c1_data
    pairs = [(int(a), int(b)) for a, b in g["pairs"]]
    mo = g["m_off"]
    client = FakeClient()
    gen = DetDescCorrespondenceGenerator(TwoWayMatcher(ratio_test_threshold=0.8), FixtureDetDesc(kps, descs), num_shards=3)
    keypoints_list, corr = gen.generate_correspondences(client, [FakeImage(i) for i in range(12)], pairs)
    assert client.count("_apply_matcher_shard") == 3 and list(corr.keys()) == pairs and len(keypoints_list) == 12
    for k, key in enumerate(pairs):
        assert corr[key].dtype == np.uint32
        np.testing.assert_array_equal(corr[key], g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32))
    # equals the per-pair plugin call of the reference's loop
    m = TwoWayMatcher(ratio_test_threshold=0.8)
    for key in (pairs[0], pairs[40]):
        np.testing.assert_array_equal(corr[key], m.match(kps[key[0]], kps[key[1]], descs[key[0]], descs[key[1]], (1, 1, 3), (1, 1, 3)))
    # no client: same dictionary
    _, corr2 = DetDescCorrespondenceGenerator(TwoWayMatcher(ratio_test_threshold=0.8), FixtureDetDesc(kps, descs), num_shards=1) \
        .generate_correspondences(None, [FakeImage(i) for i in range(12)], pairs[:5])
    for key in pairs[:5]:
        np.testing.assert_array_equal(corr2[key], corr[key])


@pytest.mark.gpu
def test_c1_run_two_view_estimator_as_futures_equals_per_pair_and_oracle(c1_data):
    from gtsfm_b200.frontend.inlier_support_processor import InlierSupportProcessor
    from gtsfm_b200.frontend.verifier.ransac import Ransac
    from gtsfm_b200.two_view_estimator import TwoViewEstimator, run_two_view_estimator_as_futures, unzip_two_view_results
    from oracle import oracle as O
    g, kps, descs, cals, cams = c1_data
    pairs = [(int(a), int(b)) for a, b in g["pairs"]]
    mo = g["m_off"]
    putative = {key: g["m_u16"][mo[k]:mo[k + 1]].astype(np.uint32) for k, key in enumerate(pairs)}
    putative[pairs[7]] = putative[pairs[7]][:4]                         # too few matches -> failure tuple
    putative[pairs[9]] = np.array([])                                   # the matcher's "no matches" value
    est = TwoViewEstimator(Ransac(True, 4.0, num_hypotheses=512, seed=3), InlierSupportProcessor(15, 0.1), False, 4.0)
    gt = list(cams)
    gt[11] = None                                                        # an image without ground truth
    client = FakeClient()
    out = run_two_view_estimator_as_futures(client, est, kps, putative, cals, {}, gt, None, num_shards=4)
    assert client.count("_apply_two_view_estimator_shard") == 4 and list(out.keys()) == pairs
    one = run_two_view_estimator_as_futures(None, est, kps, putative, cals, {}, gt, None, num_shards=1)
    for k, key in enumerate(pairs):
        R, U, v, pre, post_ba, post_isp = out[key]
        R1, U1, v1, pre1, _, _ = one[key]
        assert (R is None) == (R1 is None)
        np.testing.assert_array_equal(v, v1)                            # sharding does not change a pair's result
        if key in (pairs[7], pairs[9]):
            assert R is None and U is None and v.size == 0 and v.dtype == np.uint64 and pre.num_inliers_est_model == 0
            continue
        np.testing.assert_array_equal(R.matrix(), R1.matrix())
        assert v.dtype == np.uint32 and pre.num_inliers_est_model == len(v) and post_ba.num_inliers_est_model == len(v)
        assert abs(pre.inlier_ratio_est_model - len(v) / len(putative[key])) < 1e-12
        if 11 in key:
            assert pre.R_error_deg is None and pre.v_corr_idxs_inlier_mask_gt is None
            continue
        # the report against the numpy restatement of the reference functions
        i, j = key
        Rg = g["wRc"][j].T @ g["wRc"][i]
        tg = g["wRc"][j].T @ (g["wtc"][i] - g["wtc"][j])
        ref = O.two_view_report(kps[i].coordinates, kps[j].coordinates, v, len(putative[key]), R.matrix(), U.point3(), tuple(g["cal"][i]),
                                tuple(g["cal"][j]), Rg, tg, 4.0, 15, 0.1)
        assert abs(pre.R_error_deg - ref["R_error_deg"]) < 1e-6 and abs(pre.U_error_deg - ref["U_error_deg"]) < 1e-6
        band = np.abs(ref["dist_sq"] / 16.0 - 1.0) < 1e-9
        assert ((pre.v_corr_idxs_inlier_mask_gt == ref["gt_mask"]) | band).all()
        np.testing.assert_allclose(pre.reproj_error_gt_model, ref["dist_sq"], rtol=1e-9, atol=1e-12)   # px^2; cancellation in near-zero numerators
        assert abs(pre.num_inliers_gt_model - ref["num_inliers_gt"]) <= band.sum()
        assert ref["keep"] == (post_isp.num_inliers_est_model > 0 and R is not None)
    # per-pair entry point: a batch of one through the same code (same sampler key -> same result)
    key = pairs[15]
    single = est.run_2view(kps[key[0]], kps[key[1]], putative[key], cals[key[0]], cals[key[1]], None, gt[key[0]], gt[key[1]])
    assert single[0] is not None and abs(single[3].R_error_deg - out[key][3].R_error_deg) < 2.0
    i2Ri1, i2Ui1, v_corr, reports = unzip_two_view_results(out)
    assert pairs[7] not in i2Ri1 and pairs[9] not in i2Ri1 and len(i2Ri1) == len(reports) >= 60
