"""Generates tests/golden/*.npz by running the UNMODIFIED reference from /root/reference in the build
container (it cannot travel to the GPU box; the fixtures do).  ORACLE-side test infrastructure.

    python oracle/gen_golden.py

Needs cv2 + numpy; `gtsam` is replaced by oracle/gtsam_stub when the wheel is absent (it is absent in
this image).  Every array written here is an input or an output of reference code:

  matcher_*.npz       inputs + TwoWayMatcher(...).match outputs (gtsfm/frontend/matcher/twoway_matcher.py)
  lund_door_pair.npz  tests/data/set1_lund_door/features/* (descriptors as uint8: they are integer-valued)
                      + reference matches with ratio 0.8 / None
  verifier_*.npz      inputs + Ransac(True, thr).verify outputs (gtsfm/frontend/verifier/ransac.py), the
                      reference tests' scenes (tests/frontend/verifier/test_verifier_base.py:81-100,235-294;
                      tests/frontend/verifier/test_verifier_argoverse.py:73-119)
  sampson.npz         reference numpy Sampson^2 on random inputs (gtsfm/utils/verification.py:170-214)
"""
import os
import pickle
import sys

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
sys.path.insert(0, REF)
try:
    import gtsam  # noqa: F401
    GTSAM = "real"
except ImportError:
    sys.path.insert(0, os.path.join(HERE, "gtsam_stub"))
    GTSAM = "stub"

import cv2  # noqa: E402
from gtsam import Cal3Bundler  # noqa: E402

import gtsfm.utils.verification as verification_utils  # noqa: E402
from gtsfm.common.keypoints import Keypoints  # noqa: E402
from gtsfm.frontend.matcher.twoway_matcher import TwoWayMatcher  # noqa: E402
from gtsfm.frontend.verifier.ransac import Ransac  # noqa: E402


def ref_match(d1, d2, ratio):
    kp = Keypoints(coordinates=np.zeros((0, 2)))
    return np.asarray(TwoWayMatcher(ratio_test_threshold=ratio).match(kp, kp, d1, d2, (1, 1), (1, 1)))


def sift_like(rng, protos, noise=0.03):
    """SIFT-128-like integer-valued float32 descriptors, norm ~512."""
    d = protos + rng.normal(0, noise, protos.shape)
    d = np.abs(d)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = np.minimum(d, 0.2)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.clip(np.rint(d * 512), 0, 255).astype(np.float32)


def gen_matcher(rng):
    cases = {}
    # (a) the reference's own known-answer vectors (tests/frontend/matcher/test_twoway_with*ratiotest_matcher.py:25-49)
    d1 = np.array([0.4865, 0.3752, 0.3077, 0.9188, 0.7837, 0.1083, 0.6822, 0.3764, 0.2288, 0.8018, 1.1], np.float32).reshape(-1, 1)
    d2 = np.array([0.9995, 0.3376, 0.9005, 0.5382, 0.3162, 0.7974, 0.1785, 0.3491, 0.8658, 0.2912], np.float32).reshape(-1, 1)
    cases["dummy1d"] = (d1, d2)
    # (b) SIFT-like with shared points and clutter
    protos = np.abs(rng.normal(0, 1, (700, 128)))
    protos /= np.linalg.norm(protos, axis=1, keepdims=True)
    a = sift_like(rng, protos[rng.permutation(700)[:500]])
    b = sift_like(rng, protos[rng.permutation(700)[:450]])
    cases["sift500"] = (a, b)
    # (c) heavy ties: tiny alphabet so many distances coincide (exercises lowest-index and stable-sort rules)
    a = rng.integers(0, 3, (300, 16)).astype(np.float32)
    b = rng.integers(0, 3, (257, 16)).astype(np.float32)
    b[:40] = a[rng.permutation(300)[:40]]
    cases["ties"] = (a, b)
    # (d) SuperPoint-256-like unit-norm floats
    pr = rng.normal(0, 1, (400, 256))
    a = pr[rng.permutation(400)[:333]] + rng.normal(0, 0.1, (333, 256))
    b = pr[rng.permutation(400)[:301]] + rng.normal(0, 0.1, (301, 256))
    a /= np.linalg.norm(a, axis=1, keepdims=True)
    b /= np.linalg.norm(b, axis=1, keepdims=True)
    cases["float256"] = (a.astype(np.float32), b.astype(np.float32))
    # (e) dimension not a multiple of 16 (scalar tail of the cv2 loop) + NaN rows (twoway_matcher.py:75-76)
    a = rng.normal(0, 1, (90, 37)).astype(np.float32)
    b = rng.normal(0, 1, (70, 37)).astype(np.float32)
    b[:30] = a[10:40] + rng.normal(0, 0.05, (30, 37)).astype(np.float32)
    a[5, 3] = np.nan
    a[17, :] = np.nan
    b[2, 36] = np.nan
    cases["nan37"] = (a, b)
    # (f) ragged: a single query, and exactly two train rows
    cases["ragged_1xN"] = (rng.normal(0, 1, (1, 8)).astype(np.float32), rng.normal(0, 1, (9, 8)).astype(np.float32))
    cases["ragged_Nx2"] = (rng.normal(0, 1, (7, 8)).astype(np.float32), rng.normal(0, 1, (2, 8)).astype(np.float32))
    for name, (d1, d2) in cases.items():
        out = {"d1": d1, "d2": d2}
        for tag, ratio in (("none", None), ("r08", 0.8), ("r095", 0.95)):
            try:
                m = ref_match(d1, d2, ratio)
            except ValueError:  # twoway_matcher.py:137 cannot unpack knnMatch rows when a side has < 2 descriptors
                out["raises_" + tag] = np.array(1)
                continue
            out["m_" + tag] = m.reshape(-1, 2).astype(np.uint32) if m.size else np.zeros((0, 2), np.uint32)
        np.savez_compressed(os.path.join(OUT, f"matcher_{name}.npz"), **out)
        print("matcher", name, {k: v.shape for k, v in out.items()})


def gen_lund():
    root = os.path.join(REF, "tests", "data", "set1_lund_door", "features")
    d0 = np.load(os.path.join(root, "descriptors_0.npy"))
    d1 = np.load(os.path.join(root, "descriptors_1.npy"))
    assert np.all(d0 == np.rint(d0)) and np.all(d1 == np.rint(d1)) and d0.max() <= 255 and d1.max() <= 255
    with open(os.path.join(root, "keypoints_0.pkl"), "rb") as f:
        k0 = pickle.load(f)
    with open(os.path.join(root, "keypoints_1.pkl"), "rb") as f:
        k1 = pickle.load(f)
    out = {"d0_u8": d0.astype(np.uint8), "d1_u8": d1.astype(np.uint8),
           "kp0": np.asarray(k0.coordinates, np.float32), "kp1": np.asarray(k1.coordinates, np.float32)}
    for tag, ratio in (("none", None), ("r08", 0.8)):
        out["m_" + tag] = ref_match(d0, d1, ratio).astype(np.uint32)
    # reference verifier on the reference matches (black-box USAC_ACCURATE; recorded, compared statistically).
    # Intrinsics: Olsson loader builds f from data.mat; here a plausible Cal3Bundler for 1296x1936 images.
    cal = Cal3Bundler(2398.1, 0.0, 0.0, 628.3, 932.4)
    R, U, v, ratio = Ransac(True, 4.0).verify(k0, k1, out["m_r08"], cal, cal)
    out["ver_cal"] = np.array([2398.1, 0.0, 0.0, 628.3, 932.4])
    out["ver_R"] = R.matrix()
    out["ver_t"] = U.point3()
    out["ver_idx"] = v.astype(np.uint32)
    out["ver_ratio"] = np.float64(ratio)
    np.savez_compressed(os.path.join(OUT, "lund_door_pair.npz"), **out)
    print("lund", {k: getattr(v, "shape", v) for k, v in out.items()})


def gen_c1():
    """BASELINE config C1: the 12 Lund-door images through the reference's own front end (loader resize to
    --max_resolution 760: gtsfm/loader/loader_base.py:168-190, gtsfm/utils/images.py:101-201; SIFT top-5000:
    gtsfm/frontend/detector_descriptor/sift.py:44-54) and the unmodified TwoWayMatcher(ratio 0.8) on all 66 pairs
    (ExhaustiveRetriever).  Intrinsics / GT poses from data.mat as gtsfm/loader/olsson_loader.py:75-147 does
    (RQ decomposition, gtsfm/utils/verification.py:20-49), intrinsics rescaled as loader_base.py:214-224."""
    import glob
    import hashlib
    import scipy.io
    import scipy.linalg
    from PIL import Image as PILImage
    import gtsfm.utils.images as img_utils
    from gtsfm.common.image import Image
    from gtsfm.frontend.detector_descriptor.sift import SIFTDetectorDescriptor
    root = os.path.join(REF, "tests", "data", "set1_lund_door")
    paths = sorted(glob.glob(os.path.join(root, "images", "*.JPG")))
    P = scipy.io.loadmat(os.path.join(root, "data.mat"))["P"][0]
    assert len(paths) == len(P) == 12
    det = SIFTDetectorDescriptor(max_keypoints=5000)
    descs, kps, cals, wRc, wtc, counts = [], [], [], [], [], []
    for path, M in zip(paths, P):
        pil = PILImage.open(path)
        pil = pil.convert("RGB") if pil.mode != "RGB" else pil
        img = Image(value_array=np.asarray(pil), file_name=os.path.basename(path))
        su, sv, th, tw = img_utils.get_downsampling_factor_per_axis(img.height, img.width, 760)
        if min(img.height, img.width) > 760:
            img = img_utils.resize_image(img, new_height=th, new_width=tw)
        k, d = det.detect_and_describe(img)
        assert np.all(d == np.rint(d)) and d.max() <= 255 and d.min() >= 0
        Q, m4 = M[:3, :3], M[:, 3]
        K, cRw = scipy.linalg.rq(Q)
        T = np.diag(np.sign(np.diag(K)))
        K, cRw = K @ T, T @ cRw
        if not cals:                       # olsson_loader.py:87,135-147: every image gets the FIRST camera's K
            K0 = K / K[2, 2]
        f = min(K0[0, 0], K0[1, 1])
        cals.append([f * su, 0.0, 0.0, K0[0, 2] * su, K0[1, 2] * sv])
        wRc.append(cRw.T)
        wtc.append(np.linalg.inv(-Q) @ m4)
        counts.append(len(k))
        descs.append(d.astype(np.uint8))
        kps.append(np.asarray(k.coordinates, np.float32))
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    pairs = [(i, j) for i in range(12) for j in range(i + 1, 12)]
    m_all, m_off, sha = [], [0], []
    for i, j in pairs:
        m = ref_match(descs[i].astype(np.float32), descs[j].astype(np.float32), 0.8).reshape(-1, 2).astype(np.uint32)
        sha.append(hashlib.sha1(np.ascontiguousarray(m).tobytes()).hexdigest())
        m_all.append(m.astype(np.uint16))
        m_off.append(m_off[-1] + len(m))
    out = {"desc_u8": np.concatenate(descs), "kps": np.concatenate(kps), "img_off": off, "cal": np.array(cals),
           "wRc": np.array(wRc), "wtc": np.array(wtc), "pairs": np.array(pairs, np.int32),
           "m_u16": np.concatenate(m_all), "m_off": np.array(m_off, np.int64), "m_sha1": np.array(sha)}
    np.savez_compressed(os.path.join(OUT, "c1_lund_door_12.npz"), **out)
    print("c1", counts, "matches/pair", np.diff(m_off).min(), np.diff(m_off).max(),
          os.path.getsize(os.path.join(OUT, "c1_lund_door_12.npz")))


def gen_verifier(rng):
    sys.path.insert(0, REF)
    import tests.frontend.verifier.test_verifier_base as tvb
    np.random.seed(15)
    # (a) two-plane scene of the reference test: expects all 8 matches verified, R,t within 2 deg of GT
    k1, k2, E = tvb.simulate_two_planes_scene(4, 4)
    mi = np.vstack((np.arange(8), np.arange(8))).T
    R, U, v, ratio = Ransac(True, 0.5).verify(k1, k2, mi, Cal3Bundler(), Cal3Bundler())
    np.savez_compressed(os.path.join(OUT, "verifier_two_planes.npz"), kp1=k1.coordinates, kp2=k2.coordinates,
                        R_gt=E.rotation().matrix(), t_gt=E.direction().point3(), ref_R=R.matrix(), ref_t=U.point3(),
                        ref_idx=v, ref_ratio=np.float64(ratio))
    print("two planes: ref verified", v.shape, "ratio", ratio)
    # (b) larger two-plane scene for recover_relative_pose (tests/utils/test_verification_utils.py:17-29)
    k1, k2, E = tvb.simulate_two_planes_scene(10, 10)
    R, U = verification_utils.recover_relative_pose_from_essential_matrix(E.matrix(), k1.coordinates, k2.coordinates, Cal3Bundler(), Cal3Bundler())
    np.savez_compressed(os.path.join(OUT, "verifier_recover_pose.npz"), kp1=k1.coordinates, kp2=k2.coordinates, E=E.matrix(),
                        R_gt=E.rotation().matrix(), t_gt=E.direction().point3(), ref_R=R.matrix(), ref_t=U.point3())
    # (c) Argoverse annotated pair (tests/frontend/verifier/test_verifier_argoverse.py)
    with open(os.path.join(REF, "tests/data/argoverse/labeled_correspondences/argoverse_315975640448534784__315975643412234000.pkl"), "rb") as f:
        d = pickle.load(f)
    uv1 = np.c_[d["x1"], d["y1"]].astype(np.float32)
    uv2 = np.c_[d["x2"], d["y2"]].astype(np.float32)
    cal = (1392.1069298937407, 0.0, 0.0, 980.1759848618066, 604.3534182680304)
    mi = np.vstack([np.arange(len(uv1)), np.arange(len(uv1))]).T
    np.random.seed(0)
    R, U, v, ratio = Ransac(True, 0.5).verify(Keypoints(uv1), Keypoints(uv2), mi, Cal3Bundler(*cal), Cal3Bundler(*cal))
    np.savez_compressed(os.path.join(OUT, "verifier_argoverse.npz"), kp1=uv1, kp2=uv2, cal=np.array(cal), ref_R=R.matrix(),
                        ref_t=U.point3(), ref_idx=v, ref_ratio=np.float64(ratio),
                        gt_euler_zyx_deg=np.array([-0.37, 32.47, -0.42]), gt_i1ti2=np.array([0.21, -0.0024, 0.976]))
    print("argoverse: ref verified", v.shape, "ratio", ratio)
    # (d) synthetic scene with outliers; reference (black-box USAC_ACCURATE) outputs recorded for the statistical comparison
    n, f = 2000, 1000.0
    X = rng.uniform(-2, 2, (n, 3)) + [0, 0, 6]
    Rgt, _ = cv2.Rodrigues(np.array([0.05, -0.2, 0.03]))
    tgt = np.array([1.0, 0.1, 0.2]); tgt /= np.linalg.norm(tgt)
    x1 = X[:, :2] / X[:, 2:]
    Xc = X @ Rgt.T + tgt
    x2 = Xc[:, :2] / Xc[:, 2:]
    kp1 = x1 * f + [640, 480] + rng.normal(0, 0.5, (n, 2))
    kp2 = x2 * f + [640, 480] + rng.normal(0, 0.5, (n, 2))
    kp2[:900] = rng.uniform(0, 1, (900, 2)) * [1280, 960]
    perm = rng.permutation(n)
    mi = np.c_[perm, perm].astype(np.uint32)
    cal = (f, 0.0, 0.0, 640.0, 480.0)
    R, U, v, ratio = Ransac(True, 4.0).verify(Keypoints(kp1), Keypoints(kp2), mi, Cal3Bundler(*cal), Cal3Bundler(*cal))
    np.savez_compressed(os.path.join(OUT, "verifier_synth2000.npz"), kp1=kp1, kp2=kp2, mi=mi, cal=np.array(cal), R_gt=Rgt, t_gt=tgt,
                        ref_R=R.matrix(), ref_t=U.point3(), ref_idx=v.astype(np.uint32), ref_ratio=np.float64(ratio))
    print("synth2000: ref verified", v.shape, "ratio", ratio)


def gen_sampson(rng):
    c1 = rng.uniform(-1, 1, (64, 2))
    c2 = rng.uniform(-1, 1, (64, 2))
    F = rng.normal(0, 1, (3, 3))
    s = verification_utils.compute_epipolar_distances_sq_sampson(c1, c2, F)
    np.savez_compressed(os.path.join(OUT, "sampson.npz"), c1=c1, c2=c2, F=F, ref=s)


def gen_fmat():
    """Fundamental-matrix mode (ransac.py:84-111): cv2's 7-point solver on minimal sets, and the UNMODIFIED reference
    Ransac(use_intrinsics_in_verification=False).verify on a synthetic pair.  Own seed: `python gen_golden.py fmat`
    adds these files without touching the others."""
    rng = np.random.default_rng(20260007)
    n, f = 1500, 1000.0
    X = rng.uniform([-4, -4, 6], [4, 4, 12], (n, 3))
    Rgt = cv2.Rodrigues(np.array([0.03, 0.25, -0.02]))[0]
    tgt = np.array([1.0, 0.15, 0.25])
    K = np.array([[f, 0, 640.0], [0, f, 480.0], [0, 0, 1]])
    p1 = (K @ X.T).T
    p2 = (K @ ((Rgt @ X.T).T + tgt).T).T
    kp1 = p1[:, :2] / p1[:, 2:] + rng.normal(0, 0.6, (n, 2))
    kp2 = p2[:, :2] / p2[:, 2:] + rng.normal(0, 0.6, (n, 2))
    kp2[:500] = rng.uniform(0, 1, (500, 2)) * [1280, 960]
    perm = rng.permutation(n)
    mi = np.c_[perm, perm].astype(np.uint32)
    cal = (f, 0.0, 0.0, 640.0, 480.0)
    # minimal sets drawn from the true correspondences
    sets = np.stack([500 + rng.choice(n - 500, 7, replace=False) for _ in range(64)])
    F7 = np.full((64, 3, 3, 3), np.nan)
    n7 = np.zeros(64, np.int32)
    for k, idx in enumerate(sets):
        Fs, _ = cv2.findFundamentalMat(kp1[idx].astype(np.float32), kp2[idx].astype(np.float32), cv2.FM_7POINT)
        if Fs is not None:
            n7[k] = Fs.shape[0] // 3
            F7[k, :n7[k]] = Fs.reshape(-1, 3, 3)
    # cv2's own model and mask on all matches (the error / mask rule is pinned on these)
    Fcv, mcv = cv2.findFundamentalMat(kp1[mi[:, 0]], kp2[mi[:, 1]], method=cv2.FM_RANSAC, ransacReprojThreshold=4.0,
                                      confidence=0.999999, maxIters=1000000)
    R, U, v, ratio = Ransac(False, 4.0).verify(Keypoints(kp1), Keypoints(kp2), mi, Cal3Bundler(*cal), Cal3Bundler(*cal))
    np.savez_compressed(os.path.join(OUT, "fmat_synth1500.npz"), kp1=kp1, kp2=kp2, mi=mi, cal=np.array(cal), R_gt=Rgt, t_gt=tgt,
                        sets=sets.astype(np.int32), F7=F7, n7=n7, cv_F=Fcv, cv_mask=mcv.ravel().astype(np.uint8),
                        ref_R=R.matrix(), ref_t=U.point3(), ref_idx=v.astype(np.uint32), ref_ratio=np.float64(ratio))
    print("fmat_synth1500: ref verified", v.shape, "ratio", ratio, "7pt solutions", n7.sum())


def gen_hamming():
    """TwoWayMatcher(distance_type=HAMMING) of the unmodified reference on ORB-like 256-bit descriptors (uint8 rows)."""
    from gtsfm.frontend.matcher.twoway_matcher import MatchingDistanceType
    rng = np.random.default_rng(20260006)
    kp = Keypoints(coordinates=np.zeros((0, 2)))
    base = rng.integers(0, 256, (300, 32), dtype=np.uint8)

    def noisy(d, flips):
        d = np.unpackbits(d, axis=1)
        for r in range(d.shape[0]):
            d[r, rng.choice(256, flips, replace=False)] ^= 1
        return np.packbits(d, axis=1)
    d1 = np.concatenate([noisy(base, 20), rng.integers(0, 256, (150, 32), dtype=np.uint8), base[:25]])
    d2 = np.concatenate([noisy(base[::-1], 25), rng.integers(0, 256, (100, 32), dtype=np.uint8), base[:25]])
    out = dict(d1=d1, d2=d2)
    for tag, ratio in (("none", None), ("r08", 0.8), ("r095", 0.95)):
        m = TwoWayMatcher(distance_type=MatchingDistanceType.HAMMING, ratio_test_threshold=ratio).match(kp, kp, d1, d2, (1, 1), (1, 1))
        out["m_" + tag] = np.asarray(m).reshape(-1, 2).astype(np.uint32)
    np.savez_compressed(os.path.join(OUT, "hamming_orb256.npz"), **out)
    print("hamming:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if sys.argv[1:] == ["hamming"]:
        gen_hamming()
        with open(os.path.join(OUT, "PROVENANCE.txt"), "a") as fh:
            fh.write(f"hamming_*.npz: oracle/gen_golden.py hamming, /root/reference (commit bc04d78b), cv2 {cv2.__version__}, "
                     f"numpy {np.__version__}, gtsam={GTSAM}\n")
        sys.exit(0)
    if sys.argv[1:] == ["c1"]:
        gen_c1()
        with open(os.path.join(OUT, "PROVENANCE.txt"), "a") as fh:
            fh.write(f"c1_lund_door_12.npz: oracle/gen_golden.py c1, /root/reference (commit bc04d78b), cv2 {cv2.__version__}, "
                     f"numpy {np.__version__}, gtsam={GTSAM}\n")
        sys.exit(0)
    if sys.argv[1:] == ["fmat"]:
        gen_fmat()
        with open(os.path.join(OUT, "PROVENANCE.txt"), "a") as fh:
            fh.write(f"fmat_*.npz: oracle/gen_golden.py fmat, /root/reference (commit bc04d78b), cv2 {cv2.__version__}, "
                     f"numpy {np.__version__}, gtsam={GTSAM}\n")
        sys.exit(0)
    rng = np.random.default_rng(20260001)
    print("gtsam:", GTSAM, "cv2:", cv2.__version__)
    gen_matcher(rng)
    gen_lund()
    gen_verifier(rng)
    gen_sampson(rng)
    gen_fmat()
    gen_hamming()
    gen_c1()
    with open(os.path.join(OUT, "PROVENANCE.txt"), "w") as fh:
        fh.write(f"generated by oracle/gen_golden.py from /root/reference (commit bc04d78b), cv2 {cv2.__version__}, "
                 f"numpy {np.__version__}, gtsam={GTSAM}\n")
