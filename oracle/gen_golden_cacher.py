"""Golden vectors for the cacher key scheme, recorded from the UNMODIFIED reference cachers
(gtsfm/frontend/cacher/matcher_cacher.py:46-80, gtsfm/two_view_estimator_cacher.py:49-63) imported from /root/reference.
ORACLE-side test infrastructure; run in the build container:

    python oracle/gen_golden_cacher.py        -> tests/golden/cacher_keys.npz

The reference cachers import modules that are absent from this image (gtsam, h5py, open3d, dask, matplotlib, ...).  None
of them is touched by the key / file code, so they are replaced by MagicMock placeholders here - in this generator only.
The fixture holds inputs, the cache keys the reference computes for them, and the bytes of the .pbz2 files it writes.
"""
import importlib
import os
import sys
import tempfile
from pathlib import Path
from unittest.mock import MagicMock

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
sys.path.insert(0, REF)


def import_with_placeholders(name):
    stubbed = []
    for _ in range(100):
        try:
            return importlib.import_module(name), stubbed
        except ModuleNotFoundError as e:
            sys.modules[e.name] = MagicMock()
            stubbed.append(e.name)
    raise RuntimeError("too many missing modules")


def main():
    mc, s1 = import_with_placeholders("gtsfm.frontend.cacher.matcher_cacher")
    tc, s2 = import_with_placeholders("gtsfm.two_view_estimator_cacher")
    from gtsfm.common.keypoints import Keypoints
    print("placeholders:", s1 + s2)
    rng = np.random.default_rng(20260033)
    out = {}

    class TwoWayMatcher:                       # only type(obj).__name__ enters the key (matcher_cacher.py:34)
        def __init__(self, result):
            self.result = result

        def match(self, **kw):
            return self.result

    cases = []
    for n1, n2, dim, dtype, extras in ((25, 40, 128, np.float32, True), (7, 12, 256, np.float32, False), (30, 30, 32, np.uint8, True)):
        c1 = rng.uniform(0, 1000, (n1, 2))
        c2 = rng.uniform(0, 1000, (n2, 2)).astype(np.float32)          # mixed coordinate dtypes promote in the concatenate
        d1 = (rng.uniform(0, 255, (n1, dim))).astype(dtype)
        d2 = (rng.uniform(0, 255, (n2, dim))).astype(dtype)
        s1_ = rng.uniform(1, 8, n1).astype(np.float32) if extras else None
        r1_ = rng.uniform(0, 1, n1).astype(np.float32) if extras else None
        s2_ = rng.uniform(1, 8, n2).astype(np.float32) if extras else None
        r2_ = rng.uniform(0, 1, n2).astype(np.float32) if extras else None
        cases.append((c1, c2, d1, d2, s1_, r1_, s2_, r2_, (480, 640, 3), (760, 1135, 3)))
    with tempfile.TemporaryDirectory() as tmp:
        mc.CACHE_ROOT_PATH = Path(tmp)
        tc.CACHE_ROOT_PATH = Path(tmp)
        for k, (c1, c2, d1, d2, s1_, r1_, s2_, r2_, sh1, sh2) in enumerate(cases):
            result = rng.integers(0, 7, (9, 2)).astype(np.uint32)
            cacher = mc.MatcherCacher(matcher_obj=TwoWayMatcher(result))
            kp1 = Keypoints(coordinates=c1, scales=s1_, responses=r1_)
            kp2 = Keypoints(coordinates=c2, scales=s2_, responses=r2_)
            before = set(os.listdir(os.path.join(tmp, "matcher"))) if os.path.isdir(os.path.join(tmp, "matcher")) else set()
            got = cacher.match(keypoints_i1=kp1, keypoints_i2=kp2, descriptors_i1=d1, descriptors_i2=d2, im_shape_i1=sh1, im_shape_i2=sh2)
            assert got is result
            new = sorted(set(os.listdir(os.path.join(tmp, "matcher"))) - before)
            assert len(new) == 1
            with open(os.path.join(tmp, "matcher", new[0]), "rb") as f:
                blob = np.frombuffer(f.read(), np.uint8)
            p = f"m{k}_"
            out.update({p + "c1": c1, p + "c2": c2, p + "d1": d1, p + "d2": d2, p + "sh": np.array(sh1 + sh2), p + "result": result,
                        p + "file": np.array(new[0]), p + "blob": blob, p + "extras": np.array(s1_ is not None)})
            if s1_ is not None:
                out.update({p + "s1": s1_, p + "r1": r1_, p + "s2": s2_, p + "r2": r2_})
        # two-view estimator cacher: key only (its payload is the reference's 6-tuple of gtsam objects and reports)
        tve = tc.TwoViewEstimatorCacher(two_view_estimator_obj=None)
        keyfn = getattr(tve, "_TwoViewEstimatorCacher__generate_cache_key")
        for k, (n1, n2, m, idt) in enumerate(((50, 60, 30, np.uint32), (20, 20, 4, np.uint32), (300, 200, 120, np.int64))):
            c1 = rng.uniform(0, 1000, (n1, 2))
            c2 = rng.uniform(0, 1000, (n2, 2))
            idx = np.stack([rng.integers(0, n1, m), rng.integers(0, n2, m)], 1).astype(idt)
            key = keyfn(Keypoints(coordinates=c1), Keypoints(coordinates=c2), idx)
            p = f"t{k}_"
            out.update({p + "c1": c1, p + "c2": c2, p + "idx": idx, p + "key": np.array(key)})
    np.savez_compressed(os.path.join(OUT, "cacher_keys.npz"), **out)
    with open(os.path.join(OUT, "PROVENANCE.txt"), "a") as fh:
        fh.write("cacher_keys.npz: oracle/gen_golden_cacher.py, /root/reference (commit bc04d78b), unmodified MatcherCacher / "
                 f"TwoViewEstimatorCacher with placeholder modules for {sorted(set(s1 + s2))}, numpy {np.__version__}, python {sys.version.split()[0]}\n")
    print({k: (str(v) if v.ndim == 0 else v.shape) for k, v in out.items() if k.endswith(("file", "key"))})


if __name__ == "__main__":
    main()
