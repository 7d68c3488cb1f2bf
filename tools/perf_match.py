"""Matcher-only throughput probe: N images x K keypoints of SIFT-like descriptors, P pairs, inputs resident."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gtsfm_b200

def sift_like(rng, n, dim=128):
    p = np.abs(rng.standard_normal((n, dim))).astype(np.float32)
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    p = np.minimum(p, 0.2); p /= np.linalg.norm(p, axis=1, keepdims=True)
    return np.clip(np.rint(p * 512), 0, 255).astype(np.float32)

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 24
kps = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
rng = np.random.default_rng(0)
c = gtsfm_b200.Context(0)
c.clear()
base = sift_like(rng, kps)
for i in range(n_img):
    d = base.copy()
    idx = rng.permutation(kps)[: kps // 2]
    d[idx] = sift_like(rng, kps // 2)
    c.upload_image(i, d[rng.permutation(kps)])
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
for r in range(reps):
    t0 = time.perf_counter()
    m, off = c.match_batch(pairs, 0.8)
    dt = time.perf_counter() - t0
    P = len(pairs)
    fl = 2.0 * kps * kps * 128 * P
    print(f"rep {r}: {P} pairs of {kps}x{kps}x128 in {dt*1e3:.1f} ms -> {P/dt:.0f} pairs/s, {dt/P*1e6:.1f} us/pair, {fl/dt/1e12:.1f} TFLOP/s algorithmic, matches/pair {m.shape[0]/P:.0f}, launches {c.kernel_launches}, flagged(last chunk) {c.debug_fetch(10,1,np.int32)[0]}")
