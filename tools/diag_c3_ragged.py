"""Diagnostics: GPU verifier vs oracle on the ragged C3-shaped batch of tests/test_gpu_baseline_sizes.py."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gtsfm_b200
from gtsfm_b200 import synthetic
from oracle import oracle as O
from tests.test_gpu_verify import params, rot_angle_deg, dir_angle_deg

scene = synthetic.make_scene(n_images=18, n_kps=8192, pairs="topk", top_k=3, seed=20260003, visible=0.6)
rng = np.random.default_rng(33)
sizes = rng.integers(3000, 8193, 18)
sizes[3], sizes[11] = 8192, 3001
descs = [d[:n] for d, n in zip(scene.descriptors, sizes)]
kps = [k[:n] for k, n in zip(scene.keypoints, sizes)]
ctx = gtsfm_b200.Context(0)
ws = int(sys.argv[1]) if len(sys.argv) > 1 else (160 << 20)
ctx.set_workspace(ws)
for i, (d, k) in enumerate(zip(descs, kps)):
    ctx.upload_image(i, d, k, scene.cal)
pairs = np.ascontiguousarray(scene.pairs, np.int32)
H, seed, lo = 512, 20260003, 20
m, off = ctx.match_batch(pairs, 0.8)
out = ctx.verify_batch(pairs, params(4.0, H, lo=lo, seed=seed), total=len(m))
thr2 = (4.0 / scene.cal[0]) ** 2
for p, (i, j) in enumerate(pairs):
    want = m[off[p]:off[p + 1]]
    samples = gtsfm_b200.ransac_samples(seed, p, len(want), H)
    ref = O.verify_pair(kps[i], kps[j], want, scene.cal, scene.cal, 4.0, samples, lo_iters=lo)
    n1 = O.calibrate(kps[i][want[:, 0]], scene.cal); n2 = O.calibrate(kps[j][want[:, 1]], scene.cal)
    s = O.sampson_sq(ref["E"], n1, n2) / thr2
    sg = O.sampson_sq(out["E"][p], n1, n2) / thr2
    mask = out["mask"][off[p]:off[p + 1]]
    differ = mask != ref["mask"]
    bad = differ & (np.abs(s - 1) > 1e-6)
    Eg = out["E"][p] / np.linalg.norm(out["E"][p]); Er = ref["E"] / np.linalg.norm(ref["E"])
    dE = min(np.abs(Eg - Er).max(), np.abs(Eg + Er).max())
    if bad.any() or dE > 1e-9:
        cg = np.minimum(sg, 1).sum() * thr2; cr = np.minimum(s, 1).sum() * thr2
        print(f"pair {p} ({i},{j}) M={len(want)} bad={bad.sum()} differ={differ.sum()} dE={dE:.3e} dR={rot_angle_deg(out['R'][p], ref['R']):.4f} "
              f"dt={dir_angle_deg(out['t'][p], ref['t']):.4f} ninl {out['n_inliers'][p]} vs {ref['ninl']} best_id {ref['best_id']} "
              f"cost ref {ref['cost']:.9e} (min {ref['cost_min']:.9e}) gpu-E cost {cg:.9e} ref-E cost {cr:.9e}")
        print("   s(ref) at bad:", s[bad][:5], " s(gpu) at bad:", sg[bad][:5])
print("done")
