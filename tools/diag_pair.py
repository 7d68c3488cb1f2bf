"""Diagnostics: one pair of the ragged C3-shaped batch, GPU vs oracle for every local-optimisation budget."""
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
P = int(sys.argv[1]) if len(sys.argv) > 1 else 27
i, j = scene.pairs[P]
d1, d2 = scene.descriptors[i][:sizes[i]], scene.descriptors[j][:sizes[j]]
k1, k2 = scene.keypoints[i][:sizes[i]], scene.keypoints[j][:sizes[j]]
ctx = gtsfm_b200.Context(0)
m = ctx.match_pair(d1, d2, 0.8)
H, seed = 512, 20260003
samples = gtsfm_b200.ransac_samples(seed, P, len(m), H)
thr2 = (4.0 / scene.cal[0]) ** 2
n1 = O.calibrate(k1[m[:, 0]], scene.cal); n2 = O.calibrate(k2[m[:, 1]], scene.cal)
for lo in [0, 1, 2, 3, 4, 5, 6, 8, 10, 15, 20, 40]:
    g = ctx.verify_pair(k1, k2, m, scene.cal, scene.cal, params(4.0, H, lo=lo, samples=samples))
    ref = O.verify_pair(k1, k2, m, scene.cal, scene.cal, 4.0, samples, lo_iters=lo)
    Eg = g["E"] / np.linalg.norm(g["E"]); Er = ref["E"] / np.linalg.norm(ref["E"])
    dE = min(np.abs(Eg - Er).max(), np.abs(Eg + Er).max())
    cg = np.minimum(O.sampson_sq(g["E"], n1, n2) / thr2, 1).sum() * thr2
    cr = np.minimum(O.sampson_sq(ref["E"], n1, n2) / thr2, 1).sum() * thr2
    print(f"lo={lo:2d} dE={dE:.3e} ninl {g['n_inliers']} vs {ref['ninl']} cost gpu {cg:.10e} ref {cr:.10e} best_id {ref['best_id']} dR={rot_angle_deg(g['R'], ref['R']):.4f} dt={dir_angle_deg(g['t'], ref['t']):.4f}")
