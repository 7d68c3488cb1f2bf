"""How many rows / columns does the resolve step flag for the exact re-scan?  (C2-shaped pairs, one chunk.)"""
import sys
import numpy as np
sys.path.insert(0, ".")
import gtsfm_b200
from gtsfm_b200 import synthetic

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 12
scene = synthetic.make_scene(n_images=n_img, n_kps=8192, seed=20260002)
ctx = gtsfm_b200.Context(0)
for i in range(n_img):
    ctx.upload_image(i, scene.descriptors[i], scene.keypoints[i], scene.cal)
m, off = ctx.match_batch(scene.pairs, 0.8)
nf = ctx.debug_fetch(10, 1, np.int32)[0]
P = len(scene.pairs)
print(f"pairs {P}, flagged rows+cols {nf} = {nf / P:.1f} per pair = {nf / (P * 2 * 8192):.2e} of all rows/cols; matches/pair {off[-1] / P:.0f}")
