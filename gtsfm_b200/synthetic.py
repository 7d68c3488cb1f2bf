"""Synthetic two-view front-end workloads: random cameras observing random 3-D points, with
keypoint noise and clutter, projected to SIFT-like 128-d (integer-valued, norm ~512) or SuperPoint-like 256-d
(unit-norm float) descriptors.  Seeds are fixed per BASELINE.json config so GPU runs, the CPU baseline and the parity
tests see identical inputs.  Pure numpy: used by bench.py, tests and tools; not part of the device path.
"""
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

F, U0, V0, W, H = 1000.0, 640.0, 480.0, 1280.0, 960.0
CAL = (F, 0.0, 0.0, U0, V0)

# BASELINE.json configs: name -> (n_images, keypoints, dim, kind, visible fraction, pair rule, n_hypotheses, ratio, seed)
CONFIGS = {
    "C2": dict(n_images=100, n_kps=8192, dim=128, kind="sift", visible=0.6, pairs="exhaustive", top_k=0, n_hyp=1024, ratio=0.8, seed=20260002),
    "C3": dict(n_images=1000, n_kps=8192, dim=128, kind="sift", visible=0.6, pairs="topk", top_k=50, n_hyp=1024, ratio=0.8, seed=20260003),
    "C4": dict(n_images=2000, n_kps=2048, dim=256, kind="superpoint", visible=0.6, pairs="topk", top_k=50, n_hyp=1024, ratio=None, seed=20260004),
    "C5": dict(n_images=426, n_kps=20480, dim=128, kind="sift", visible=0.5, pairs="topk", top_k=50, n_hyp=10000, ratio=0.8, seed=20260005, max_pairs=20000, clutter_hidden=1.0),
}


@dataclass
class Scene:
    descriptors: List[np.ndarray]      # per image (n, dim) float32
    keypoints: List[np.ndarray]        # per image (n, 2) float64 pixels
    point_ids: List[np.ndarray]        # per image (n,) int64: 3-D point id, -1 for clutter
    R: List[np.ndarray]                # world -> camera rotations
    t: List[np.ndarray]                # world -> camera translations
    pairs: np.ndarray                  # (P, 2) int32, i < j
    cal: Tuple[float, float, float, float, float] = CAL
    ratio: Optional[float] = 0.8
    n_hyp: int = 1024
    threshold_px: float = 4.0

    def relative_pose(self, i: int, j: int):
        """i2Ri1, i2ti1 (unit) for pair (i1, i2) = (i, j)."""
        R = self.R[j] @ self.R[i].T
        t = self.t[j] - R @ self.t[i]
        return R, t / np.linalg.norm(t)


def _sift_quantise(p: np.ndarray) -> np.ndarray:
    p = np.abs(p)
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    return np.clip(np.rint(p * 512.0), 0, 255).astype(np.float32)


def sift_prototypes(rng, n: int, dim: int = 128) -> np.ndarray:
    p = np.abs(rng.standard_normal((n, dim), dtype=np.float32))
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    p = np.minimum(p, 0.2)
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    return p


def sift_observe(rng, protos: np.ndarray, noise: float = 0.03) -> np.ndarray:
    return _sift_quantise(protos + rng.normal(0.0, noise, protos.shape).astype(np.float32))


def superpoint_observe(rng, protos: np.ndarray, noise: float = 0.1) -> np.ndarray:
    p = protos + rng.normal(0.0, noise, protos.shape).astype(np.float32)
    return (p / np.linalg.norm(p, axis=1, keepdims=True)).astype(np.float32)


def _look_at_origin(rng, angle: float, radius: float = 10.0, jitter_deg: float = 5.0):
    c = np.array([radius * np.cos(angle), radius * np.sin(angle), rng.uniform(-0.5, 0.5)])
    z = -c / np.linalg.norm(c)
    x = np.cross(z, [0.0, 0.0, 1.0])
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z])
    w = np.radians(rng.normal(0.0, jitter_deg / 3.0, 3))
    th = np.linalg.norm(w)
    if th > 0:
        k = w / th
        K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        R = (np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K) @ R
    return R, -R @ c


def make_pairs(n_images: int, rule: str, top_k: int = 50, angles: Optional[np.ndarray] = None, max_pairs: Optional[int] = None) -> np.ndarray:
    """exhaustive: all i < j (retriever/exhaustive_retriever.py:15-19); topk: the top_k nearest later images by camera
    angle (mirrors top-K on the upper triangle, retriever/netvlad_retriever.py:164-170,213-228)."""
    if rule == "exhaustive":
        pairs = [(i, j) for i in range(n_images) for j in range(i + 1, n_images)]
    else:
        pairs = []
        for i in range(n_images - 1):
            later = np.arange(i + 1, n_images)
            d = np.abs(np.angle(np.exp(1j * (angles[later] - angles[i]))))
            for j in later[np.argsort(d, kind="stable")[:top_k]]:
                pairs.append((i, int(j)))
        pairs.sort()
    pairs = np.asarray(pairs, np.int32).reshape(-1, 2)
    return pairs[:max_pairs] if max_pairs else pairs


def make_scene(n_images: int, n_kps: int, dim: int = 128, kind: str = "sift", visible: float = 0.7, pairs: str = "exhaustive",
               top_k: int = 50, n_hyp: int = 1024, ratio: Optional[float] = 0.8, seed: int = 0, max_pairs: Optional[int] = None,
               noise_px: float = 0.5, arc: Optional[float] = None, clutter_hidden: float = 0.75,
               only_images: Optional[Sequence[int]] = None) -> Scene:
    """only_images: generate the features of these images only (a rank of a sharded job needs just the images its pairs
    name); every image draws from its own generator (seed, image index), so the subset equals the full scene's entries and the
    other entries are None.  Cameras and the pair list are always complete."""
    rng = np.random.default_rng(seed)
    n_pts = n_kps
    X = np.c_[rng.uniform(-4, 4, (n_pts, 2)), rng.uniform(-1, 1, n_pts)]
    if kind == "sift":
        protos = sift_prototypes(rng, n_pts, dim)
    else:
        protos = rng.standard_normal((n_pts, dim), dtype=np.float32)
    # cameras on an arc of the circle: neighbours see the scene from nearby directions
    arc = arc if arc is not None else min(2 * np.pi, 0.02 * n_images * np.pi)
    angles = np.sort(rng.uniform(0, arc, n_images))
    n_vis = int(round(visible * n_kps))
    descs, kps, ids, Rs, ts = [], [], [], [], []
    cams = [_look_at_origin(rng, angles[i]) for i in range(n_images)]
    wanted = None if only_images is None else set(int(i) for i in only_images)
    for i in range(n_images):
        R, t = cams[i]
        Rs.append(R)
        ts.append(t)
        if wanted is not None and i not in wanted:
            descs.append(None); kps.append(None); ids.append(None)
            continue
        rng = np.random.default_rng([seed, i])           # per-image stream: an image does not depend on which others are generated
        Xc = X @ R.T + t
        uv = Xc[:, :2] / Xc[:, 2:] * F + [U0, V0]
        ok = np.flatnonzero((Xc[:, 2] > 0.1) & (uv[:, 0] >= 0) & (uv[:, 0] < W) & (uv[:, 1] >= 0) & (uv[:, 1] < H))
        pick = rng.permutation(ok)[:n_vis]
        n_clut = n_kps - len(pick)
        uv_i = np.concatenate([uv[pick] + rng.normal(0, noise_px, (len(pick), 2)), rng.uniform(0, 1, (n_clut, 2)) * [W, H]])
        # Clutter keypoints sit at random pixels but carry the appearance of 3-D points this image does NOT observe, so
        # they form confident, geometrically wrong matches with the images that do (the outliers RANSAC must reject).
        hidden = np.setdiff1d(np.arange(n_pts), pick)
        clut = rng.choice(hidden if len(hidden) else np.arange(n_pts), n_clut, replace=len(hidden) < n_clut)
        cp = protos[clut].copy()
        fresh = rng.random(n_clut) >= clutter_hidden          # the rest of the clutter looks like nothing else
        nf = int(fresh.sum())
        if nf:
            cp[fresh] = sift_prototypes(rng, nf, dim) if kind == "sift" else rng.standard_normal((nf, dim), dtype=np.float32)
        if kind == "sift":
            d_i = np.concatenate([sift_observe(rng, protos[pick]), sift_observe(rng, cp)])
        else:
            d_i = np.concatenate([superpoint_observe(rng, protos[pick]), superpoint_observe(rng, cp)])
        id_i = np.concatenate([pick, -np.ones(n_clut, np.int64)])
        perm = rng.permutation(n_kps)
        descs.append(np.ascontiguousarray(d_i[perm]))
        kps.append(np.ascontiguousarray(uv_i[perm]))
        ids.append(id_i[perm])
    pr = make_pairs(n_images, pairs, top_k, angles, max_pairs)
    return Scene(descs, kps, ids, Rs, ts, pr, CAL, ratio, n_hyp)


def config_pairs(cfg: dict) -> np.ndarray:
    """The pair list of a config without generating any features (cameras are drawn first from the scene's generator)."""
    return make_scene(**cfg, only_images=[]).pairs


def make_config(name: str, n_images: Optional[int] = None, n_kps: Optional[int] = None, max_pairs: Optional[int] = None) -> Scene:
    """A BASELINE.json config (C2..C5), optionally shrunk (tests use the same recipe at smaller sizes)."""
    cfg = dict(CONFIGS[name])
    if n_images:
        cfg["n_images"] = n_images
    if n_kps:
        cfg["n_kps"] = n_kps
    if max_pairs:
        cfg["max_pairs"] = max_pairs
    return make_scene(**cfg)


def load_c1(path: str) -> Scene:
    """BASELINE config C1 from the committed fixture (tests/golden/c1_lund_door_12.npz, written by oracle/gen_golden.py c1):
    the reference's own 12 Lund-door images through its SIFT front end, intrinsics and GT poses as the Olsson loader
    builds them (gtsfm/loader/olsson_loader.py:75-147), all 66 pairs (ExhaustiveRetriever)."""
    g = np.load(path)
    off = g["img_off"]
    n = len(off) - 1
    descs = [g["desc_u8"][off[i]:off[i + 1]].astype(np.float32) for i in range(n)]
    kps = [g["kps"][off[i]:off[i + 1]].astype(np.float64) for i in range(n)]
    Rs = [g["wRc"][i].T for i in range(n)]
    ts = [-(g["wRc"][i].T @ g["wtc"][i]) for i in range(n)]
    ids = [np.full(len(d), -1, np.int64) for d in descs]
    return Scene(descs, kps, ids, Rs, ts, np.ascontiguousarray(g["pairs"], np.int32), tuple(float(v) for v in g["cal"][0]), 0.8, 1024)
