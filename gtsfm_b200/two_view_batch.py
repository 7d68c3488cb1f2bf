"""Batched two-view front-end: the caller-side replacement of the reference's per-pair submission loops
(SURVEY.md 8(a) row a12):

  DetDescCorrespondenceGenerator.generate_correspondences   det_desc_correspondence_generator.py:66-85  (match per pair)
  run_two_view_estimator_as_futures                         two_view_estimator.py:567-587               (verify per pair)

Instead of one Dask task per pair, one call per GPU shard uploads the shard's images once, matches every pair, keeps
the matches on the device and verifies them there.  Outputs use the reference's dictionary conventions
(gtsfm_runner_base.py:415-440): {(i1, i2): value}; pairs whose verification failed carry None rotations.
"""
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from gtsfm_b200 import _native
from gtsfm_b200.common.types import Rot3, Unit3, cal_to_tuple


@dataclass
class TwoViewBatchResult:
    pairs: np.ndarray                     # (P, 2) int32
    match_offsets: np.ndarray             # (P + 1,) int64
    matches: np.ndarray                   # (total, 2) uint32 putative correspondences (TwoWayMatcher order)
    inlier_mask: np.ndarray               # (total,) uint8
    R: np.ndarray                         # (P, 3, 3) i2Ri1
    t: np.ndarray                         # (P, 3) i2Ui1
    n_inliers: np.ndarray                 # (P,)
    status: np.ndarray                    # (P,) 0 ok, 1 too few matches, 2 no model
    # Lifetime: with pinned_outputs=True the arrays are views of the context's page-locked arena and are overwritten by the
    # next batch on that context (the memory itself stays valid until the context is closed); copy what must outlive it.

    def _need_matches(self):
        if len(self.matches) != int(self.match_offsets[-1]):
            raise ValueError("this result was produced with fetch_matches=False: the match rows stayed on the device, so the "
                             "correspondence dictionaries / reports cannot be built from it")

    def putative_corr_idxs_dict(self) -> Dict[Tuple[int, int], np.ndarray]:
        self._need_matches()
        o = self.match_offsets
        return {(int(a), int(b)): self.matches[o[p]:o[p + 1]] for p, (a, b) in enumerate(self.pairs)}

    def two_view_dicts(self):
        """(i2Ri1_dict, i2Ui1_dict, v_corr_idxs_dict, inlier_ratio_dict), built lazily from the structure-of-arrays."""
        self._need_matches()
        o = self.match_offsets
        Rd, td, vd, rd = {}, {}, {}, {}
        for p, (a, b) in enumerate(self.pairs):
            key = (int(a), int(b))
            m = self.matches[o[p]:o[p + 1]]
            if self.status[p] != 0:
                Rd[key], td[key], vd[key], rd[key] = None, None, np.array([], dtype=np.uint64), 0.0
                continue
            mask = self.inlier_mask[o[p]:o[p + 1]]
            Rd[key], td[key] = Rot3(self.R[p].copy()), Unit3(self.t[p].copy())
            vd[key] = m[mask == 1]
            rd[key] = float(mask.mean()) if mask.size else 0.0
        return Rd, td, vd, rd


def two_view_reports(result: TwoViewBatchResult, ctx: "_native.Context", corr_metric_dist_threshold: float,
                     min_num_inliers_est_model: int, min_inlier_ratio_est_model: float,
                     gt_i2Ri1: Optional[np.ndarray] = None, gt_i2ti1: Optional[np.ndarray] = None,
                     gt_cal: Optional[np.ndarray] = None):
    """The stage after verify() in TwoViewEstimator.run_2view (two_view_estimator.py:298-349) for every pair of the batch the
    context verified last: GT correspondence metrics, relative-pose errors, report aggregates and the InlierSupportProcessor
    decision, computed on the device in one launch.  Returns (reports, keep): a TwoViewEstimationReport per pair (the
    pre-ISP report), the boolean keep mask and the raw arrays.  gt_i2Ri1 rows of NaN mark pairs without ground truth; gt_cal
    (P, 10) are the GT cameras' calibrations when they differ from the ones the images were uploaded with."""
    from gtsfm_b200.common.two_view_estimation_report import TwoViewEstimationReport
    result._need_matches()
    P = len(result.pairs)
    o = result.match_offsets
    rep = ctx.two_view_report_batch(P, int(o[-1]), corr_metric_dist_threshold, min_num_inliers_est_model,
                                    min_inlier_ratio_est_model, gt_i2Ri1, gt_i2ti1, gt_cal=gt_cal)
    # a pair without ground truth (one of its gt cameras is None, two_view_estimator.py:131) carries a NaN row
    gt_rows = None if gt_i2Ri1 is None else ~np.isnan(np.asarray(gt_i2Ri1, np.float64).reshape(P, 9)[:, 0])
    reports = []
    for p in range(P):
        have_gt = gt_rows is not None and bool(gt_rows[p])
        m = result.matches[o[p]:o[p + 1]]
        ok = result.status[p] == 0
        mask = result.inlier_mask[o[p]:o[p + 1]] == 1 if ok else np.zeros(len(m), bool)
        v = m[mask] if ok else np.array([], dtype=np.uint64)
        ratio = float(mask.mean()) if (ok and mask.size) else 0.0
        if have_gt and ok and len(v) > 0:
            gm = rep["gt_mask"][o[p]:o[p + 1]][mask].astype(bool)
            err = rep["dist_sq"][o[p]:o[p + 1]][mask]
            n_gt = int(rep["num_inliers_gt"][p])
            reports.append(TwoViewEstimationReport(
                v_corr_idxs=v, num_inliers_est_model=len(v), inlier_ratio_est_model=ratio, num_inliers_gt_model=n_gt,
                inlier_ratio_gt_model=n_gt / len(v), v_corr_idxs_inlier_mask_gt=gm, R_error_deg=float(rep["R_error_deg"][p]),
                U_error_deg=float(rep["U_error_deg"][p]), reproj_error_gt_model=err,
                inlier_avg_reproj_error_gt_model=float(rep["inlier_avg_err"][p]),
                outlier_avg_reproj_error_gt_model=float(rep["outlier_avg_err"][p])))
        else:
            reports.append(TwoViewEstimationReport(
                v_corr_idxs=v, num_inliers_est_model=len(v), inlier_ratio_est_model=ratio, num_inliers_gt_model=0,
                inlier_ratio_gt_model=float("nan"), R_error_deg=float(rep["R_error_deg"][p]) if (have_gt and ok) else None,
                U_error_deg=float(rep["U_error_deg"][p]) if (have_gt and ok) else None,
                inlier_avg_reproj_error_gt_model=float("nan"), outlier_avg_reproj_error_gt_model=float("nan")))
    return reports, rep["keep"].astype(bool), rep


def num_gpu_shards() -> int:
    """Shards a batched caller deals its pair list into: GTSFM_B200_NUM_SHARDS when set, else the visible CUDA devices."""
    import os
    env = os.environ.get("GTSFM_B200_NUM_SHARDS")
    if env:
        return max(1, int(env))
    try:
        import torch
        return max(1, torch.cuda.device_count())
    except Exception:  # noqa: BLE001 - torch is plumbing here; without it there is one shard
        return 1


def num_gpu_devices() -> int:
    try:
        import torch
        return max(1, torch.cuda.device_count())
    except Exception:  # noqa: BLE001
        return 1


def shard_pairs(pairs: np.ndarray, n_shards: int, cost: Optional[np.ndarray] = None) -> List[np.ndarray]:
    """Contiguous, cost-balanced split of an i1-sorted pair list: returns index arrays per shard.
    Contiguity keeps each shard's image set compact, so only those images have to be placed on that GPU."""
    P = len(pairs)
    cost = np.ones(P) if cost is None else np.asarray(cost, float)
    order = np.lexsort((pairs[:, 1], pairs[:, 0])) if P else np.zeros(0, np.int64)
    csum = np.cumsum(cost[order])
    total = csum[-1] if P else 0.0
    # shard k ends with the first pair whose running cost reaches k / n_shards of the total
    bounds = [min(P, int(np.searchsorted(csum, total * k / n_shards * (1 - 1e-12), side="left")) + 1) if P else 0
              for k in range(1, n_shards)]
    bounds = list(np.maximum.accumulate(bounds)) if bounds else bounds
    return [order[a:b] for a, b in zip([0] + bounds, bounds + [P])]


def run_two_view_batch(descriptors: Sequence[np.ndarray], keypoints: Sequence[np.ndarray], intrinsics: Sequence, pairs: np.ndarray,
                       ratio_test_threshold: Optional[float] = 0.8, estimation_threshold_px: float = 4.0, num_hypotheses: int = 1024,
                       seed: int = 0, lo_iters: int = 20, device: int = 0, ctx: Optional[_native.Context] = None,
                       upload: bool = True, fetch_matches: bool = True, pinned_outputs: bool = False) -> TwoViewBatchResult:
    """Match + verify every pair of one shard on one GPU.  Only the images a shard's pairs name are uploaded; image
    indices in `pairs` refer to the caller's lists.  keypoints: (N, 2) pixel arrays; intrinsics: Cal3Bundler-likes or
    (f, k1, k2, u0, v0) tuples.  pinned_outputs: the result arrays are views of the context's page-locked arena (device -> host
    copies at full PCIe rate instead of through pageable staging: ~55 ms less per 5,000 C2-sized pairs); they stay valid until
    the next batch on the same context, so copy what must outlive it."""
    pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
    ctx = ctx or _native.default_context(device)
    if upload:
        used = np.unique(pairs)
        remap = -np.ones(len(descriptors), np.int64)
        remap[used] = np.arange(len(used))
        ctx.clear()
        for k, i in enumerate(used):
            ctx.upload_image(k, descriptors[i], keypoints[i], cal_to_tuple(intrinsics[i]))
        local = remap[pairs].astype(np.int32)
    else:
        local = pairs
    p = _native.VerifyParams()
    p.threshold_px, p.n_hypotheses, p.lo_iters, p.seed = float(estimation_threshold_px), int(num_hypotheses), int(lo_iters), int(seed)
    keys = (pairs[:, 0].astype(np.uint64) << np.uint64(32)) | pairs[:, 1].astype(np.uint64)   # sampler keys = global pair ids
    p.pair_keys, p.samples = keys.ctypes.data, None
    matches, off = ctx.match_batch(local, ratio_test_threshold, fetch=fetch_matches, reuse_outputs=pinned_outputs)
    out = ctx.verify_batch(local, p, total=int(off[-1]), reuse_outputs=pinned_outputs)
    return TwoViewBatchResult(pairs, off, matches, out["mask"], out["R"], out["t"], out["n_inliers"], out["status"])


# ------------------------------------------------------------------------------------------------ multi-GPU result exchange
RECORD_WIDTH = 16   # R (9) | t (3) | n_inliers | status | n_matches | global pair id


def pack_records(result: TwoViewBatchResult, pair_ids: np.ndarray) -> np.ndarray:
    """Fixed-width per-pair header: the only data that crosses GPUs.  pair_ids are the positions of the
    shard's pairs in the global pair list, so the receiver can put every record back in place."""
    P = len(result.pairs)
    rec = np.zeros((P, RECORD_WIDTH), np.float64)
    rec[:, 0:9] = result.R.reshape(P, 9)
    rec[:, 9:12] = result.t
    rec[:, 12] = result.n_inliers
    rec[:, 13] = result.status
    rec[:, 14] = np.diff(result.match_offsets)
    rec[:, 15] = pair_ids
    return rec


def gather_records(rec, n_global: int, dist=None, device=None) -> np.ndarray:
    """All-gather the shards' record blocks (NCCL over NVLink on the GPU box, gloo in the CPU tests) and scatter them
    into global pair order.  Shards may differ in length: blocks are padded to the longest one, padding rows carry id -1."""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        blocks = [np.asarray(rec)]
    else:
        world = dist.get_world_size()
        n = torch.tensor([len(rec)], dtype=torch.int64, device=device)
        ns = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(ns, n)
        longest = max(int(x) for x in ns)
        pad = torch.full((longest, RECORD_WIDTH), -1.0, dtype=torch.float64, device=device)
        pad[:len(rec)] = torch.as_tensor(rec, dtype=torch.float64, device=device)
        out = [torch.zeros_like(pad) for _ in range(world)]
        dist.all_gather(out, pad)
        blocks = [o[:int(k)].cpu().numpy() for o, k in zip(out, ns)]
    table = np.full((n_global, RECORD_WIDTH), np.nan)
    for blk in blocks:
        ids = blk[:, 15].astype(np.int64)
        table[ids] = blk
    return table
