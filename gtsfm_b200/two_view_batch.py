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
                     gt_cal: Optional[np.ndarray] = None, after_ba: bool = False, inlier_ratios: Optional[np.ndarray] = None):
    """The stage after verify() in TwoViewEstimator.run_2view (two_view_estimator.py:298-349) for every pair of the batch the
    context verified last: GT correspondence metrics, relative-pose errors, report aggregates and the InlierSupportProcessor
    decision, computed on the device in one launch.  Returns (reports, keep): a TwoViewEstimationReport per pair (the
    pre-ISP report), the boolean keep mask and the raw arrays.  gt_i2Ri1 rows of NaN mark pairs without ground truth; gt_cal
    (P, 10) are the GT cameras' calibrations when they differ from the ones the images were uploaded with.  after_ba: `result`
    holds the outputs of two_view_ba_batch (post-BA report); inlier_ratios then carries the pre-BA ratios the reference keeps
    (two_view_estimator.py:325-329)."""
    from gtsfm_b200.common.two_view_estimation_report import TwoViewEstimationReport
    result._need_matches()
    P = len(result.pairs)
    o = result.match_offsets
    rep = ctx.two_view_report_batch(P, int(o[-1]), corr_metric_dist_threshold, min_num_inliers_est_model,
                                    min_inlier_ratio_est_model, gt_i2Ri1, gt_i2ti1, gt_cal=gt_cal, after_ba=after_ba)
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
        if inlier_ratios is not None:
            ratio = float(inlier_ratios[p])
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
# one process per GPU, no data-path collective; the only exchange is the result of every pair -- "per-pair
# fixed header {n_matches, n_inliers, status, R[9], t[3], ratio} plus variable-length index payloads packed with a prefix-sum
# offset table (two-phase: gather counts -> gather payload)".  The block is built on the device (csrc/result_pack.cu), moved
# with one NCCL all-gather from the library's buffers and lands in pinned host memory with one copy; nothing is bounced
# through the host in between.  What travels is what unzip_two_view_results hands to the back end
# (gtsfm/runner/gtsfm_runner_base.py:415-440): i2Ri1, i2Ui1, v_corr_idxs, inlier ratio.
HEADER_DTYPE = np.dtype([("pair_id", "<i8"), ("n_matches", "<i4"), ("n_inliers", "<i4"), ("status", "<i4"), ("pad", "<i4"),
                         ("R", "<f8", (9,)), ("t", "<f8", (3,)), ("inlier_ratio", "<f8")])
HEADER_BYTES = 128
assert HEADER_DTYPE.itemsize == HEADER_BYTES


def block_layout(max_pairs: int, max_rows: int) -> Tuple[int, int]:
    """(payload offset, block size) in bytes of a rank's [headers | payload] block, padded to the largest rank."""
    payload_offset = HEADER_BYTES * int(max_pairs)
    return payload_offset, payload_offset + 8 * int(max_rows) + 16


def pack_block(result: TwoViewBatchResult, pair_ids: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Host mirror of the device pack (gtsfm_result_pack / gtsfm_result_pack_copy): (headers (P,) HEADER_DTYPE, payload
    (rows, 2) uint32).  Used by the CPU (gloo) tests and to check the device kernels; the GPU path never calls it."""
    result._need_matches()
    P = len(result.pairs)
    o = result.match_offsets
    hdr = np.zeros(P, HEADER_DTYPE)
    ok = np.asarray(result.status) == 0
    hdr["pair_id"] = pair_ids
    hdr["n_matches"] = np.diff(o)
    hdr["status"] = result.status
    hdr["n_inliers"] = np.where(ok, result.n_inliers, 0)
    hdr["R"] = np.where(ok[:, None], np.asarray(result.R, np.float64).reshape(P, 9), 0.0)
    hdr["t"] = np.where(ok[:, None], np.asarray(result.t, np.float64).reshape(P, 3), 0.0)
    with np.errstate(all="ignore"):
        hdr["inlier_ratio"] = np.where(ok & (hdr["n_matches"] > 0), hdr["n_inliers"] / np.maximum(hdr["n_matches"], 1), 0.0)
    keep = np.zeros(int(o[-1]), bool)
    for p in np.flatnonzero(ok):
        keep[o[p]:o[p + 1]] = result.inlier_mask[o[p]:o[p + 1]] == 1
    return hdr, np.ascontiguousarray(result.matches[keep], np.uint32).reshape(-1, 2)


def block_bytes(hdr: np.ndarray, payload: np.ndarray, max_pairs: int, max_rows: int) -> np.ndarray:
    """One rank's block as bytes, in the layout of block_layout (host mirror of gtsfm_result_pack_copy)."""
    off, size = block_layout(max_pairs, max_rows)
    buf = np.zeros(size, np.uint8)
    buf[: hdr.nbytes] = hdr.view(np.uint8)
    buf[off: off + payload.nbytes] = payload.reshape(-1).view(np.uint8)
    return buf


class GatheredTwoViewResults:
    """The result table of a whole job after the exchange.  It keeps the received blocks as they are -- per (sub-batch, rank)
    a header view and a payload view into the pinned receive buffer, no copy of up to gigabytes of indices -- and builds the
    global-order header table and the pair -> (block, first payload row) index on first use, so that finishing an exchange
    costs no host work proportional to the job."""

    def __init__(self, blocks: List[Tuple[np.ndarray, np.ndarray]], n_global: int):
        self._blocks = blocks                 # [(headers (n,) HEADER_DTYPE, payload (rows, 2) uint32)]
        self.n_global = int(n_global)
        self._headers = None
        self._where = None

    def _index(self):
        if self._headers is None:
            headers = np.zeros(self.n_global, HEADER_DTYPE)
            headers["pair_id"] = -1                     # pairs nobody reported
            where = np.zeros((self.n_global, 2), np.int64)
            for k, (h, _) in enumerate(self._blocks):
                if len(h) == 0:
                    continue
                ids = h["pair_id"]
                headers[ids] = h
                where[ids, 0] = k
                where[ids, 1] = np.concatenate([[0], np.cumsum(h["n_inliers"][:-1], dtype=np.int64)])
            self._headers, self._where = headers, where
        return self._headers, self._where

    @property
    def headers(self) -> np.ndarray:
        """(P_total,) HEADER_DTYPE, row = global pair id; pair_id -1 where no rank reported the pair."""
        return self._index()[0]

    def v_corr_idxs(self, pair_id: int) -> np.ndarray:
        headers, where = self._index()
        k, o = where[pair_id]
        return self._blocks[int(k)][1][int(o): int(o) + int(headers[pair_id]["n_inliers"])]

    def two_view_dicts(self, pairs: np.ndarray):
        """(i2Ri1_dict, i2Ui1_dict, v_corr_idxs_dict, inlier_ratio_dict) in the reference's conventions; failed pairs carry
        (None, None, empty uint64, 0.0) like VerifierBase's failure tuple (verifier_base.py:56)."""
        headers, _ = self._index()
        Rd, td, vd, rd = {}, {}, {}, {}
        for g, (a, b) in enumerate(np.asarray(pairs).reshape(-1, 2)):
            key, h = (int(a), int(b)), headers[g]
            if h["pair_id"] < 0 or h["status"] != 0:
                Rd[key], td[key], vd[key], rd[key] = None, None, np.array([], dtype=np.uint64), 0.0
                continue
            Rd[key], td[key] = Rot3(h["R"].reshape(3, 3).copy()), Unit3(h["t"].copy())
            vd[key], rd[key] = self.v_corr_idxs(g), float(h["inlier_ratio"])
        return Rd, td, vd, rd


def unpack_blocks(buf: np.ndarray, counts: np.ndarray, n_global: int) -> GatheredTwoViewResults:
    """buf: (world, block size) uint8 as received; counts: (world, 2) [pairs, payload rows] per rank.  Views only."""
    off, _ = block_layout(counts[:, 0].max(), counts[:, 1].max())
    blocks = []
    for r in range(buf.shape[0]):
        n, rows = int(counts[r, 0]), int(counts[r, 1])
        blocks.append((buf[r, : HEADER_BYTES * n].view(HEADER_DTYPE), buf[r, off: off + 8 * rows].view(np.uint32).reshape(-1, 2)))
    return GatheredTwoViewResults(blocks, n_global)


def _all_gather_bytes(block, dist, world: int):
    """All-gather equal-sized uint8 tensors: NCCL moves them as one collective straight between the GPUs' buffers; gloo (CPU
    tests) goes through the list form."""
    import torch
    out = torch.empty((world, block.numel()), dtype=torch.uint8, device=block.device)
    if block.is_cuda:
        dist.all_gather_into_tensor(out.view(-1), block)
    else:
        parts = [torch.empty_like(block) for _ in range(world)]
        dist.all_gather(parts, block)
        for r, p in enumerate(parts):
            out[r] = p
    return out


class ResultExchange:
    """Two-phase exchange of one (sub-)batch: phase 1 announces [pairs, payload rows] per rank, phase 2 all-gathers the
    [headers | payload] blocks.  On the GPU the block is written by the library straight into a torch tensor, the collective is
    enqueued behind it without a host synchronisation, and the gathered table is copied once into pinned host memory on
    `root` (all ranks when root is None); `finish()` waits for that copy.  With one process there is no collective."""

    def __init__(self, dist=None, root: Optional[int] = 0):
        self.dist = dist if (dist is not None and dist.is_initialized() and dist.get_world_size() > 1) else None
        self.world = self.dist.get_world_size() if self.dist else 1
        self.rank = self.dist.get_rank() if self.dist else 0
        self.root = root
        self._pinned = {}

    def _counts(self, n_pairs: int, rows: int, device):
        import torch
        mine = torch.tensor([n_pairs, rows], dtype=torch.int64, device=device)
        if not self.dist:
            return mine.cpu().numpy().reshape(1, 2)
        if mine.is_cuda:
            out = torch.empty(2 * self.world, dtype=torch.int64, device=device)
            self.dist.all_gather_into_tensor(out, mine)
        else:
            parts = [torch.empty_like(mine) for _ in range(self.world)]
            self.dist.all_gather(parts, mine)
            out = torch.cat(parts)
        return out.cpu().numpy().reshape(self.world, 2)

    def start_device(self, ctx: "_native.Context", n_pairs: int, pair_ids: np.ndarray, slot: int = 0):
        """Pack the last verify batch of `ctx` on its device and start the exchange; returns a handle for finish()."""
        import torch
        rows = ctx.result_pack(n_pairs, pair_ids)
        dev = torch.device("cuda", ctx.device)
        counts = self._counts(n_pairs, rows, dev)
        off, size = block_layout(counts[:, 0].max(), counts[:, 1].max())
        block = torch.empty(size, dtype=torch.uint8, device=dev)
        ctx.result_pack_copy(block.data_ptr(), size, off)
        stream = torch.cuda.current_stream(dev)
        ctx.stream_wait(stream.cuda_stream)                       # the collective is ordered behind the pack kernels, no host sync
        block.record_stream(stream)
        gathered = _all_gather_bytes(block, self.dist, self.world) if self.dist else block.view(1, -1)
        host = None
        if self.root is None or self.rank == self.root:
            host = self._host_buffer(slot, gathered.numel()).view(self.world, size)
            host.copy_(gathered, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(stream)
        return dict(counts=counts, host=host, event=ev, keep=(block, gathered), d2h_bytes=0 if host is None else gathered.numel())

    def start_host(self, result: TwoViewBatchResult, pair_ids: np.ndarray):
        """The same exchange from host arrays (CPU / gloo tests): the block is built by the host mirror of the device pack."""
        import torch
        hdr, payload = pack_block(result, pair_ids)
        counts = self._counts(len(hdr), len(payload), torch.device("cpu"))
        block = torch.from_numpy(block_bytes(hdr, payload, counts[:, 0].max(), counts[:, 1].max()))
        gathered = _all_gather_bytes(block, self.dist, self.world) if self.dist else block.view(1, -1)
        return dict(counts=counts, host=gathered, event=None, keep=None, d2h_bytes=0)

    def _host_buffer(self, slot: int, nbytes: int):
        import torch
        buf = self._pinned.get(slot)
        if buf is None or buf.numel() < nbytes:
            buf = self._pinned[slot] = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, pin_memory=True)
        return buf[:nbytes]

    @staticmethod
    def finish(handle, n_global: int) -> Optional[GatheredTwoViewResults]:
        """Wait for the exchange; the gathered table on the root (None elsewhere).  The arrays are views of the exchange's
        pinned buffer (slot), valid until that slot is used again."""
        if handle["event"] is not None:
            handle["event"].synchronize()
        if handle["host"] is None:
            return None
        return unpack_blocks(handle["host"].numpy(), handle["counts"], n_global)


def merge_gathered(parts: Sequence[GatheredTwoViewResults]) -> GatheredTwoViewResults:
    """Sub-batches of one job (disjoint pair ids) -> one table (the blocks are shared, nothing is copied)."""
    return GatheredTwoViewResults([b for part in parts for b in part._blocks], parts[0].n_global)
