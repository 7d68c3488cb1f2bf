#!/usr/bin/env python
"""bench.py -- image pairs/s of the two-view front-end hot path (TwoWayMatcher + Ransac verifier) on B200.

  python bench.py --gpus N --steps K --warmup W            the B200 path (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference's CPU path (cv2, one task per pair, all cores)

Workload (BASELINE.json): config C3 = synthetic 1,000 images x 8192 SIFT-128 keypoints, retriever top-50 = 48,725 pairs, ratio
test 0.8, 1,024 five-point hypotheses per pair, 4 px threshold -- the configuration BASELINE quotes "sharded 1/2/4/8 B200".  A
"step" is one pass over ALL pairs of that one scene: match + verify + the result table (i2Ri1, i2Ui1, v_corr_idxs, inlier
ratio of every pair) in pinned host memory of rank 0.  N = 1 runs the whole pair list on one GPU; N > 1 deals the SAME list
over the ranks with shard_pairs (strong scaling), every rank uploads only the images its shard names, and the result records
are packed on the device and moved with one NCCL all-gather per sub-batch (gtsfm_b200.two_view_batch.ResultExchange).
`value` times the pass with descriptors / keypoints already resident in HBM; `e2e` times it from pinned host buffers, uploads
included.  `--config C2` (100 images, 4,950 pairs, one GPU) and `--scaling weak` (every rank its own scene) remain available.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "image_pairs_per_sec_match_plus_ransac_verify"
_REAL_STDOUT = sys.stdout
UNIT = "pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C3")
    ap.add_argument("--images", type=int, default=0, help="shrink the config (debugging only; the line then says so)")
    ap.add_argument("--kps", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-pairs", type=int, default=0, help="pairs per reference step (default: 2 per host core)")
    ap.add_argument("--verifier", default="essential", choices=["essential", "fundamental"],
                    help="essential = Ransac(use_intrinsics_in_verification=True) (BASELINE configs); fundamental = the "
                         "reference's F-matrix mode (ransac.py:84-111), a SURVEY 8(f) row measured with the same harness")
    ap.add_argument("--scaling", choices=("weak", "strong"), default="strong",
                    help="N > 1: strong (default) = ONE scene, its pair list dealt over the ranks by shard_pairs, every rank uploads only "
                         "the images its shard names; weak = every rank its own scene of the configured size")
    ap.add_argument("--sub-batches", type=int, default=0,
                    help="sub-batches per step whose result exchange overlaps the next sub-batch's kernels (0 = 1 on one GPU, 4 otherwise)")
    ap.add_argument("--ratio", type=float, default=None, help="override the config's ratio-test threshold (e.g. C4 with the ratio test on)")
    ap.add_argument("--workspace-gib", type=float, default=0.0, help="matcher / verifier work-buffer budget per chunk (0 = library default)")
    ap.add_argument("--float-screening", default="on", choices=["on", "off"],
                    help="float descriptors, ratio test off (C4): fp16 tensor-core screening (default) or the exhaustive fp32 kernel")
    return ap.parse_args()


def load_traffic(args, cfg, pairs_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum of one match_screen_kernel launch, from the committed `ncu --set full` capture
    of a full workspace chunk of the C3 workload (profiles/r2_screen_ncu.json: 512 pairs of 8192 x 8192 x 128).  C3 and C2 have the
    same pair shape, so the capture's bytes per pair are scaled to this run's average pairs per launch (the figure `achieved` is
    quoted on); None for any other workload.  Returns (bytes, provenance)."""
    name = "r2_screen_ncu.json"
    if args.config not in ("C3", "C2") or args.images or args.kps or not os.path.exists(os.path.join(ROOT, "profiles", name)):
        return None, None
    with open(os.path.join(ROOT, "profiles", name)) as f:
        rec = json.load(f)
    cap_pairs = rec.get("pairs_per_launch", 512)
    src = (f"profiles/{name}: {rec.get('dram_bytes_per_launch', 0) / 1e9:.2f} GB for {cap_pairs} pairs per launch, captured at commit "
           f"{rec.get('commit', '?')}; scaled to this run's {pairs_per_launch:.0f} pairs per launch")
    return rec.get("dram_bytes_per_launch") * pairs_per_launch / cap_pairs, src


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(bf16_burst=p["bf16_tflops"], bf16_sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]), hbm=p["hbm_gbs"], source="measured")
    except Exception:  # noqa: BLE001
        return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:  # noqa: BLE001
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_scene(args, rank: int, shard=None):
    """shard = (rank, world) of a strong-scaled run: only the images that rank's pairs name are generated (the scene's cameras
    and pair list are always complete, and an image's features do not depend on which other images are generated)."""
    from gtsfm_b200 import synthetic
    if args.config == "C1":            # the reference's own CPU-runnable case: real SIFT features of its 12 Lund-door images
        scene = synthetic.load_c1(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "c1_lund_door_12.npz"))
        scene.fundamental = args.verifier == "fundamental"
        return scene, dict(n_images=12, n_kps=5000, kind="lund-door-sift", dim=128, pairs="exhaustive", n_hyp=1024, ratio=0.8, seed=0)
    cfg = dict(synthetic.CONFIGS[args.config])
    if args.images:
        cfg["n_images"] = args.images
    if args.kps:
        cfg["n_kps"] = args.kps
    cfg["seed"] = cfg["seed"] + 1000 * rank          # weak scaling: every rank its own scene of the same shape
    only = None
    if shard is not None and shard[1] > 1:
        from gtsfm_b200.two_view_batch import shard_pairs
        all_pairs = synthetic.config_pairs(cfg)
        only = np.unique(all_pairs[shard_pairs(all_pairs, shard[1])[shard[0]]])
    scene = synthetic.make_scene(**cfg, only_images=only)
    scene.fundamental = args.verifier == "fundamental"
    if args.ratio is not None:
        scene.ratio = args.ratio
    return scene, cfg


def workload_name(args, cfg, n_pairs):
    s = f"{args.config}: {'reference fixture' if args.config == 'C1' else 'synthetic'} {cfg['n_images']} images x {cfg['n_kps']} {cfg['kind']}-{cfg['dim']} keypoints, {n_pairs} pairs ({cfg['pairs']})"
    if args.verifier == "fundamental":
        s += " [verifier in fundamental-matrix mode - not the BASELINE configuration]"
    if args.ratio is not None:
        s += f" [ratio test {args.ratio} - not the BASELINE configuration]"
    if args.images or args.kps:
        s += " [SHRUNK - not the BASELINE configuration]"
    return s


# ------------------------------------------------------------------------------------------------ reference arm (CPU)
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_cv2
    scene, cfg = make_scene(args, 0)
    cores = len(os.sched_getaffinity(0))
    per_step = args.ref_pairs or 2 * cores
    pool = ref_cv2.PairPool(scene, cores)
    rng = np.random.default_rng(0)
    n_pairs = len(scene.pairs)

    def step():
        ids = rng.choice(n_pairs, min(per_step, n_pairs), replace=False)
        t0 = time.perf_counter()
        pool.run(ids)
        return len(ids), time.perf_counter() - t0

    for _ in range(args.warmup):
        step()
    done, secs = 0, 0.0
    for _ in range(args.steps):
        n, dt = step()
        done += n; secs += dt
    pool.close()
    value = done / secs
    sample = f"{per_step} random pairs per step, one task per pair on {cores} single-threaded cv2 workers"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "reference fixture (Lund door SIFT features)" if args.config == "C1" else "synthetic", "config": {"workload": workload_name(args, cfg, n_pairs),
                       "n_hypotheses": "cv2 FM_RANSAC (adaptive, maxIters 1e6)" if args.verifier == "fundamental" else "cv2 USAC_ACCURATE default (maxIters 1000)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def cpu_baseline_sample(scene, cfg):
    """The reference's CPU path (oracle/ref_cv2.py) on a bounded sample: 2 pairs per host core, all cores."""
    from oracle import ref_cv2
    cores = len(os.sched_getaffinity(0))
    n = min(len(scene.pairs), 2 * cores)
    pool = ref_cv2.PairPool(scene, cores, keep_poses=True)
    ids = np.random.default_rng(1).choice(len(scene.pairs), n, replace=False)
    pool.run(ids[:cores])                      # warm the workers (fork, cv2 import)
    t0 = time.perf_counter()
    res = pool.run(ids)
    dt = time.perf_counter() - t0
    pool.close()
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} random pairs of the same workload, one task per pair on {cores} single-threaded cv2 workers ({dt:.1f} s)"}, (ids, res)


def agreement_vs_cv2(table, sample, scene=None):
    """agreement of the device verifier with black-box cv2 (USAC_ACCURATE / FM_RANSAC) on the sampled pairs, as
    a distribution: rotation / translation-direction angle between the two poses, IoU of the two verified-correspondence sets."""
    ids, res = sample
    r_deg, t_deg, iou = [], [], []
    gt = {"device_R": [], "device_t": [], "cv2_R": [], "cv2_t": []}
    ang = lambda A, B: float(np.degrees(np.arccos(np.clip((np.trace(A.T @ B) - 1) / 2, -1, 1))))
    dirang = lambda a, b: float(np.degrees(np.arccos(np.clip(a @ b / np.linalg.norm(a) / np.linalg.norm(b), -1, 1))))
    for g, out in zip(ids, res):
        if len(out) < 5 or out[2] is None:
            continue
        h = table.headers[int(g)]
        if h["status"] != 0:
            continue
        R, t = h["R"].reshape(3, 3), h["t"]
        r_deg.append(float(np.degrees(np.arccos(np.clip((np.trace(R.T @ out[2]) - 1) / 2, -1, 1)))))
        t_deg.append(float(np.degrees(np.arccos(np.clip(t @ out[3] / np.linalg.norm(out[3]), -1, 1)))))
        a = set(map(tuple, table.v_corr_idxs(int(g)).tolist()))
        b = set(map(tuple, out[4].tolist()))
        iou.append(len(a & b) / max(len(a | b), 1))
        if scene is not None:          # both against the synthetic scene's ground truth: which of the two is off when they disagree
            Rg, tg = scene.relative_pose(int(scene.pairs[int(g)][0]), int(scene.pairs[int(g)][1]))
            gt["device_R"].append(ang(R, Rg)); gt["device_t"].append(dirang(t, tg))
            gt["cv2_R"].append(ang(out[2], Rg)); gt["cv2_t"].append(dirang(out[3], tg))
    if not r_deg:
        return None
    q = lambda v, p: float(np.percentile(v, p))
    return {"pairs": len(r_deg), "R_deg": {"median": q(r_deg, 50), "p95": q(r_deg, 95), "max": max(r_deg)},
            "t_deg": {"median": q(t_deg, 50), "p95": q(t_deg, 95), "max": max(t_deg)},
            "inlier_set_iou": {"median": q(iou, 50), "p5": q(iou, 5), "min": min(iou)},
            "vs_ground_truth_deg": {k: {"median": q(v, 50), "p95": q(v, 95)} for k, v in gt.items() if v}}


def run_b200(args):
    import torch
    import gtsfm_b200
    from gtsfm_b200 import _native

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    strong = args.scaling == "strong" or world == 1
    scene, cfg = make_scene(args, 0 if strong else rank, shard=(rank, world) if strong else None)
    from gtsfm_b200.two_view_batch import ResultExchange, merge_gathered, shard_pairs
    if strong:                                           # contiguous cost-balanced blocks of the i1-sorted pair list
        mine = shard_pairs(scene.pairs, world)[rank]
        pairs_g = np.ascontiguousarray(scene.pairs[mine])
        used = np.unique(pairs_g)
        remap = -np.ones(len(scene.descriptors), np.int64)
        remap[used] = np.arange(len(used))
        pairs = np.ascontiguousarray(remap[pairs_g].astype(np.int32))
        P_total = len(scene.pairs)
        pair_ids = np.ascontiguousarray(mine, np.int64)                    # positions in the global pair list
        # the rank works through its pairs in the order of the LAST image a pair names: sub-batch 0 then needs a prefix of the
        # (in-order) image uploads only and starts while the rest is still arriving.  Results go to pair_ids, sampler keys are
        # global pair ids, so the order changes nothing but the schedule.
        order = np.argsort(pairs.max(axis=1), kind="stable")
        pairs, pairs_g, pair_ids = np.ascontiguousarray(pairs[order]), np.ascontiguousarray(pairs_g[order]), np.ascontiguousarray(pair_ids[order])
    else:
        pairs_g = pairs = scene.pairs
        used = np.arange(len(scene.descriptors))
        P_total = world * len(pairs)
        pair_ids = rank * len(pairs) + np.arange(len(pairs), dtype=np.int64)   # weak scaling: every rank owns P pairs of the job
    P = len(pairs)
    n_img = len(used)
    # CPU baseline first: the worker pool is forked before this process creates its CUDA context
    cpu, cpu_sample = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, cpu_sample = cpu_baseline_sample(scene, cfg)
    ctx = gtsfm_b200.Context(local)
    ctx.set_float_screening(args.float_screening == "on")
    if args.workspace_gib > 0:
        ctx.set_workspace(int(args.workspace_gib * (1 << 30)))

    # pinned host copies of the step's inputs (the e2e leg uploads from these every step)
    desc_pin = [torch.from_numpy(scene.descriptors[i]).pin_memory() for i in used]
    kps_pin = [torch.from_numpy(scene.keypoints[i]).pin_memory() for i in used]
    h2d_bytes = sum(t.numel() * t.element_size() for t in desc_pin + kps_pin)

    prm = _native.VerifyParams()
    prm.threshold_px, prm.n_hypotheses, prm.lo_iters, prm.seed = scene.threshold_px, scene.n_hyp, 20, 20260002
    keys = (pairs_g[:, 0].astype(np.uint64) << np.uint64(32)) | pairs_g[:, 1].astype(np.uint64)     # sampler keys = global pair ids
    prm.samples = None
    prm.estimate_fundamental = 1 if args.verifier == "fundamental" else 0

    desc_np, kps_np, cal_np = [t.numpy() for t in desc_pin], [t.numpy() for t in kps_pin], np.tile(np.asarray(scene.cal, np.float64), (n_img, 1))

    def upload():
        ctx.clear()
        ctx.upload_images(0, desc_np, kps_np, cal_np)           # one library call: copies + prepare kernels enqueued on the copy stream

    # sub-batches of the rank's pair list (contiguous, equal pair counts): the exchange of sub-batch k -- device pack, NCCL
    # all-gather, copy to rank 0's pinned memory -- runs on torch's stream while the library's stream computes sub-batch k + 1
    # (the e2e leg always pipelines: sub-batch 0 starts as soon as ITS images are on the device, the rest of the upload runs
    # on the copy stream behind it)
    # a sub-batch costs about a millisecond of fixed work (two host synchronisations, the exchange), so a rank with few pairs
    # (C1: 66) runs them as one batch
    small = P < 1024
    n_sub = max(1, min(args.sub_batches or (1 if small else 4), P))
    n_sub_e2e = max(1, min(args.sub_batches or (1 if small else 4 if world == 1 else 6), P))

    def sub_bounds(n):
        # equal sub-batches except a last one of half the size: its exchange is the only one nothing overlaps
        w = np.ones(n)
        if n > 1:
            w[-1] = 0.5
        return np.concatenate([[0], np.round(np.cumsum(w) / w.sum() * P)]).astype(np.int64)
    exchange = ResultExchange(dist, root=0)
    last_total = [0]

    trace = [] if os.environ.get("GTSFM_B200_BENCH_TRACE") else None      # host timestamps per sub-batch (diagnostics)

    def compute(n_sub=n_sub):
        bounds = sub_bounds(n_sub)
        handles = []
        for k in range(n_sub):
            a, b = int(bounds[k]), int(bounds[k + 1])
            sub_keys = np.ascontiguousarray(keys[a:b])
            prm.pair_keys = sub_keys.ctypes.data
            t0 = time.perf_counter()
            _, off = ctx.match_batch(pairs[a:b], scene.ratio, fetch=False)               # putative matches stay on the device
            t1 = time.perf_counter()
            ctx.verify_batch(pairs[a:b], prm, total=int(off[-1]), fetch=False)           # so do masks and poses
            t2 = time.perf_counter()
            handles.append(exchange.start_device(ctx, b - a, pair_ids[a:b], slot=k))
            if trace is not None:
                trace.append((k, t1 - t0, t2 - t1, time.perf_counter() - t2))
            last_total[0] = int(off[-1])
        t3 = time.perf_counter()
        parts = [exchange.finish(h, P_total) for h in handles]                           # rank 0: the job's result table on the host
        table = merge_gathered(parts) if parts[0] is not None else None
        if trace is not None:
            trace.append((-1, time.perf_counter() - t3, 0.0, 0.0))
        return table, sum(h["d2h_bytes"] for h in handles)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """K steps bracketed by barrier + synchronize; device time from CUDA events on the library's stream, max over ranks."""
        sync_all()
        ctx.event_record(0)
        t0 = time.perf_counter()
        for _ in range(steps):
            last = fn()
        ctx.event_record(1)
        sync_all()
        wall = time.perf_counter() - t0
        ms = ctx.event_elapsed_ms(0, 1)
        ms = max(ms, 0.0)
        if world > 1:
            t = torch.tensor([ms, wall * 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall, last

    # ---- resident leg: inputs already in HBM
    upload()
    for _ in range(args.warmup):
        compute()
    ctx.profile_enable(not os.environ.get("GTSFM_B200_BENCH_NOPROF"))      # per-kernel CUDA events (their cost stays inside `value`)
    ctx.verify_stats(reset=True)
    launches0 = ctx.kernel_launches
    clocks = ClockSampler(local)
    clocks.start()
    ms, wall, (table, _) = timed(compute, args.steps)
    clk = clocks.stop()
    launches = ctx.kernel_launches - launches0
    if trace is not None:
        print(f"[trace rank {rank}] (sub-batch, match s, verify s, exchange-start s) of the timed resident steps; -1 = finish:",
              [(k, round(a * 1e3, 2), round(b * 1e3, 2), round(c * 1e3, 2)) for k, a, b, c in trace[-(n_sub + 1) * args.steps:]], file=sys.stderr)
    prof = ctx.profile_get()
    ctx.profile_enable(False)
    n_cand, n_hyp_solved, n_evals, n_evals_exec = ctx.verify_stats(reset=True)
    # the device-event time excludes nothing but host gaps after the last kernel; use the larger of the two clocks
    step_ms = max(ms, wall * 1e3) / args.steps
    value = P_total / (step_ms / 1e3)
    # per-rank kernel time of the timed steps (load balance of the static split; clocks differ between GPUs under the power cap)
    kern_ms = sum(v[0] for v in prof.values()) / args.steps
    per_rank_kernel_ms = [kern_ms]
    if world > 1:
        t = torch.tensor([kern_ms], dtype=torch.float64, device="cuda")
        allk = torch.empty(world, dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allk, t)
        per_rank_kernel_ms = [float(x) for x in allk.cpu()]
    stats = None
    if rank == 0:
        h = table.headers
        assert (h["pair_id"] == np.arange(P_total)).all(), "the gathered table misses pairs"
        okp = h["status"] == 0
        stats = dict(matches_per_pair=float(h["n_matches"].mean()), inliers_per_pair=float(h["n_inliers"].mean()), verified_ok=int(okp.sum()),
                     payload_rows=int(h["n_inliers"].sum()))
        # spot check of the exchange: the verified correspondences of a few pairs are rows of plausible keypoint indices
        for g in (0, P_total // 2, P_total - 1):
            v = table.v_corr_idxs(g)
            assert len(v) == h["n_inliers"][g] and (len(v) == 0 or int(v.max()) < cfg["n_kps"])
        if cpu_sample is not None:
            stats["agreement_vs_cv2"] = agreement_vs_cv2(table, cpu_sample, scene)

    # ---- the stage after verify() in run_2view (GT metrics + InlierSupportProcessor), timed on its own
    # over the pairs of the last sub-batch: it is not part of BASELINE's metric, so it stays outside `value` / `e2e`
    bounds = sub_bounds(n_sub)
    a, b = int(bounds[n_sub - 1]), int(bounds[n_sub])
    gt = [scene.relative_pose(int(i), int(j)) for i, j in pairs_g[a:b]]
    gt_R, gt_t = np.stack([g[0] for g in gt]), np.stack([g[1] for g in gt])
    rep = ctx.two_view_report_batch(b - a, last_total[0], scene.threshold_px, 15, 0.1, gt_R, gt_t, reuse_outputs=True)
    ctx.event_record(2)
    rep = ctx.two_view_report_batch(b - a, last_total[0], scene.threshold_px, 15, 0.1, gt_R, gt_t, reuse_outputs=True)
    ctx.event_record(3)
    report_ms = ctx.event_elapsed_ms(2, 3)
    report_extra = {"pairs": b - a, "ms_per_pass_incl_result_copies": report_ms, "pairs_kept_by_inlier_support": int(rep["keep"].sum()),
                    "median_R_error_deg": float(np.nanmedian(rep["R_error_deg"])) if np.isfinite(rep["R_error_deg"]).any() else None,
                    "median_U_error_deg": float(np.nanmedian(rep["U_error_deg"])) if np.isfinite(rep["U_error_deg"]).any() else None}

    # ---- two-view triangulation + bundle adjustment of the same pairs (the default configuration's step
    # between verify() and the inlier-support decision), timed on its own for the same reason; then the post-BA report
    n_verified = ctx.two_view_ba_batch(b - a, last_total[0], 1 << 30)["num_valid"].copy()    # floor above every pair: the verifier's counts
    ba = ctx.two_view_ba_batch(b - a, last_total[0], 15, 100, 100.0, 0.5, 1.345)
    ctx.event_record(2)
    ba = ctx.two_view_ba_batch(b - a, last_total[0], 15, 100, 100.0, 0.5, 1.345)
    ctx.event_record(3)
    rep_ba = ctx.two_view_report_batch(b - a, last_total[0], scene.threshold_px, 15, 0.1, gt_R, gt_t, reuse_outputs=True, after_ba=True)
    ran = ba["status"] == 0
    ba_extra = {"pairs": b - a, "pairs_adjusted": int(ran.sum()), "ms_per_pass_incl_result_copies": ctx.event_elapsed_ms(2, 3),
                "median_lm_iterations": float(np.median(ba["iterations"][ran])) if ran.any() else None,
                "tracks_kept_fraction": float(ba["num_valid"][ran].sum() / max(1, int(n_verified[ran].sum()))) if ran.any() else None,
                "median_R_error_deg_post_ba": float(np.nanmedian(rep_ba["R_error_deg"])) if np.isfinite(rep_ba["R_error_deg"]).any() else None,
                "median_U_error_deg_post_ba": float(np.nanmedian(rep_ba["U_error_deg"])) if np.isfinite(rep_ba["U_error_deg"]).any() else None}

    # ---- the retriever step in front of the path, outside the timed steps: similarity + top-50 of
    # 2,000 images x 4,096-d global descriptors (C4's image count, NetVLAD's dimension), host arrays in, pair list out
    retr_extra = None
    if rank == 0:
        rg = np.random.default_rng(5).standard_normal((2000, 4096)).astype(np.float32)
        rg /= np.linalg.norm(rg, axis=1, keepdims=True)
        ctx.retriever_similarity(rg, 50, fetch=False)
        ctx.retriever_topk(None, 2000, 50, 0.0)
        t0 = time.perf_counter()
        ctx.retriever_similarity(rg, 50, fetch=False)
        ridx, _ = ctx.retriever_topk(None, 2000, 50, 0.0)
        retr_extra = {"workload": "2000 images x 4096-d, top-50, min_score 0", "ms_incl_copies": (time.perf_counter() - t0) * 1e3,
                      "pairs": int((ridx >= 0).sum())}

    # ---- end-to-end leg: pinned host buffers -> device -> results on the host, every step
    def e2e_step():
        upload()
        return compute(n_sub_e2e)
    for _ in range(min(args.warmup, 1)):
        e2e_step()
    ems, ewall, (_, d2h_bytes) = timed(e2e_step, args.steps)
    if world > 1:                                          # whole-job copy volumes: uploads of every rank, downloads on rank 0
        t = torch.tensor([float(h2d_bytes)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        h2d_bytes = int(t[0])
    e2e_ms = max(ems, ewall * 1e3) / args.steps
    e2e_value = P_total / (e2e_ms / 1e3)

    if rank == 0:
        peaks = load_peaks()
        n_kps, dim = cfg["n_kps"], cfg["dim"]
        scr_ms, scr_calls = prof["screen"]
        flops_per_launch = 2.0 * n_kps * n_kps * dim * P * args.steps / max(scr_calls, 1)
        achieved = flops_per_launch / (scr_ms / max(scr_calls, 1) * 1e-3) / 1e12 if scr_ms > 0 else 0.0
        total_kernel_ms = sum(v[0] for v in prof.values())
        roofline = {"bound": "tensor", "kernel": "match_screen_kernel (tcgen05 distance GEMM + fused row/column minima)",
                    "achieved": achieved, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s", "frac": achieved / peaks["bf16_sustained"],
                    "peak_source": peaks["source"] + " (sustained bf16; burst %.1f)" % peaks["bf16_burst"], "traffic": None,
                    "algorithmic_flops_per_launch": flops_per_launch, "launch_ms": scr_ms / max(scr_calls, 1),
                    "share_of_step": scr_ms / total_kernel_ms if total_kernel_ms else None,
                    "kernel_ms_per_step": {k: v[0] / args.steps for k, v in prof.items()}}
        # MSAC scoring against the FP32 SIMT roofline (34 FLOP per candidate x correspondence)
        sc_ms, sc_calls = prof["score"]
        sm_mhz = clk.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
        sc_achieved = 34.0 * n_evals / (sc_ms * 1e-3) / 1e12 if sc_ms > 0 else 0.0
        sc_name = ("count_kernel (inlier count of every 7-point candidate over all correspondences, fp32 test with a rigorous error band, fp64 inside the band)"
                   if args.verifier == "fundamental" else "score_kernel (MSAC cost of every 5-point candidate over all correspondences, FFMA2)")
        roofline["scoring"] = {"bound": "fp32", "kernel": sc_name,
                               "achieved": sc_achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": sc_achieved / fp32_peak,
                               "peak_source": "148 SM x 128 lanes x 2 x %.0f MHz (median SM clock sampled during the timed region)" % sm_mhz,
                               "sampson_evaluations_per_step": n_evals / args.steps,
                               "executed_evaluations_per_step": n_evals_exec / args.steps,
                               "executed_tflops": 34.0 * n_evals_exec / (sc_ms * 1e-3) / 1e12 if sc_ms > 0 else 0.0, "candidates_per_hypothesis": n_cand / max(n_hyp_solved, 1),
                               "share_of_step": sc_ms / total_kernel_ms if total_kernel_ms else None}
        roofline["traffic"], roofline["traffic_source"] = load_traffic(args, cfg, P * args.steps / max(scr_calls, 1))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "bf16 GEMM / f32 scoring / f64 finalize",
                "data": "reference fixture (Lund door SIFT features)" if args.config == "C1" else "synthetic",
                "config": {"workload": workload_name(args, cfg, P_total if strong else P), "pairs_per_gpu": P, "ratio_test": scene.ratio, "n_hypotheses": scene.n_hyp,
                           "threshold_px": scene.threshold_px, "lo_iters": 20, "matches_per_pair": stats["matches_per_pair"],
                           "inliers_per_pair": stats["inliers_per_pair"], "verified_ok": stats["verified_ok"],
                           "l2": "inputs exceed L2 (descriptor store %.0f MB on rank 0)" % (n_img * n_kps * (dim * 7 + 96) / 1e6),
                           "result": "per pair a 128-byte header (R, t, counts, ratio) + v_corr_idxs, %d rows per step, in pinned host memory of rank 0"
                                     % stats["payload_rows"],
                           "sub_batches_per_step": {"resident": n_sub, "e2e": n_sub_e2e},
                           "parallelism": ((f"one pair list dealt over {world} GPUs (shard_pairs), {n_img} of {len(scene.descriptors)} images on rank 0, "
                                            "device-packed result records, one NCCL all-gather per sub-batch") if world > 1 else "1 GPU") if strong else
                                          f"every rank its own scene ({world} GPUs), device-packed result records, one NCCL all-gather per sub-batch"},
                "roofline": roofline, "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes)},
                "gpu_launches": int(launches), "clocks": clk, "device_event_ms_per_step": ms / args.steps,
                "agreement_vs_cv2": stats.get("agreement_vs_cv2"), "per_rank_kernel_ms_per_step": per_rank_kernel_ms, "two_view_report": report_extra, "two_view_ba": ba_extra, "retriever": retr_extra}
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def _claim_stdout():
    """Libraries (NCCL's version banner, torchrun children) print to fd 1; the contract is ONE JSON line on stdout.
    Keep a private copy of the real stdout for that line and point fd 1 at stderr for everything else."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    _REAL_STDOUT = _claim_stdout()
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
