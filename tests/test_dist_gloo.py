"""The N > 1 path on CPU: world_size 2, gloo.  Pairs are sharded with shard_pairs, every rank holds the TwoViewBatchResult
of its own shard (match rows, inlier masks, poses), and the two-phase exchange (counts, then [headers | payload] blocks)
reassembles the whole job's result -- verified correspondences included -- exactly as a single process produces it.  Same
code path as bench.py over NCCL, except that the block is built by the host mirror of the device pack kernels (the GPU test
tests/test_gpu_result_pack.py checks that mirror against the kernels byte for byte)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gtsfm_b200 import synthetic
from gtsfm_b200.two_view_batch import ResultExchange, TwoViewBatchResult, merge_gathered, pack_block, shard_pairs


def result_for(pairs: np.ndarray) -> TwoViewBatchResult:
    """Deterministic stand-in for the device output of one shard: every field is a function of the global pair, so a pair
    gets the same result whichever shard computes it."""
    P = len(pairs)
    counts, ms, masks, Rs, ts, ninl, status = [], [], [], [], [], [], []
    for a, b in pairs:
        rng = np.random.default_rng(int(a) * 1000 + int(b))
        n = int(rng.integers(0, 40))
        m = rng.integers(0, 5000, (n, 2)).astype(np.uint32)
        mask = (rng.random(n) < 0.5).astype(np.uint8)
        st = 1 if n < 6 else (2 if rng.random() < 0.1 else 0)
        counts.append(n); ms.append(m); masks.append(mask); status.append(st)
        Rs.append(rng.normal(size=(3, 3))); ts.append(rng.normal(size=3)); ninl.append(int(mask.sum()))
    off = np.zeros(P + 1, np.int64); np.cumsum(counts, out=off[1:])
    return TwoViewBatchResult(pairs.astype(np.int32), off, np.concatenate(ms) if P else np.zeros((0, 2), np.uint32),
                              np.concatenate(masks) if P else np.zeros(0, np.uint8), np.stack(Rs), np.stack(ts),
                              np.array(ninl, np.int32), np.array(status, np.int32))


def check_table(table, pairs: np.ndarray) -> bool:
    whole = result_for(pairs)
    hdr, payload = pack_block(whole, np.arange(len(pairs)))
    ok = bool(np.array_equal(table.headers, hdr))
    o = np.concatenate([[0], np.cumsum(hdr["n_inliers"])])
    for g in range(len(pairs)):
        ok &= bool(np.array_equal(table.v_corr_idxs(g), payload[o[g]:o[g + 1]]))
    Rd, td, vd, rd = table.two_view_dicts(pairs)
    Rw, tw, vw, rw = whole.two_view_dicts()
    for key in Rw:
        ok &= (Rd[key] is None) == (Rw[key] is None) and bool(np.array_equal(vd[key], vw[key])) and abs(rd[key] - rw[key]) < 1e-15
        if Rw[key] is not None:
            ok &= bool(np.array_equal(Rd[key].matrix(), Rw[key].matrix())) and bool(np.allclose(td[key].point3(), tw[key].point3()))
    return ok


def _worker(rank: int, world: int, port: int, n_images: int, n_sub: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pairs = synthetic.make_pairs(n_images, "exhaustive")
        mine = shard_pairs(pairs, world)[rank]
        ex = ResultExchange(dist, root=None)
        parts = []
        for sub in np.array_split(mine, n_sub):                     # sub-batches, as bench.py pipelines them
            handle = ex.start_host(result_for(pairs[sub]), sub)
            parts.append(ex.finish(handle, len(pairs)))
        table = merge_gathered(parts)
        ok = check_table(table, pairs)
        tmax = torch.tensor([float(rank + 1)], dtype=torch.float64)   # the max-over-ranks timing reduction of bench.py
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        q.put((rank, ok, len(mine), float(tmax)))
    finally:
        dist.destroy_process_group()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_single_process_exchange_is_the_identity():
    pairs = synthetic.make_pairs(8, "exhaustive")
    ex = ResultExchange(None)
    table = ex.finish(ex.start_host(result_for(pairs), np.arange(len(pairs))), len(pairs))
    assert check_table(table, pairs)


@pytest.mark.timeout(180)
@pytest.mark.parametrize("n_images,n_sub", [(9, 1), (12, 3)])          # 36 pairs (even split) and 66 pairs in 3 sub-batches
def test_sharded_results_with_payload_gather_to_the_single_process_result_world2(n_images, n_sub):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_images, n_sub, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=150) for _ in range(world)]
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    n_pairs = n_images * (n_images - 1) // 2
    assert sorted(r for r, *_ in got) == [0, 1]
    assert all(ok for _, ok, _, _ in got)
    assert sum(n for _, _, n, _ in got) == n_pairs
    assert all(t == 2.0 for *_, t in got)
