"""The N > 1 path on CPU: world_size 2, gloo.  Pairs are sharded with shard_pairs, every rank produces the records of
its own shard, and gather_records reassembles the global table on every rank -- the same code bench.py runs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gtsfm_b200 import synthetic
from gtsfm_b200.two_view_batch import TwoViewBatchResult, gather_records, pack_records, shard_pairs


def _result_for(pairs: np.ndarray) -> TwoViewBatchResult:
    """Deterministic stand-in for the device output of one shard: every field is a function of the global pair."""
    P = len(pairs)
    key = pairs[:, 0].astype(np.float64) * 1000.0 + pairs[:, 1]
    counts = (key % 17).astype(np.int64)
    off = np.zeros(P + 1, np.int64); np.cumsum(counts, out=off[1:])
    R = np.tile(np.eye(3), (P, 1, 1)) * key[:, None, None]
    t = np.stack([key, key + 1, key + 2], 1)
    return TwoViewBatchResult(pairs.astype(np.int32), off, np.zeros((off[-1], 2), np.uint32), np.zeros(off[-1], np.uint8), R, t,
                              (counts // 2).astype(np.int32), (counts < 6).astype(np.int32))


def _worker(rank: int, world: int, port: int, n_images: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pairs = synthetic.make_pairs(n_images, "exhaustive")
        shards = shard_pairs(pairs, world)
        mine = shards[rank]
        res = _result_for(pairs[mine])
        table = gather_records(pack_records(res, mine), len(pairs), dist)
        want = pack_records(_result_for(pairs), np.arange(len(pairs)))
        ok = bool(np.array_equal(table, want))
        # max-over-ranks timing reduction used by bench.py
        tmax = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        q.put((rank, ok, len(mine), float(tmax)))
    finally:
        dist.destroy_process_group()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(180)
@pytest.mark.parametrize("n_images", [9, 12])          # 36 pairs (even split) and 66 pairs
def test_sharded_records_gather_to_global_order_world2(n_images):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_images, q)) for r in range(world)]
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
