"""CPU tests of the multi-GPU plumbing: row sharding (no collective on the data path) and,
with two gloo ranks, the shard -> process -> gather round trip plus the max-over-ranks
timing reduction bench.py uses.  Each rank processes its shard with the CPU oracle here
(there is no GPU in this container); on the GPU box the same shards go to
embed_truncate_norm."""
import os
import socket

import numpy as np
import pytest

from llm_mcp_b200 import embedgpu as eg


def test_shard_rows_covers_every_row_once():
    for batch in (0, 1, 7, 8, 1024, 1048576 + 3):
        for world in (1, 2, 3, 8):
            sh = eg.shard_rows(batch, world)
            assert len(sh) == world and sh[0][0] == 0 and sh[-1][1] == batch
            assert all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
            sizes = [b - a for a, b in sh]
            assert max(sizes) - min(s for s in sizes if s or batch < world) <= -(-batch // world)
            assert max(sizes) == (-(-batch // world) if batch else 0)


def test_shard_ragged_balances_bytes_not_rows():
    rng = np.random.default_rng(5)
    ks = np.array([64, 128, 256, 512, 1024])
    kpr = np.concatenate([np.full(rng.integers(1, 2049), rng.choice(ks)) for _ in range(256)])
    for world in (1, 2, 4, 8):
        sh = eg.shard_ragged(kpr, 4096, world)
        assert sh[0][0] == 0 and sh[-1][1] == len(kpr) and all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
        loads = np.array([kpr[a:b].sum() for a, b in sh], dtype=np.float64)
        assert loads.max() / loads.mean() < 1.01, loads
    rows = np.array([b - a for a, b in eg.shard_ragged(kpr, 4096, 8)])
    assert rows.max() / rows.min() > 1.05          # equal bytes, unequal row counts
    assert eg.shard_ragged(np.array([5000, 5000]), 100, 2) == [(0, 1), (1, 2)]   # k clamps to D


def test_shard_ragged_equals_the_oracle_cut(oracle):
    """etn_shard_ragged (C++, the cut embed_truncate_norm_ragged_host makes itself) against the
    plain-Python restatement, on random and on degenerate length patterns."""
    rng = np.random.default_rng(11)
    cases = [rng.integers(1, 1500, size=n) for n in (0, 1, 2, 7, 100, 5000)]
    cases += [np.ones(50, dtype=np.int64), np.full(64, 1024), np.array([1, 1, 900, 1, 1, 900, 1]), np.array([10 ** 6] * 5)]
    for kpr in cases:
        for world in (1, 2, 3, 8):
            assert eg.shard_ragged(kpr, 1024, world) == oracle.shard_ragged_ref(kpr, 1024, world), (kpr[:8], world)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from oracle import etn_oracle as oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        batch, D, k = 1001, 768, 256
        lo, hi = eg.shard_rows(batch, world)[rank]
        x = oracle.fill(2, hi - lo, D, np.float32, row0=lo)       # every rank generates its own shard
        y = oracle.renorm_exact(x, k)                              # no data-path collective
        elapsed = torch.tensor([0.010 * (rank + 1)], dtype=torch.float64)
        dist.barrier()
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)             # bench.py's max over ranks
        # C5-style: a ragged batch cut by kept bytes, every rank renormalises its own range
        rng = np.random.default_rng(5)
        kpr = rng.choice([64, 128, 256, 512], size=batch).astype(np.int32)
        rlo, rhi = eg.shard_ragged(kpr, D, world)[rank]
        xr = oracle.fill(5, rhi - rlo, D, np.float64, row0=rlo)
        yr, _ = oracle.ragged(xr, kpr[rlo:rhi], 1)
        gathered = [None] * world
        dist.all_gather_object(gathered, (lo, hi, y, rlo, rhi, yr))  # test-only gather
        if rank == 0:
            full = np.concatenate([g[2] for g in sorted(gathered, key=lambda g: g[0])])
            want = oracle.renorm_exact(oracle.fill(2, batch, D, np.float32), k)
            rag = np.concatenate([g[5] for g in sorted(gathered, key=lambda g: g[3])])
            rag_want, _ = oracle.ragged(oracle.fill(5, batch, D, np.float64), kpr, 1)
            same = bool(np.array_equal(full, want)) and bool(np.array_equal(rag, rag_want))
            q.put((same, float(elapsed.item()), [g[:2] for g in gathered]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_gloo_ranks_shard_and_reduce(oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    same, tmax, ranges = q.get(timeout=90)
    [p.join(30) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert same, "sharded result differs from the single-rank result"
    assert tmax == pytest.approx(0.020)
    assert sorted(ranges) == [(0, 501), (501, 1001)]
