"""Multi-process host logic on CPU (gloo, world_size 2): the two partitions of SURVEY.md §8e.
The arithmetic in these tests is the ORACLE's (no GPU here); what is under test is the
sharding / gather / recombination logic of fourier_b200/shard.py that bench.py uses over NCCL:
a column-sharded MSM recombines to the unsharded commitment, and row-sharded results come
back in global row order."""
import os
import random
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from fourier_b200 import shard
    from oracle import bls12_381 as bls
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        R = bls.R
        rng = random.Random(5)  # same inputs on every rank
        T, M = 12, 5            # deliberately not divisible by the world size
        ss = [rng.randrange(R) for _ in range(T)]
        ks = [rng.randrange(R) for _ in range(T)]
        pts = [bls.g1_mul(bls.G1_GEN, s) for s in ss]
        # (1) one MSM over column ranges: 96-byte partials, gathered and added
        c0, nc = shard.split_range(T, world, rank)
        part = bls.g1_msm(pts[c0:c0 + nc], ks[c0:c0 + nc])
        parts = shard.all_gather_bytes(dist, bls.g1_serialize(part))
        total = bls.G1_INF
        for b in parts:
            total = bls.g1_add(total, bls.g1_deserialize(b))
        ok_msm = bls.g1_compress(total) == bls.g1_compress(bls.g1_msm(pts, ks))
        # (2) rows: every rank produces (commit, y, proof) for its rows; gather keeps global order
        counts = [shard.split_range(M, world, r)[1] for r in range(world)]
        r0, nr = shard.split_range(M, world, rank)
        rows = list(range(r0, r0 + nr))
        cs = [bytes([i]) * 48 for i in rows]
        ys = [bytes([100 + i]) * 32 for i in rows]
        ps = [bytes([200 + i]) * 48 for i in rows]
        blob = shard.pack_rows(cs, ys, ps)
        blob += bytes(max(counts) * 128 - len(blob))  # equal length for the collective
        gc, gy, gp = shard.unpack_rows(shard.all_gather_bytes(dist, blob), counts)
        ok_rows = (gc == [bytes([i]) * 48 for i in range(M)] and gy == [bytes([100 + i]) * 32 for i in range(M)]
                   and gp == [bytes([200 + i]) * 48 for i in range(M)])
        q.put((rank, ok_msm, ok_rows))
    finally:
        dist.destroy_process_group()


def test_split_range_covers_everything():
    sys.path.insert(0, ROOT)
    from fourier_b200 import shard
    for total in (0, 1, 7, 8, 1 << 20):
        for world in (1, 2, 3, 8):
            spans = [shard.split_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (s0, c0), (s1, _) in zip(spans, spans[1:]):
                assert s0 + c0 == s1
    with pytest.raises(ValueError):
        shard.split_range(8, 2, 2)


@pytest.mark.timeout(300)
def test_column_and_row_shards_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 300)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert sorted(r for r, _, _ in res) == [0, 1]
    assert all(a and b for _, a, b in res)
