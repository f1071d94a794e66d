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


class OracleShardBackend:
    """The sharded commit + open protocol (begin -> states -> finish -> partials) restated on the ORACLE's big-int
    arithmetic with a known-tau SRS (m = 0): test infrastructure for the host logic of shard.sharded_commit_open."""

    def __init__(self, o, bls, col0, ncols):
        self.o, self.bls, self.col0, self.ncols = o, bls, col0, ncols
        self.slice = None

    def _span(self, length):
        lo, hi = min(self.col0, length), min(self.col0 + self.ncols, length)
        return lo, hi

    def commit_open_shard_begin(self, i, coeffs_slice, length, alpha):
        R = self.bls.R
        lo, hi = self._span(length)
        assert len(coeffs_slice) == hi - lo
        w = self.bls.root_of_unity(length.bit_length() - 1)
        s1 = s0 = ck = 0
        k = -1
        self.inv = []
        for jj, c in enumerate(coeffs_slice):
            wj = pow(w, lo + jj, R)
            d = (wj - alpha) % R
            if d == 0:
                k, ck = lo + jj, c
                self.inv.append(0)
                continue
            inv = pow(d, -1, R)
            self.inv.append(inv)
            s1 = (s1 + c * wj * inv) % R
            s0 = (s0 + wj * inv) % R
        self.slice, self.length = list(coeffs_slice), length
        return b"".join(v.to_bytes(32, "big") for v in (s1, s0, ck)) + (k % (1 << 64)).to_bytes(8, "big") + bytes(24)

    def commit_open_shard_finish(self, i, states, alpha):
        bls, R, length = self.bls, self.bls.R, self.length
        lo, hi = self._span(length)
        s1 = s0 = ck = 0
        k = -1
        for p in range(len(states) // 128):
            st = states[128 * p:128 * p + 128]
            s1 = (s1 + int.from_bytes(st[:32], "big")) % R
            s0 = (s0 + int.from_bytes(st[32:64], "big")) % R
            kk = int.from_bytes(st[96:104], "big")
            if kk != (1 << 64) - 1 and k < 0:
                k, ck = kk, int.from_bytes(st[64:96], "big")
        w = bls.root_of_unity(length.bit_length() - 1)
        if k < 0:
            y = (-(pow(alpha, length, R) - 1) * pow(length, -1, R) * s1) % R
        else:
            y = ck
        qhat = []
        for jj, c in enumerate(self.slice):
            if lo + jj == k:
                qhat.append((-pow(w, -k, R) * (s1 - c * s0)) % R)
            else:
                qhat.append((c - y) * self.inv[jj] % R)
        lag = self.o.settings.l_at_tau  # length == T here
        part = lambda v: bls.g1_serialize(bls.g1_mul(bls.G1_GEN, sum(a * lag[lo + jj] for jj, a in enumerate(v)) % R))
        return part(self.slice), y.to_bytes(32, "big"), part(qhat)

    def g1_sum(self, blob, fmt=96):
        total = self.bls.G1_INF
        for k in range(len(blob) // fmt):
            total = self.bls.g1_add(total, self.bls.g1_deserialize(blob[fmt * k:fmt * (k + 1)]))
        return self.bls.g1_compress(total)


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
        # (3) commit + open with the field half sharded too: begin -> all-gather states -> finish -> all-gather partials
        from oracle import piano as opiano
        n = 4
        o = opiano.PianoOracle(n, 0)
        o.generate_trusted_setup(bytes([1] * 32), bytes([2] * 32), with_points=False)
        row = [rng.randrange(R) for _ in range(o.T)]
        ok_shard = True
        for alpha in (rng.randrange(R), pow(bls.root_of_unity(n), 11, R)):  # outside the domain; on it (index 11 -> rank 1's slice)
            col0, ncols = shard.split_range(o.T, world, rank)
            be = OracleShardBackend(o, bls, col0, ncols)
            got = shard.sharded_commit_open(be, dist, 0, row[col0:col0 + ncols], o.T, alpha)
            oy, opi = o.worker_open_known_tau(0, row, alpha)
            want = (bls.g1_compress(o.worker_commit_known_tau(0, row)), oy.to_bytes(32, "big"), bls.g1_compress(opi))
            ok_shard = ok_shard and got == want
        q.put((rank, ok_msm, ok_rows and ok_shard))
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
