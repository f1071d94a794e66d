"""How the path shards over the GPUs of one node (one process per GPU, torch.distributed).

Two natural partitions, neither with bulk traffic:

* point (column) ranges of ONE MSM: rank k owns columns [k*T/G, (k+1)*T/G) of every
  resident SRS row, runs the MSM over its slice and contributes one uncompressed 96-byte
  partial sum; the partials are all-gathered and added (fb_g1_sum, format 96);
* independent bivariate rows: rank k owns rows [k*M/G, (k+1)*M/G); per-row results
  (48-byte commitment, 32-byte evaluation, 48-byte proof) are all-gathered and the master
  step (size M) runs on rank 0.

The helpers below are pure host logic (used by bench.py / bench_workloads.py and covered by
world_size-2 gloo tests on CPU); the collective is whatever backend the process group has.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def split_range(total: int, world: int, rank: int) -> Tuple[int, int]:
    """(start, count) of rank's contiguous share; the remainder goes to the low ranks."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, rem = divmod(total, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def all_gather_bytes(dist, blob: bytes, device=None) -> List[bytes]:
    """All-gather one equal-length byte string per rank -> list indexed by rank."""
    import torch
    world = dist.get_world_size()
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    if device is not None:
        mine = mine.to(device)
    out = torch.empty(world * len(blob), dtype=torch.uint8, device=mine.device)
    dist.all_gather_into_tensor(out, mine)
    flat = bytes(out.cpu().numpy())
    return [flat[k * len(blob):(k + 1) * len(blob)] for k in range(world)]


def sharded_commit_open(be, dist, i: int, coeffs_slice, length: int, alpha, device=None, combine=None):
    """Commit + open of row i with BOTH halves sharded by columns: every rank passes only its slice of the row.

    begin (partial sums over the slice) -> all-gather of the 128-byte states -> finish (y, partial commitment and partial
    proof over the slice) -> all-gather of the 224-byte partials -> sum.  `be` is this rank's column-sharded backend;
    `combine(parts)` adds the gathered partials (default: the backend's host-side sum).  Returns (commitment48, y32, proof48),
    identical on every rank and to the unsharded worker_commit / worker_open."""
    state = be.commit_open_shard_begin(i, coeffs_slice, length, alpha)
    states = b"".join(all_gather_bytes(dist, state, device))
    c96, y, p96 = be.commit_open_shard_finish(i, states, alpha)
    parts = all_gather_bytes(dist, c96 + y + p96, device)
    if combine is not None:
        return combine(parts)
    return (be.g1_sum(b"".join(p[:96] for p in parts), fmt=96), y, be.g1_sum(b"".join(p[128:224] for p in parts), fmt=96))


def pack_rows(commits: Sequence[bytes], ys: Sequence[bytes], proofs: Sequence[bytes]) -> bytes:
    """Per-rank row results as one blob: all commitments, then all evaluations, then all proofs."""
    assert len(commits) == len(ys) == len(proofs)
    return b"".join(commits) + b"".join(ys) + b"".join(proofs)


def unpack_rows(blobs: Sequence[bytes], rows_per_rank: Sequence[int]):
    """Inverse of pack_rows over the gathered blobs, rows back in global order."""
    cs: List[bytes] = []
    ys: List[bytes] = []
    ps: List[bytes] = []
    for blob, rows in zip(blobs, rows_per_rank):
        assert len(blob) >= rows * 128
        cs += [blob[i * 48:(i + 1) * 48] for i in range(rows)]
        ys += [blob[rows * 48 + i * 32:rows * 48 + (i + 1) * 32] for i in range(rows)]
        ps += [blob[rows * 80 + i * 48:rows * 80 + (i + 1) * 48] for i in range(rows)]
    return cs, ys, ps
