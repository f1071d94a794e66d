"""Deterministic synthetic inputs: SplitMix64 -> uniform 255-bit
values reduced into [0, r), emitted as 32-byte big-endian Fr strings.  numpy only;
used by bench.py and the tests to build identical inputs on every rank.
"""
from __future__ import annotations

import numpy as np

R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
_R_WORDS = [(R >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)]  # little-endian u64 words


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def fr_words(seed: int, n: int, start: int = 0) -> np.ndarray:
    """(n, 4) uint64 little-endian words of values in [0, r)."""
    idx = np.arange(start, start + n, dtype=np.uint64)
    base = (np.uint64((seed & 0xFFFFFF) << 40) + idx * np.uint64(4))
    w = np.stack([_splitmix64(base + np.uint64(k)) for k in range(4)], axis=1)
    w[:, 3] &= np.uint64((1 << 63) - 1)  # 255 bits
    # subtract r once where value >= r (r > 2^254, so once is enough)
    ge = np.zeros(n, dtype=bool)
    decided = np.zeros(n, dtype=bool)
    for k in (3, 2, 1, 0):
        rk = np.uint64(_R_WORDS[k])
        gt = (w[:, k] > rk) & ~decided
        lt = (w[:, k] < rk) & ~decided
        ge |= gt
        decided |= gt | lt
    ge |= ~decided  # equal to r
    borrow = np.zeros(n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k in range(4):
            rk = np.uint64(_R_WORDS[k])
            a = w[:, k]
            d = a - rk - borrow
            nb = ((a < rk) | ((a == rk) & (borrow > 0))).astype(np.uint64)
            w[:, k] = np.where(ge, d, a)
            borrow = nb
    return w


def fr_bytes(seed: int, n: int, start: int = 0) -> np.ndarray:
    """n Fr elements as an (n, 32) uint8 array, big-endian (the wire encoding)."""
    w = fr_words(seed, n, start)
    be = w[:, ::-1].astype(">u8")  # most significant word first, big-endian bytes
    return np.ascontiguousarray(be).view(np.uint8).reshape(n, 32)


def fr_ints(seed: int, n: int, start: int = 0):
    b = fr_bytes(seed, n, start)
    return [int.from_bytes(b[i].tobytes(), "big") for i in range(n)]
