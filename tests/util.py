"""Shared helpers for the parity tests (the ORACLE side lives in oracle/)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import bls12_381 as bls  # noqa: E402
from oracle import piano as opiano  # noqa: E402

R = bls.R


def fr_list_to_bytes(v):
    return b"".join(int(x).to_bytes(32, "big") for x in v)


def bytes_to_fr_list(b):
    return [int.from_bytes(b[i:i + 32], "big") for i in range(0, len(b), 32)]


SECRET_X = bytes([1] * 32)
SECRET_Y = bytes([2] * 32)
