"""CPU tests of the PRODUCT's field / curve routines (fourier_b200/csrc/*.cuh)
compiled for the host, where each PTX carry-chain primitive is emulated
(limbs.cuh).  They check the very code the kernels inline - the even/odd
carry-chain Montgomery multiplier, XYZZ group law, wire codecs - against the
Python big-int oracle, without a GPU."""
import ctypes
import os
import random
import subprocess

import pytest

from util import ROOT, bls

SRC = os.path.join(ROOT, "tests", "host_arith.cpp")
OUT = os.path.join(ROOT, "tests", "_build", "libhost_arith.so")


@pytest.fixture(scope="module")
def L():
    hdrs = [os.path.join(ROOT, "fourier_b200", "csrc", f) for f in ("limbs.cuh", "field.cuh", "field_mul.cuh", "g1.cuh")]
    newest = max(os.path.getmtime(p) for p in hdrs + [SRC])
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-x", "c++", "-shared", "-fPIC", "-o", OUT, SRC])
    return ctypes.CDLL(OUT)


def arr(v, n):
    return (ctypes.c_uint32 * n)(*[(v >> (32 * i)) & 0xFFFFFFFF for i in range(n)])


def val(a):
    return sum(int(x) << (32 * i) for i, x in enumerate(a))


@pytest.mark.parametrize("name,m,n", [("fr", bls.R, 8), ("fp", bls.P, 12)])
def test_field_ops(L, name, m, n):
    rng = random.Random(n)
    RR = 1 << (32 * n)
    Ri = pow(RR, -1, m)
    mul, mulp, op = getattr(L, "h_%s_mul" % name), getattr(L, "h_%s_mul_portable" % name), getattr(L, "h_%s_op" % name)
    cases = [(0, 0), (1, 1), (m - 1, m - 1), (m - 1, 1), (RR % m, RR % m), (m - 1, 0)]
    cases += [(rng.randrange(m), rng.randrange(m)) for _ in range(2000)]
    cases += [((1 << k) % m, m - 1 - ((1 << j) % m)) for k in range(0, 32 * n, 37) for j in range(0, 32 * n, 53)]
    # rows whose Montgomery quotient digit is 0 (low limbs of the running sum vanish): the m == 0 branch of
    # Fr's multiplier-free low limbs (fe_mont_step)
    cases += [(1 << (32 * i), (1 << (32 * j)) + (rng.randrange(m) >> (32 * k) << (32 * k))) for i in range(n - 1) for j in range(n - 1) for k in (1, 3)]
    cases = [(a % m, b % m) for a, b in cases]
    r = (ctypes.c_uint32 * n)()
    for a, b in cases:
        e = a * b * Ri % m
        mul(arr(a, n), arr(b, n), r)
        assert val(r) == e
        mulp(arr(a, n), arr(b, n), r)
        assert val(r) == e
        op(0, arr(a, n), arr(b, n), r)
        assert val(r) == (a + b) % m
        op(1, arr(a, n), arr(b, n), r)
        assert val(r) == (a - b) % m
        op(2, arr(a, n), arr(b, n), r)
        assert val(r) == (-a) % m
    # dedicated Montgomery squaring (fe_sqr): every operand of the list above, plus limb patterns that drive the
    # carries between the two product arrays and the lone-limb carry of the reduction rows
    sqr = getattr(L, "h_%s_sqr" % name)
    sq_cases = [a for a, _ in cases] + [b for _, b in cases]
    sq_cases += [((1 << (32 * k)) - 1) % m for k in range(1, n + 1)] + [(m - (1 << (32 * k))) % m for k in range(n)]
    sq_cases += [int("ffffffff00000000" * (n // 2), 16) % m, int("00000000ffffffff" * (n // 2), 16) % m,
                 int("ffffffff" * n, 16) % m, int("80000000" * n, 16) % m, int("00000001" * n, 16) % m]
    for a in sq_cases:
        sqr(arr(a, n), r)
        assert val(r) == a * a * Ri % m, hex(a)
    for a, b in cases[:30]:
        op(3, arr(a, n), arr(b, n), r)
        assert val(r) == (pow(a * Ri % m, -1, m) * RR % m if a else 0)
        op(4, arr(a, n), arr(b, n), r)
        assert val(r) == a * RR % m
        op(5, arr(a, n), arr(b, n), r)
        assert val(r) == a * Ri % m
    # single-thread inversions (shift-and-subtract, op 6; division-step batches, op 7) == Fermat == big-int
    edge = [2, 3, m - 2, (m - 1) // 2, (m + 1) // 2, 1 << 30, (1 << 30) - 1, 1 << 60, (1 << (32 * n - 3)) % m,
            RR % m, Ri, (m - RR) % m, pow(2, 32 * n - 1, m)]
    edge += [(1 << k) % m for k in range(1, 32 * n, 29)] + [(m - (1 << k)) % m for k in range(0, 32 * n - 4, 31)]
    inv_cases = cases[:2500] + [(k, 0) for k in edge]
    for a, b in inv_cases[:600] + inv_cases[-60:]:
        op(6, arr(a, n), arr(b, n), r)
        assert val(r) == (pow(a * Ri % m, -1, m) * RR % m if a else 0), hex(a)
    for a, b in inv_cases:
        op(7, arr(a, n), arr(b, n), r)
        assert val(r) == (pow(a * Ri % m, -1, m) * RR % m if a else 0), hex(a)


def cb(b):
    return (ctypes.c_uint8 * len(b))(*b)


def test_g1_codecs_and_group_law(L):
    rng = random.Random(4)
    o = (ctypes.c_uint8 * 48)()
    L.h_g1_generator(o)
    assert bytes(o) == bls.g1_compress(bls.G1_GEN)
    pts = [bls.g1_mul(bls.G1_GEN, rng.randrange(bls.R)) for _ in range(8)] + [bls.G1_INF]
    for p in pts:
        b = bls.g1_compress(p)
        assert L.h_g1_decompress_compress(cb(b), o) == 0 and bytes(o) == b
        o96 = (ctypes.c_uint8 * 96)()
        assert L.h_g1_serialize_roundtrip(cb(b), o96, o) == 0
        assert bytes(o96) == bls.g1_serialize(p) and bytes(o) == b
    for p in pts:
        for q in pts + [p, bls.g1_neg(p)]:
            for mode in (0, 1):
                assert L.h_g1_add(mode, cb(bls.g1_compress(p)), cb(bls.g1_compress(q)), o) == 0
                assert bytes(o) == bls.g1_compress(bls.g1_add(p, q))
        L.h_g1_add(2, cb(bls.g1_compress(p)), cb(bls.g1_compress(p)), o)
        assert bytes(o) == bls.g1_compress(bls.g1_double(p))
        k = rng.randrange(bls.R)
        L.h_g1_mul(cb(bls.g1_compress(p)), arr(k, 8), o)
        assert bytes(o) == bls.g1_compress(bls.g1_mul(p, k))


def test_g1_decode_rejects(L):
    o = (ctypes.c_uint8 * 48)()
    assert L.h_g1_decompress_compress(cb(bytes(48)), o) != 0  # compression bit missing
    bad_inf = bytes([0xE0]) + bytes(47)  # infinity with the sign bit
    assert L.h_g1_decompress_compress(cb(bad_inf), o) != 0
    too_big = bytes([0x9F]) + bytes([0xFF] * 47)  # x >= p
    assert L.h_g1_decompress_compress(cb(too_big), o) != 0
    # find an x that is not on the curve
    for x in range(1, 50):
        enc = bytearray(x.to_bytes(48, "big"))
        enc[0] |= 0x80
        try:
            bls.g1_decompress(bytes(enc))
        except ValueError:
            assert L.h_g1_decompress_compress(cb(bytes(enc)), o) != 0
            break
