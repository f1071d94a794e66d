"""CPU tests that PIN the oracle (no GPU): the reference's only numeric KAT, the
well-known generator encoding, the algebraic identities the reference's own tests
assert (src/engine/piano.rs:1276-1390, 1414-1459, 1478-1563, 1604-1722), and the C
oracle against the Python big-int oracle."""
import json
import os
import random

import pytest

from util import R, SECRET_X, SECRET_Y, bls, bytes_to_fr_list, fr_list_to_bytes, opiano
from oracle import cref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_parameters_rederive():
    z = bls.Z_PARAM
    assert bls.R == z**4 - z**2 + 1
    assert bls.P == (z - 1) ** 2 * bls.R // 3 + z
    assert bls.P % 4 == 3


def test_generator_encoding_and_order():
    g = bls.g1_compress(bls.G1_GEN)
    assert g.hex().startswith("97f1d3a73197d7942695638c4fa9ac0f") and g.hex().endswith("fb3af00adb22c6bb")
    assert bls.g1_on_curve_affine((bls.G1_X, bls.G1_Y))
    assert bls.g1_is_inf(bls.g1_mul(bls.G1_GEN, R - 1)) is False
    assert bls.g1_eq(bls.g1_add(bls.g1_mul(bls.G1_GEN, R - 1), bls.G1_GEN), bls.G1_INF)
    assert bls.g1_decompress(g)[:2] == (bls.G1_X, bls.G1_Y)
    assert bls.g2_on_curve(bls.G2_GEN)


def test_reference_kat_pins_byte_order_and_modulus():
    """blst.rs:304-324: the stale KAT still pins big-endian hex mod r + Horner.
    The golden file holds the vector as transcribed, with the generating script."""
    with open(os.path.join(GOLDEN, "blst_kat.json")) as f:
        kat = json.load(f)
    poly = [int(h, 16) for h in kat["poly"]]
    x = int(kat["point"], 16)
    y = int(kat["eval"], 16)
    assert all(c < R for c in poly) and x < R and y < R
    assert opiano.poly_eval(poly, x) == y
    # little-endian interpretation must NOT satisfy it
    le = [int.from_bytes(bytes.fromhex(h)[::-1], "big") % R for h in kat["poly"]]
    assert opiano.poly_eval(le, int.from_bytes(bytes.fromhex(kat["point"])[::-1], "big") % R) != \
        int.from_bytes(bytes.fromhex(kat["eval"])[::-1], "big") % R
    # the C oracle agrees
    got = cref.evaluate(fr_list_to_bytes(poly), x.to_bytes(32, "big"))
    assert int.from_bytes(got, "big") == y


def test_root_of_unity_constants():
    # scale-2 root is the c-kzg constant 0x8d51ccce...
    assert hex(bls.root_of_unity(2)).startswith("0x8d51ccce760304d0ec030002760300000001000000000000")
    for k in range(0, 12):
        w = bls.root_of_unity(k)
        assert pow(w, 1 << k, R) == 1 and (k == 0 or pow(w, 1 << (k - 1), R) == R - 1)


def test_fft_matches_definition_and_lagrange_delta():
    rng = random.Random(1)
    for log_n in range(0, 7):
        d = [rng.randrange(R) for _ in range(1 << log_n)]
        for inv in (False, True):
            assert opiano.fft_fr(d, inv, log_n) == opiano.fft_fr_slow(d, inv)
    # piano.rs:1336-1390
    o = opiano.PianoOracle(4, 2)
    w = bls.root_of_unity(2)
    for i in range(4):
        poly = o.left_lagrange_poly(i)
        assert [opiano.poly_eval(poly, pow(w, k, R)) for k in range(4)] == [int(k == i) for k in range(4)]


def test_lagrange_closed_form_equals_reference_algorithm():
    for tau in (0, 1, 12345, bls.root_of_unity(3), R - 1):
        for log_n in (0, 1, 3, 5):
            assert opiano.lagrange_at(tau, log_n) == opiano.lagrange_at_literal(tau, log_n)


def test_msm_equals_naive_and_known_tau():
    # manual_commit_test, piano.rs:1414-1459
    o = opiano.PianoOracle(4, 1)
    o.generate_trusted_setup(SECRET_X, SECRET_Y)
    rng = random.Random(2)
    for i in range(o.M):
        c = [rng.randrange(R) for _ in range(o.T)]
        naive = bls.g1_msm_naive(o.settings.u[i], c)
        assert bls.g1_eq(o.worker_commit(i, c), naive)
        assert bls.g1_eq(o.worker_commit_known_tau(i, c), naive)
        y, pi = o.worker_open(i, c, 77)
        y2, pi2 = o.worker_open_known_tau(i, c, 77)
        assert y == y2 and bls.g1_eq(pi, pi2)


def test_protocol_verifies_and_rejects():
    # pianist_test / test_fake_poly, piano.rs:1604-1722, 1898-1955 (one small shape)
    o = opiano.PianoOracle(3, 1)
    o.generate_trusted_setup(SECRET_X, SECRET_Y)
    rng = random.Random(3)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha, beta = rng.randrange(R), rng.randrange(R)
    commits = [o.worker_commit(i, rows[i]) for i in range(o.M)]
    opens = [o.worker_open(i, rows[i], alpha) for i in range(o.M)]
    assert o.worker_verify(1, commits[1], alpha, opens[1][0], opens[1][1])
    assert not o.worker_verify(1, commits[1], alpha, (opens[1][0] + 1) % R, opens[1][1])
    com = o.master_commit(commits)
    z, pf = o.master_open([y for y, _ in opens], [p for _, p in opens], beta)
    assert o.master_verify(com, beta, alpha, z, pf)
    # z == f(beta, alpha): reconstruction tests piano.rs:1478-1563
    polys = [o.fft_left(r, True) for r in rows]
    lag = [o.right_lagrange_poly(i) for i in range(o.M)]
    assert z == sum(opiano.poly_eval(lag[i], beta) * opiano.poly_eval(polys[i], alpha) for i in range(o.M)) % R
    # identity commitment / proof verify for the zero polynomial (piano.rs:1392-1412)
    assert o.worker_verify(0, bls.G1_INF, alpha, 0, bls.G1_INF)


def test_setup_bytes_roundtrip():
    o = opiano.PianoOracle(3, 1)
    s = o.generate_trusted_setup(SECRET_X, SECRET_Y)
    for compressed in (True, False):
        blob = opiano.setup_to_bytes(s, compressed)
        s2 = opiano.setup_from_bytes(blob, compressed)
        assert opiano.setup_to_bytes(s2, compressed) == blob
        assert opiano.setup_to_bytes(s2, not compressed) == opiano.setup_to_bytes(s, not compressed)


def test_golden_vectors_match_python_oracle():
    """tests/golden/piano_small.json was produced by tests/golden/make_golden.py from
    the Python oracle; recompute a sample so the file cannot drift."""
    with open(os.path.join(GOLDEN, "piano_small.json")) as f:
        gold = json.load(f)
    o = opiano.PianoOracle(gold["n"], gold["m"])
    o.generate_trusted_setup(bytes.fromhex(gold["secret_x"]), bytes.fromhex(gold["secret_y"]))
    case = gold["rows"][0]
    coeffs = [int(h, 16) for h in case["coeffs"]]
    assert bls.g1_compress(o.worker_commit(case["i"], coeffs)).hex() == case["commit"]
    y, pi = o.worker_open(case["i"], coeffs, int(gold["alpha"], 16))
    assert "%064x" % y == case["y"] and bls.g1_compress(pi).hex() == case["proof"]


# ------------------------------------------------------------------ C oracle
@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 10, 13])
def test_c_ntt_matches_python(log_n):
    rng = random.Random(log_n)
    d = [rng.randrange(R) for _ in range(1 << log_n)]
    for inv in (False, True):
        assert bytes_to_fr_list(cref.ntt(fr_list_to_bytes(d), inv)) == opiano.fft_fr(d, inv, log_n)


def test_c_open_scalars_eval_lagrange_match_python():
    rng = random.Random(5)
    o = opiano.PianoOracle(9, 0)
    for n in (1, 2, 64, 512):
        c = [rng.randrange(R) for _ in range(n)]
        a = rng.randrange(R)
        oo = opiano.PianoOracle(n.bit_length() - 1, 0)
        y, q = oo.open_scalars(c, a)
        cy, cq = cref.open_scalars(fr_list_to_bytes(c), a.to_bytes(32, "big"))
        assert int.from_bytes(cy, "big") == y and bytes_to_fr_list(cq) == q
        assert int.from_bytes(cref.evaluate(fr_list_to_bytes(c), a.to_bytes(32, "big")), "big") == opiano.poly_eval(c, a)
    for tau in (0, 5, bls.root_of_unity(4), rng.randrange(R)):
        assert bytes_to_fr_list(cref.lagrange_at(tau.to_bytes(32, "big"), 4)) == opiano.lagrange_at(tau, 4)
    with pytest.raises(ValueError):
        cref.ntt((R).to_bytes(32, "big"))


@pytest.mark.parametrize("n", [1, 5, 8, 100, 3000])
def test_c_msm_and_points_match_python(n):
    rng = random.Random(n)
    ss = [rng.randrange(R) for _ in range(n)]
    if n >= 5:
        ss[1] = ss[0]
        ss[2] = (R - ss[0]) % R
        ss[3] = 0
    ks = [rng.randrange(R) for _ in range(n)]
    if n >= 5:
        ks[0] = ks[1] = ks[2] = 9
        ks[4] = R - 1
    pts96 = cref.g1_mul_gen(fr_list_to_bytes(ss))
    for i in range(min(n, 6)):
        assert pts96[96 * i:96 * i + 96] == bls.g1_serialize(bls.g1_mul(bls.G1_GEN, ss[i]))
    want = bls.g1_compress(bls.g1_mul(bls.G1_GEN, sum(s * k for s, k in zip(ss, ks)) % R))
    assert cref.msm(pts96, fr_list_to_bytes(ks)) == want
    pts48 = b"".join(cref.g1_recode(pts96[96 * i:96 * i + 96], 48) for i in range(min(n, 20)))
    assert cref.g1_sum(pts48) == bls.g1_compress(bls.g1_mul(bls.G1_GEN, sum(ss[:20]) % R))


def test_c_bench_entry_points_are_consistent():
    rng = random.Random(11)
    n = 256
    tau = opiano.bls.hash_to_bls_field(SECRET_X)
    l = cref.lagrange_at(tau.to_bytes(32, "big"), 8)
    pts = cref.g1_mul_gen(l)
    coeffs = [rng.randrange(R) for _ in range(n)]
    alpha = rng.randrange(R)
    tc, to, c, y, p = cref.bench_commit_open(pts, fr_list_to_bytes(coeffs), alpha.to_bytes(32, "big"), 1)
    o = opiano.PianoOracle(8, 0)
    o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=False)
    assert c == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
    oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
    assert y == oy.to_bytes(32, "big") and p == bls.g1_compress(opi)
    assert tc > 0 and to > 0
    t, out = cref.bench_msm(pts, fr_list_to_bytes(coeffs), 1)
    assert out == c and t > 0
    assert cref.bench_ntt(fr_list_to_bytes(coeffs), False, 1) > 0
