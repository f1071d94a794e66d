"""The host-side pairing product behind worker_verify / master_verify
(reference src/engine/piano.rs:358-388, 422-464: blst raw_aggregate + final_exp +
fp12_is_one).  It is host code inside libfourier_b200.so, so it is checked here
without a GPU, against the oracle's independent restatement (flat Fp12, affine
Miller loop, generic (p^12-1)/r power)."""
import random

import pytest

from util import R, SECRET_X, SECRET_Y, bls, opiano

P = bls.P


def _lib_or_skip():
    try:
        from fourier_b200 import engine
        from fourier_b200 import _lib
        _lib.load()
        return engine
    except OSError as e:  # pragma: no cover
        pytest.skip("libfourier_b200.so not built: %s" % e)


def s1(p):
    return bls.g1_serialize(p)


def s2(q):
    return bls.g2_serialize(q)


def test_bilinearity_and_rejection():
    eng = _lib_or_skip()
    rng = random.Random(5)
    g1, g2 = bls.G1_GEN, bls.G2_GEN
    for _ in range(2):
        a, b = rng.randrange(1, R), rng.randrange(1, R)
        pa, qb = bls.g1_mul(g1, a), bls.g2_mul(g2, b)
        # e(aP, bQ) = e(abP, Q)
        assert eng.pairing_check([s1(pa), s1(bls.g1_neg(bls.g1_mul(g1, a * b % R)))], [s2(qb), s2(g2)])
        assert not eng.pairing_check([s1(pa), s1(bls.g1_neg(bls.g1_mul(g1, (a * b + 1) % R)))], [s2(qb), s2(g2)])
    # three factors: e(P,Q)^a e(P,Q)^b e(-(a+b)P, Q) = 1
    a, b = 17, R - 5
    assert eng.pairing_check([s1(bls.g1_mul(g1, a)), s1(g1), s1(bls.g1_neg(bls.g1_mul(g1, (a + b) % R)))],
                             [s2(g2), s2(bls.g2_mul(g2, b)), s2(g2)])
    assert not eng.pairing_check([s1(g1)], [s2(g2)])  # e(g1, g2) generates GT


def test_identity_factors_are_one():
    eng = _lib_or_skip()
    assert eng.pairing_check([], [])
    assert eng.pairing_check([s1(bls.G1_INF)], [s2(bls.G2_GEN)])
    assert eng.pairing_check([s1(bls.G1_GEN)], [s2(None)])
    # an identity factor does not disturb the others
    a = 99
    assert eng.pairing_check([s1(bls.G1_INF), s1(bls.g1_mul(bls.G1_GEN, a)), s1(bls.g1_neg(bls.G1_GEN))],
                             [s2(bls.G2_GEN), s2(bls.G2_GEN), s2(bls.g2_mul(bls.G2_GEN, a))])


def test_value_matches_oracle_pairing():
    """(e(P,Q))^3 coefficient by coefficient.  The oracle's Miller loop ignores the
    sign of the curve parameter, so its GT value is the inverse (= conjugate)."""
    eng = _lib_or_skip()
    a, b = 0x1234567890ABCDEF1234567890ABCDEF, 0xFEDCBA0987654321
    p1, q2 = bls.g1_mul(bls.G1_GEN, a), bls.g2_mul(bls.G2_GEN, b)
    v = bls.f12_pow(bls.miller_loop(bls.g1_to_affine(p1), q2), bls.FINAL_EXP)
    vi = [(c if k % 2 == 0 else (-c) % P) for k, c in enumerate(v)]  # w -> -w
    assert bls.f12_mul(v, vi) == bls.F12_ONE
    want = bls.f12_mul(bls.f12_mul(vi, vi), vi)
    out = eng.pairing_value([s1(p1)], [s2(q2)])
    flat = [0] * 12  # tower coefficient (c0 + c1 u) of w^k -> Fp[w]/(w^12 - 2w^6 + 2), u = w^6 - 1
    for k in range(6):
        c0 = int.from_bytes(out[96 * k:96 * k + 48], "big")
        c1 = int.from_bytes(out[96 * k + 48:96 * k + 96], "big")
        flat[k] = (flat[k] + c0 - c1) % P
        flat[k + 6] = (flat[k + 6] + c1) % P
    assert flat == want


def test_kzg_equations_on_oracle_proofs():
    """The two verify equations of piano.rs:309-356 and :393-420 fed with the
    oracle's commitments and proofs: valid ones pass, tampered ones fail."""
    eng = _lib_or_skip()
    o = opiano.PianoOracle(4, 1)
    o.generate_trusted_setup(SECRET_X, SECRET_Y)
    s = o.settings
    rng = random.Random(21)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha, beta = rng.randrange(R), rng.randrange(R)
    commits = [o.worker_commit(i, rows[i]) for i in range(o.M)]
    opens = [o.worker_open(i, rows[i], alpha) for i in range(o.M)]
    g2x = bls.g2_sub(s.g2_tau_x, bls.g2_mul(s.g2, alpha))
    g2y = bls.g2_sub(s.g2_tau_y, bls.g2_mul(s.g2, beta))
    # worker: e(-(C - y R_i(tau_Y) g), g2) e(pi, g2^{tau_X - alpha}) == 1
    i = 1
    y, pi = opens[i]
    r_i = bls.g1_msm(s.g_tau_y, o.right_lagrange_poly(i))
    for yy, want in ((y, True), ((y + 1) % R, False)):
        lhs = bls.g1_add(commits[i], bls.g1_neg(bls.g1_mul(r_i, yy)))
        assert eng.pairing_check([s1(bls.g1_neg(lhs)), s1(pi)], [s2(s.g2), s2(g2x)]) is want
    # master
    com = o.master_commit(commits)
    z, (pi0, pi1) = o.master_open([y for y, _ in opens], [p for _, p in opens], beta)
    for zz, want in ((z, True), ((z + 1) % R, False)):
        lhs = bls.g1_add(com, bls.g1_neg(bls.g1_mul(s.g, zz)))
        assert eng.pairing_check([s1(bls.g1_neg(lhs)), s1(pi0), s1(pi1)], [s2(s.g2), s2(g2x), s2(g2y)]) is want


def test_malformed_points_are_errors():
    eng = _lib_or_skip()
    from fourier_b200.engine import FourierError
    bad1 = bytearray(s1(bls.G1_GEN))
    bad1[95] ^= 1  # off the curve
    with pytest.raises(FourierError):
        eng.pairing_check([bytes(bad1)], [s2(bls.G2_GEN)])
    bad2 = bytearray(s2(bls.G2_GEN))
    bad2[191] ^= 1
    with pytest.raises(FourierError):
        eng.pairing_check([s1(bls.G1_GEN)], [bytes(bad2)])
    with pytest.raises(FourierError):
        eng.pairing_check([s1(bls.G1_GEN)], [b"\x00" * 96])
