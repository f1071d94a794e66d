"""GPU parity: Fr NTT / evaluate / quotient against the oracle (bit-exact).
Mirrors the reference's own FFT-domain tests (src/engine/piano.rs:1276-1390:
roots of unity, Lagrange delta property) and the evaluate RPC test
(src/rpc.rs:713-758)."""
import random

import pytest

from util import R, bls, bytes_to_fr_list, fr_list_to_bytes, opiano

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be(gpu):
    from fourier_b200 import PianoBackend
    b = PianoBackend(12, 2)
    yield b
    b.close()


@pytest.mark.parametrize("log_n", list(range(0, 13)))
def test_ntt_matches_oracle(be, log_n):
    rng = random.Random(100 + log_n)
    n = 1 << log_n
    data = [rng.randrange(R) for _ in range(n)]
    for inverse in (False, True):
        got = bytes_to_fr_list(be.ntt(fr_list_to_bytes(data), inverse=inverse))
        want = opiano.fft_fr(data, inverse, log_n)
        assert got == want, (log_n, inverse)


def test_ntt_edge_values(be):
    n = 64
    for data in ([0] * n, [R - 1] * n, [1] + [0] * (n - 1), [0] * (n - 1) + [R - 1]):
        for inverse in (False, True):
            got = bytes_to_fr_list(be.ntt(fr_list_to_bytes(data), inverse=inverse))
            assert got == opiano.fft_fr(data, inverse, 6)


def test_fft_left_right_domains(be):
    # PianoFFTSettings: left = 2^(n-m) = 1024, right = 2^m = 4 (piano.rs:1061-1077)
    rng = random.Random(5)
    d = [rng.randrange(R) for _ in range(4)]
    assert bytes_to_fr_list(be.fft(fr_list_to_bytes(d), inverse=True, left=False)) == opiano.fft_fr(d, True, 2)
    d = [rng.randrange(R) for _ in range(256)]  # shorter than the left width: strided roots
    assert bytes_to_fr_list(be.fft(fr_list_to_bytes(d), inverse=False, left=True)) == opiano.fft_fr(d, False, 10)
    from fourier_b200 import FourierError
    with pytest.raises(FourierError):
        be.fft(fr_list_to_bytes([1] * 8), left=False)  # longer than the right domain
    with pytest.raises(FourierError):
        be.fft(fr_list_to_bytes([1] * 3))  # not a power of two


def test_lagrange_delta_property(be):
    # piano.rs:1336-1390: iFFT(e_i) evaluates to delta_ik on w^k
    n = 16
    w = bls.root_of_unity(4)
    for i in (0, 3, 15):
        e = [0] * n
        e[i] = 1
        poly = bytes_to_fr_list(be.ntt(fr_list_to_bytes(e), inverse=True))
        for k in range(n):
            y = int.from_bytes(be.evaluate(fr_list_to_bytes(poly), pow(w, k, R)), "big")
            assert y == (1 if k == i else 0)


def test_invalid_scalar_rejected(be):
    from fourier_b200 import FourierError
    bad = (R).to_bytes(32, "big")
    with pytest.raises(FourierError) as ei:
        be.ntt(bad + bytes(32))
    assert "Invalid scalar" in str(ei.value)


@pytest.mark.parametrize("n", [1, 2, 3, 31, 32, 33, 1000, 1024, 1025, 40000])
def test_evaluate_matches_horner(be, n):
    rng = random.Random(n)
    poly = [rng.randrange(R) for _ in range(n)]
    for x in (0, 1, R - 1, rng.randrange(R)):
        got = int.from_bytes(be.evaluate(fr_list_to_bytes(poly), x), "big")
        assert got == opiano.poly_eval(poly, x)


def test_evaluate_rpc_case(be):
    # rpc.rs:713-758: sum i*x^i at x = 1
    poly = list(range(1, 9))
    assert int.from_bytes(be.evaluate(fr_list_to_bytes(poly), 1), "big") == sum(poly)


@pytest.mark.parametrize("log_n", [16, 18, 20])
def test_large_ntt_roundtrip_and_spot(be, log_n):
    from fourier_b200 import synth
    n = 1 << log_n
    data = synth.fr_bytes(11, n)
    fwd = be.ntt(data, inverse=False)
    back = be.ntt(fwd, inverse=True)
    assert back == data.tobytes()
    # spot-check a few outputs against the definition out[k] = p(w^k)
    ints = bytes_to_fr_list(data.tobytes())
    w = bls.root_of_unity(log_n)
    for k in (0, 1, n // 2 + 3, n - 1):
        want = opiano.poly_eval(ints, pow(w, k, R))
        assert int.from_bytes(fwd[32 * k:32 * k + 32], "big") == want
