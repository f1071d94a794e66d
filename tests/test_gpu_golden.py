"""GPU parity against the committed golden fixtures (tests/golden/*.json) and, at
sizes the Python oracle cannot reach, against the C oracle (oracle/c/oracle.c)."""
import json
import os

import pytest

from util import R, SECRET_X, SECRET_Y, bls, fr_list_to_bytes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_golden_piano_small(gpu):
    from fourier_b200 import PianoBackend
    import hashlib
    g = json.load(open(os.path.join(GOLDEN, "piano_small.json")))
    b = PianoBackend(g["n"], g["m"])
    b.setup(bytes.fromhex(g["secret_x"]), bytes.fromhex(g["secret_y"]))
    assert hashlib.sha256(b.setup_to_bytes(True)).hexdigest() == g["setup_compressed_sha256"]
    alpha, beta = bytes.fromhex(g["alpha"]), bytes.fromhex(g["beta"])
    commits, ys, pis = [], [], []
    for row in g["rows"]:
        coeffs = b"".join(bytes.fromhex(h) for h in row["coeffs"])
        c = b.worker_commit(row["i"], coeffs)
        y, pi = b.worker_open(row["i"], coeffs, alpha)
        assert c.hex() == row["commit"] and y.hex() == row["y"] and pi.hex() == row["proof"]
        commits.append(c)
        ys.append(y)
        pis.append(pi)
    assert b.master_commit(commits).hex() == g["master_commit"]
    z, (p0, p1) = b.master_open(ys, pis, beta)
    assert z.hex() == g["z"] and p0.hex() == g["pi0"] and p1.hex() == g["pi1"]
    fin = b"".join(bytes.fromhex(h) for h in g["fft16_in"])
    assert b.ntt(fin).hex() == "".join(g["fft16_out"])
    b.close()


def test_reference_kat_on_gpu(gpu):
    from fourier_b200 import PianoBackend
    kat = json.load(open(os.path.join(GOLDEN, "blst_kat.json")))
    b = PianoBackend(4, 0)
    poly = b"".join(bytes.fromhex(h) for h in kat["poly"])
    assert b.evaluate(poly, bytes.fromhex(kat["point"])).hex() == kat["eval"]
    b.close()


@pytest.mark.parametrize("log_n", [14, 20, 22])
def test_ntt_vs_c_oracle(gpu, log_n):
    from fourier_b200 import PianoBackend, synth
    from oracle import cref
    n = 1 << log_n
    b = PianoBackend(log_n, 0)
    data = synth.fr_bytes(40 + log_n, n)
    for inv in (False, True):
        assert b.ntt(data, inverse=inv) == cref.ntt(data, inv)
    b.close()


def test_msm_2_18_vs_c_oracle(gpu):
    """arbitrary (compressed-wire) points, untabulated and tabulated, vs the C Pippenger"""
    from fourier_b200 import PianoBackend, synth
    from oracle import cref
    n = 1 << 18
    b = PianoBackend(4, 0)
    pts96 = cref.g1_mul_gen(synth.fr_bytes(51, n))
    ks = synth.fr_bytes(52, n)
    want = cref.msm(pts96, ks)
    bases = b.bases_from_points(pts96, 96)
    assert b.msm(bases, ks) == want
    bases.precompute()
    assert b.msm(bases, ks) == want
    b.close()
