"""Regenerates the golden fixtures in this directory.  Run in the BUILD container
(where /root/reference exists); the fixtures are committed because the reference
does not travel to the GPU box.

  blst_kat.json     - the only numeric known-answer vector in the reference,
                      transcribed by regex from src/engine/blst.rs:304-324
                      (TEST_POLY, TEST_POINT, TEST_EVAL).
  piano_small.json  - seeded commit/open/master outputs of the Python oracle at
                      n=5, m=2 (the reference itself has no golden G1 bytes and
                      cannot be run here: no Rust toolchain).
"""
import json
import os
import random
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import bls12_381 as bls  # noqa: E402
from oracle import piano  # noqa: E402

REF = "/root/reference/src/engine/blst.rs"


def kat():
    src = open(REF).read()
    block = src[src.index("const TEST_POLY"):src.index("];", src.index("const TEST_POLY"))]
    poly = re.findall(r'"([0-9A-Fa-f]{64})"', block)
    point = re.search(r'const TEST_POINT: &str = "([0-9a-fA-F]{64})"', src).group(1)
    ev = re.search(r'const TEST_EVAL: &str = "([0-9a-fA-F]{64})"', src).group(1)
    assert len(poly) == 16
    out = {"source": "reference src/engine/blst.rs:304-324", "poly": poly, "point": point, "eval": ev}
    json.dump(out, open(os.path.join(HERE, "blst_kat.json"), "w"), indent=1)


def piano_small():
    n, m = 5, 2
    sx, sy = bytes([1] * 32), bytes([2] * 32)
    o = piano.PianoOracle(n, m)
    o.generate_trusted_setup(sx, sy)
    rng = random.Random(20261104)
    alpha, beta = rng.randrange(bls.R), rng.randrange(bls.R)
    rows = []
    commits, opens = [], []
    for i in range(o.M):
        coeffs = [rng.randrange(bls.R) for _ in range(o.T)]
        c = o.worker_commit(i, coeffs)
        y, pi = o.worker_open(i, coeffs, alpha)
        commits.append(c)
        opens.append((y, pi))
        rows.append({"i": i, "coeffs": ["%064x" % v for v in coeffs], "commit": bls.g1_compress(c).hex(),
                     "y": "%064x" % y, "proof": bls.g1_compress(pi).hex()})
    com = o.master_commit(commits)
    z, (pi0, pi1) = o.master_open([y for y, _ in opens], [p for _, p in opens], beta)
    assert o.master_verify(com, beta, alpha, z, (pi0, pi1))
    out = {"n": n, "m": m, "secret_x": sx.hex(), "secret_y": sy.hex(), "alpha": "%064x" % alpha, "beta": "%064x" % beta,
           "rows": rows, "master_commit": bls.g1_compress(com).hex(), "z": "%064x" % z,
           "pi0": bls.g1_compress(pi0).hex(), "pi1": bls.g1_compress(pi1).hex(),
           "setup_compressed_sha256": __import__("hashlib").sha256(piano.setup_to_bytes(o.settings, True)).hexdigest(),
           "fft16_in": ["%064x" % v for v in range(1, 17)],
           "fft16_out": ["%064x" % v for v in piano.fft_fr(list(range(1, 17)), False, 4)]}
    json.dump(out, open(os.path.join(HERE, "piano_small.json"), "w"), indent=1)


if __name__ == "__main__":
    kat()
    piano_small()
    print("golden fixtures written")
