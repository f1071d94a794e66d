"""GPU parity: the PianoBackend protocol (setup, worker/master commit + open, setup
file round trip) against the oracle's restatement of src/engine/piano.rs, bit-exact
on the reference's wire encodings.  Follows the reference's own end-to-end tests
(piano.rs:1414-1459 manual_commit_test, :1604-1722 pianist_test, :1789-1833
test_save_and_load_setup, :1392-1412 identity inputs)."""
import random

import pytest

from util import R, SECRET_X, SECRET_Y, bls, fr_list_to_bytes, opiano

pytestmark = pytest.mark.gpu


def make(n, m, window_bits=0):
    from fourier_b200 import PianoBackend
    b = PianoBackend(n, m, window_bits=window_bits)
    b.setup(SECRET_X, SECRET_Y)
    o = opiano.PianoOracle(n, m)
    o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=(n <= 6))
    return b, o


@pytest.fixture(scope="module")
def small(gpu):
    b, o = make(5, 2)
    yield b, o
    b.close()


def test_setup_bytes_match_oracle(small):
    b, o = small
    for compressed in (True, False):
        assert b.setup_to_bytes(compressed) == opiano.setup_to_bytes(o.settings, compressed)


def test_worker_commit_open_match_oracle(small):
    b, o = small
    rng = random.Random(9)
    for i in range(o.M):
        coeffs = [rng.randrange(R) for _ in range(o.T)]
        alpha = rng.randrange(R)
        c = b.worker_commit(i, coeffs)
        assert c == bls.g1_compress(o.worker_commit(i, coeffs))
        y, pi = b.worker_open(i, coeffs, alpha)
        oy, opi = o.worker_open(i, coeffs, alpha)
        assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)
        if i == 0:  # pairing check (slow in the oracle): one row is enough
            assert o.worker_verify(i, bls.g1_decompress(c), alpha, oy, bls.g1_decompress(pi))


def test_full_protocol_and_master_verify(small):
    b, o = small
    rng = random.Random(10)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha, beta = rng.randrange(R), rng.randrange(R)
    flat = [x for row in rows for x in row]
    commits = b.commit_rows(0, o.M, flat)
    assert commits == [b.worker_commit(i, rows[i]) for i in range(o.M)]
    ys, pis = b.open_rows(0, o.M, flat, alpha)
    com = b.master_commit(commits)
    z, (pi0, pi1) = b.master_open(ys, pis, beta)
    oc = o.master_commit([o.worker_commit(i, rows[i]) for i in range(o.M)])
    opens = [o.worker_open(i, rows[i], alpha) for i in range(o.M)]
    oz, (opi0, opi1) = o.master_open([y for y, _ in opens], [p for _, p in opens], beta)
    assert com == bls.g1_compress(oc)
    assert z == oz.to_bytes(32, "big")
    assert pi0 == bls.g1_compress(opi0) and pi1 == bls.g1_compress(opi1)
    # z == f(beta, alpha) (piano.rs:1478-1563) and the pairing equation holds
    polys = [o.fft_left(r, True) for r in rows]
    lag = [o.right_lagrange_poly(i) for i in range(o.M)]
    want = sum(opiano.poly_eval(lag[i], beta) * opiano.poly_eval(polys[i], alpha) for i in range(o.M)) % R
    assert oz == want
    assert o.master_verify(bls.g1_decompress(com), beta, alpha, oz, (bls.g1_decompress(pi0), bls.g1_decompress(pi1)))
    assert not o.master_verify(bls.g1_decompress(com), beta, alpha, (oz + 1) % R,
                               (bls.g1_decompress(pi0), bls.g1_decompress(pi1)))


def test_identity_and_degenerate_inputs(small):
    b, o = small
    ident = bls.g1_compress(bls.G1_INF)
    zero = [0] * o.T
    assert b.worker_commit(1, zero) == ident
    y, pi = b.worker_open(1, zero, 5)
    assert y == bytes(32) and pi == ident
    ones = [1] * o.T  # constant evaluations -> constant polynomial, zero quotient
    y, pi = b.worker_open(2, ones, 123456789)
    assert y == (1).to_bytes(32, "big") and pi == ident
    assert b.master_commit([ident] * o.M) == ident
    z, (p0, p1) = b.master_open([0] * o.M, [ident] * o.M, 7)
    assert z == bytes(32) and p0 == ident and p1 == ident
    # alpha on the evaluation domain
    w = bls.root_of_unity(o.t)
    rng = random.Random(4)
    coeffs = [rng.randrange(R) for _ in range(o.T)]
    y, pi = b.worker_open(0, coeffs, w)
    oy, opi = o.worker_open(0, coeffs, w)
    assert y == oy.to_bytes(32, "big") == coeffs[1].to_bytes(32, "big") and pi == bls.g1_compress(opi)


def test_errors(small):
    from fourier_b200 import FourierError
    b, o = small
    with pytest.raises(FourierError):
        b.worker_commit(o.M, [1] * o.T)  # i >= M
    with pytest.raises(FourierError):
        b.worker_open(0, [1] * 3, 1)  # not a power of two
    with pytest.raises(FourierError) as ei:
        b.worker_commit(0, [R] + [0] * (o.T - 1))
    assert "Invalid scalar" in str(ei.value)
    with pytest.raises(FourierError):
        PianoBackendBad()


def PianoBackendBad():
    from fourier_b200 import PianoBackend
    return PianoBackend(2, 3)  # m > n (piano.rs:1062-1064)


def test_setup_file_roundtrip(small, tmp_path):
    from fourier_b200 import PianoBackend
    b, o = small
    rng = random.Random(2)
    coeffs = [rng.randrange(R) for _ in range(o.T)]
    for compressed in (True, False):
        p = str(tmp_path / ("setup_%d" % compressed))
        b.save_setup_to_file(p, compressed)
        b2 = PianoBackend(o.n, o.m)
        b2.load_setup_from_file(p, compressed)
        assert b2.setup_to_bytes(True) == b.setup_to_bytes(True)
        assert b2.worker_commit(3, coeffs) == b.worker_commit(3, coeffs)
        # a row shard only holds its rows
        b3 = PianoBackend(o.n, o.m)
        b3.load_setup_from_file(p, compressed, row0=2, nrows=1)
        assert b3.worker_commit(2, coeffs) == b.worker_commit(2, coeffs)
        from fourier_b200 import FourierError
        with pytest.raises(FourierError):
            b3.worker_commit(0, coeffs)
        b2.close()
        b3.close()
    # an oracle-written file loads too
    blob = opiano.setup_to_bytes(o.settings, True)
    b4 = PianoBackend(o.n, o.m)
    b4.load_setup_from_bytes(blob, True)
    assert b4.worker_commit(1, coeffs) == bls.g1_compress(o.worker_commit(1, coeffs))
    b4.close()


def test_tau_zero_setup(gpu):
    # test_piano_setup uses the all-zero secret (piano.rs:1461-1476): tau = 0
    from fourier_b200 import PianoBackend
    b = PianoBackend(4, 1)
    b.setup(bytes(32), bytes(32))
    o = opiano.PianoOracle(4, 1)
    o.generate_trusted_setup(bytes(32), bytes(32))
    assert b.setup_to_bytes(True) == opiano.setup_to_bytes(o.settings, True)
    rng = random.Random(8)
    coeffs = [rng.randrange(R) for _ in range(o.T)]
    assert b.worker_commit(1, coeffs) == bls.g1_compress(o.worker_commit(1, coeffs))
    y, pi = b.worker_open(1, coeffs, 3)
    oy, opi = o.worker_open(1, coeffs, 3)
    assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)
    b.close()


@pytest.mark.parametrize("n,m,wb", [(12, 0, 0), (12, 0, -1), (14, 4, 0), (16, 0, 0)])
def test_known_tau_commit_open(gpu, n, m, wb):
    """Sizes where the oracle cannot afford the SRS points: C = G*(R_i * sum c_j L_j)."""
    b, o = make(n, m, window_bits=wb)
    rng = random.Random(n * 31 + m)
    for i in sorted({0, o.M - 1, o.M // 2}):
        coeffs = [rng.randrange(R) for _ in range(o.T)]
        alpha = rng.randrange(R)
        assert b.worker_commit(i, coeffs) == bls.g1_compress(o.worker_commit_known_tau(i, coeffs))
        y, pi = b.worker_open(i, coeffs, alpha)
        oy, opi = o.worker_open_known_tau(i, coeffs, alpha)
        assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)
    b.close()


def test_c2_degree_2_20_known_tau(gpu):
    """BASELINE config 2: commit + open of a degree-2^20 polynomial, bit-exact."""
    from fourier_b200 import synth
    b, o = make(20, 0)
    coeffs_b = synth.fr_bytes(31, o.T)
    coeffs = synth.fr_ints(31, o.T)
    alpha = synth.fr_ints(33, 1)[0]
    assert b.worker_commit(0, coeffs_b) == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
    y, pi = b.worker_open(0, coeffs_b, alpha)
    oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
    assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)
    b.close()


def test_device_resident_rows_match_host_api(small):
    """fb_dev_commit_rows / fb_dev_open_rows (operands in HBM, batched iNTT + scan + NTT + MSM
    over all rows) return the bytes of the per-row host calls."""
    import torch
    b, o = small
    rng = random.Random(12)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha = rng.randrange(R)
    flat = fr_list_to_bytes([x for row in rows for x in row])
    dev = torch.device("cuda", 0)
    d_be = torch.frombuffer(bytearray(flat), dtype=torch.uint8).to(dev)
    d_fr = torch.empty_like(d_be)
    b.dev_fr_from_be(d_be, d_fr, o.M * o.T)
    d_a_be = torch.frombuffer(bytearray(alpha.to_bytes(32, "big")), dtype=torch.uint8).to(dev)
    d_alpha = torch.empty_like(d_a_be)
    b.dev_fr_from_be(d_a_be, d_alpha, 1)
    d_c = torch.zeros(o.M * 48, dtype=torch.uint8, device=dev)
    d_y = torch.zeros(o.M * 32, dtype=torch.uint8, device=dev)
    d_p = torch.zeros(o.M * 48, dtype=torch.uint8, device=dev)
    b.dev_commit_rows(0, o.M, d_fr, o.T, d_c)
    b.dev_open_rows(0, o.M, d_fr, o.T, d_alpha, d_y, d_p)
    b.sync()
    cs, ys, ps = bytes(d_c.cpu().numpy()), bytes(d_y.cpu().numpy()), bytes(d_p.cpu().numpy())
    for i in range(o.M):
        assert cs[i * 48:(i + 1) * 48] == b.worker_commit(i, rows[i])
        y, pi = b.worker_open(i, rows[i], alpha)
        assert ys[i * 32:(i + 1) * 32] == y and ps[i * 48:(i + 1) * 48] == pi


# ---------------------------------------------------------------- verify (piano.rs:309-465)
def test_worker_and_master_verify(small):
    """pianist_test (piano.rs:1604-1722): every worker opening verifies, the master
    opening verifies, and any altered input is rejected; answers equal the oracle's."""
    b, o = small
    rng = random.Random(31)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha, beta = rng.randrange(R), rng.randrange(R)
    flat = [x for row in rows for x in row]
    commits = b.commit_rows(0, o.M, flat)
    ys, pis = b.open_rows(0, o.M, flat, alpha)
    ys = [ys[32 * i:32 * i + 32] for i in range(o.M)] if isinstance(ys, (bytes, bytearray)) else list(ys)
    for i in range(o.M):
        assert b.worker_verify(i, commits[i], alpha, ys[i], pis[i])
    i = 2
    y_int = int.from_bytes(ys[i], "big")
    assert not b.worker_verify(i, commits[i], alpha, (y_int + 1) % R, pis[i])
    assert not b.worker_verify(i, commits[i], (alpha + 1) % R, ys[i], pis[i])
    assert not b.worker_verify(i, commits[i - 1], alpha, ys[i], pis[i])
    assert not b.worker_verify(i, commits[i], alpha, ys[i], pis[i - 1])
    assert not b.worker_verify(i - 1, commits[i], alpha, ys[i], pis[i])  # another machine's R_i
    # the oracle agrees on one accepted and one rejected case
    oc, opi = bls.g1_decompress(commits[i]), bls.g1_decompress(pis[i])
    assert o.worker_verify(i, oc, alpha, y_int, opi)
    assert not o.worker_verify(i, oc, alpha, (y_int + 1) % R, opi)

    com = b.master_commit(commits)
    z, (pi0, pi1) = b.master_open(ys, pis, beta)
    z_int = int.from_bytes(z, "big")
    assert b.master_verify(com, beta, alpha, z, (pi0, pi1))
    assert not b.master_verify(com, beta, alpha, (z_int + 1) % R, (pi0, pi1))
    assert not b.master_verify(com, alpha, beta, z, (pi0, pi1))  # alpha and beta swapped
    assert not b.master_verify(com, beta, alpha, z, (pi1, pi0))
    assert not b.master_verify(commits[0], beta, alpha, z, (pi0, pi1))


def test_verify_degenerate_and_errors(small):
    from fourier_b200 import FourierError
    b, o = small
    ident = bls.g1_compress(bls.G1_INF)
    # constant row: zero quotient, proof = identity, still a valid opening
    ones = [1] * o.T
    c = b.worker_commit(2, ones)
    y, pi = b.worker_open(2, ones, 123456789)
    assert pi == ident and b.worker_verify(2, c, 123456789, y, pi)
    assert not b.worker_verify(2, c, 123456789, 2, pi)
    # zero polynomial: commitment, value and proof all trivial
    assert b.worker_verify(0, ident, 5, 0, ident)
    assert b.master_verify(ident, 7, 5, 0, (ident, ident))
    assert not b.master_verify(ident, 7, 5, 1, (ident, ident))
    with pytest.raises(FourierError) as ei:
        b.worker_verify(0, ident, R, 0, ident)
    assert "Invalid scalar" in str(ei.value)
    with pytest.raises(FourierError) as ei:
        b.worker_verify(0, ident, 1, R, ident)
    assert "Invalid scalar" in str(ei.value)
    with pytest.raises(FourierError):
        b.worker_verify(o.M, ident, 1, 1, ident)
    with pytest.raises(FourierError):
        b.worker_verify(0, b"\x00" * 48, 1, 1, ident)  # not a compressed point
    with pytest.raises(FourierError):
        b.master_verify(ident, R + 1, 5, 0, (ident, ident))


@pytest.mark.parametrize("n,m", [(4, 0), (6, 3)])
def test_verify_other_shapes_and_loaded_setup(gpu, n, m, tmp_path):
    """M = 1 (no right NTT) and a setup that came from an uncompressed file
    (test_save_and_load_setup, piano.rs:1789-1833)."""
    from fourier_b200 import PianoBackend
    b0, o = make(n, m)
    p = str(tmp_path / "setup.bin")
    b0.save_setup_to_file(p, False)
    b0.close()
    b = PianoBackend(n, m)
    b.load_setup_from_file(p, False)
    rng = random.Random(n * 10 + m)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    alpha, beta = rng.randrange(R), rng.randrange(R)
    commits = [b.worker_commit(i, rows[i]) for i in range(o.M)]
    opens = [b.worker_open(i, rows[i], alpha) for i in range(o.M)]
    for i in range(o.M):
        assert b.worker_verify(i, commits[i], alpha, opens[i][0], opens[i][1])
    com = b.master_commit(commits)
    z, pi_f = b.master_open([y for y, _ in opens], [q for _, q in opens], beta)
    assert b.master_verify(com, beta, alpha, z, pi_f)
    assert not b.master_verify(com, beta, alpha, (int.from_bytes(z, "big") + 1) % R, pi_f)
    if o.M > 1:
        assert not b.worker_verify(0, commits[1], alpha, opens[1][0], opens[1][1])
    b.close()


# ---------------------------------------------------------------- evaluation-form opening (evalform.cu)
def test_open_alpha_at_every_domain_point(small):
    """The transform-free opening has a separate formula when alpha = w^k: every k,
    single rows and the batched row loop, against the oracle's literal
    iNTT / eval / divide / NTT sequence (piano.rs:193-236)."""
    b, o = small
    rng = random.Random(77)
    w = bls.root_of_unity(o.t)
    rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
    flat = [x for row in rows for x in row]
    for k in range(o.T):
        alpha = pow(w, k, R)
        y, pi = b.worker_open(1, rows[1], alpha)
        oy, opi = o.worker_open(1, rows[1], alpha)
        assert oy == rows[1][k]
        assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi), k
    alpha = pow(w, 5, R)
    ys, pis = b.open_rows(0, o.M, flat, alpha)
    for i in range(o.M):
        oy, opi = o.worker_open(i, rows[i], alpha)
        assert ys[i] == oy.to_bytes(32, "big") and pis[i] == bls.g1_compress(opi)


@pytest.mark.parametrize("length", [1, 2, 4])
def test_open_short_rows(small, length):
    """Fewer evaluations than columns (a smaller domain), alpha inside and outside it."""
    b, o = small
    rng = random.Random(length)
    coeffs = [rng.randrange(R) for _ in range(length)]
    lg = length.bit_length() - 1
    for alpha in (rng.randrange(R), pow(bls.root_of_unity(lg), length - 1, R) if lg else 1):
        y, pi = b.worker_open(0, coeffs, alpha)
        oy, opi = o.worker_open(0, coeffs, alpha)
        assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)


def test_concurrent_calls_on_one_backend(gpu):
    """rpc.rs:165, 414-416: the backend sits in an Arc and every connection's handler calls it from its own runtime
    thread, so concurrent calls on one backend are legal.  Eight host threads commit, open, transform and verify on
    one context at once; every result must equal the one computed serially."""
    import threading
    from fourier_b200 import PianoBackend
    n, m = 12, 2
    b = PianoBackend(n, m)
    b.setup(SECRET_X, SECRET_Y)
    T, M = 1 << (n - m), 1 << m
    rng = random.Random(77)
    jobs = []
    for k in range(8):
        coeffs = fr_list_to_bytes([rng.randrange(R) for _ in range(T)])
        jobs.append((k % M, coeffs, rng.randrange(R)))

    def run(job):
        i, coeffs, alpha = job
        c = b.worker_commit(i, coeffs)
        y, pi = b.worker_open(i, coeffs, alpha)
        f = b.fft(coeffs, inverse=True, left=True)
        ok = b.worker_verify(i, c, alpha, y, pi)
        return c, y, pi, f, ok

    serial = [run(j) for j in jobs]
    assert all(r[4] for r in serial)
    for _ in range(3):
        out = [None] * len(jobs)
        errs = []

        def work(k):
            try:
                out[k] = run(jobs[k])
            except Exception as e:  # noqa: BLE001 - reported below
                errs.append(e)

        th = [threading.Thread(target=work, args=(k,)) for k in range(len(jobs))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errs, errs
        assert out == serial
    b.close()
