"""GPU parity at the BASELINE.json configurations and on the paths round 1 left untested:

  C1  n=12, m=0 commit + open + verify through the RPC shim
  C3  Fr NTT 2^24, forward and inverse, against the C oracle
  C4  G1 MSM 2^24, known discrete logs, tabulated and untabulated
  C5  n=22, m=10: 2^10 rows x 2^12 through commit_rows / open_rows, master open + verify
  column shards (fb_ctx_set_col_shard + the *_partial calls), 2 / 3 / 4 shards on one device
  the reference's protocol sweep pianist_test (piano.rs:1604-1722): every n in 2..8, m in 1..n-1
  the literal opening (iNTT -> eval -> divide -> NTT) against the C oracle's open_scalars
  the fused commit+open call, lanes, and setup from several threads / contexts at once
"""
import base64
import json
import random
import threading
import urllib.request

import pytest

from util import R, SECRET_X, SECRET_Y, bls, fr_list_to_bytes, opiano

pytestmark = pytest.mark.gpu


def known_tau_oracle(n, m):
    o = opiano.PianoOracle(n, m)
    o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=False)
    return o


# ------------------------------------------------------------------ column shards
@pytest.mark.parametrize("n,shards", [(12, 2), (12, 3), (12, 4), (16, 2), (16, 4)])
def test_column_shards_sum_to_the_unsharded_result(gpu, n, shards):
    """One context per shard on cuda:0 (the bench runs them on separate GPUs): partial commits and
    proofs added with g1_sum(fmt=96) equal the unsharded call and the known-tau oracle.  3 shards of a
    power-of-two row are unequal (split_range gives the remainder to the low ranks)."""
    from fourier_b200 import PianoBackend, shard
    T = 1 << n
    o = known_tau_oracle(n, 0)
    rng = random.Random(n * 7 + shards)
    coeffs = [rng.randrange(R) for _ in range(T)]
    cb = fr_list_to_bytes(coeffs)
    alpha = rng.randrange(R)
    full = PianoBackend(n, 0)
    full.setup(SECRET_X, SECRET_Y)
    want_c = full.worker_commit(0, cb)
    want_y, want_p = full.worker_open(0, cb, alpha)
    assert want_c == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
    oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
    assert want_y == oy.to_bytes(32, "big") and want_p == bls.g1_compress(opi)
    parts_c, parts_p, parts_fc, parts_fp = [], [], [], []
    for k in range(shards):
        col0, ncols = shard.split_range(T, shards, k)
        b = PianoBackend(n, 0)
        b.set_col_shard(col0, ncols)
        b.setup(SECRET_X, SECRET_Y)
        parts_c.append(b.worker_commit_partial(0, cb))
        y, p = b.worker_open_partial(0, cb, alpha)
        assert y == want_y
        parts_p.append(p)
        fc, fy, fp = b.worker_commit_open_partial(0, cb, alpha)
        assert fy == want_y
        parts_fc.append(fc)
        parts_fp.append(fp)
        from fourier_b200 import FourierError
        with pytest.raises(FourierError):  # a column shard cannot answer the whole-row call
            b.worker_commit(0, cb)
        b.close()
    assert full.g1_sum(b"".join(parts_c), fmt=96) == want_c
    assert full.g1_sum(b"".join(parts_p), fmt=96) == want_p
    assert parts_fc == parts_c and parts_fp == parts_p
    full.close()


@pytest.mark.parametrize("nparts", [1, 2, 5, 16, 17, 40])
def test_device_combine_of_many_parts(gpu, nparts):
    """fb_dev_combine_commit_open on arbitrary 224-byte records (96-byte point | y | 96-byte point): the kernel deserialises
    the parts in parallel (16 per kind at a time) and adds them as a tree.  Parts include the identity, a repeated point
    (P + P inside the tree) and a point with its negative; more parts than one sweep of the warp (17, 40)."""
    import torch
    from fourier_b200 import PianoBackend
    rng = random.Random(900 + nparts)
    be = PianoBackend(4, 0)
    ss = [rng.randrange(1, R) for _ in range(nparts)]
    ts = [rng.randrange(1, R) for _ in range(nparts)]
    if nparts >= 5:
        ss[1] = ss[0]              # P + P
        ss[2] = 0                  # identity
        ts[3] = (R - ts[4]) % R    # Q and -Q
    if nparts >= 17:
        ss[16] = ss[0]             # meets part 0 again, in the second sweep of the same lane
    y = rng.randrange(R).to_bytes(32, "big")
    rec = b"".join(bls.g1_serialize(bls.g1_mul(bls.G1_GEN, s)) + y + bls.g1_serialize(bls.g1_mul(bls.G1_GEN, t))
                   for s, t in zip(ss, ts))
    dev = torch.device("cuda", 0)
    d_parts = torch.frombuffer(bytearray(rec), dtype=torch.uint8).to(dev)
    d_out = torch.zeros(128, dtype=torch.uint8, device=dev)
    be.dev_combine_commit_open(d_parts, nparts, d_out)
    be.sync()
    want = (bls.g1_compress(bls.g1_mul(bls.G1_GEN, sum(ss) % R)) + y + bls.g1_compress(bls.g1_mul(bls.G1_GEN, sum(ts) % R)))
    assert bytes(d_out.cpu().numpy()) == want
    # a part that is not on the curve is rejected
    bad = bytearray(rec)
    bad[95] ^= 1
    d_bad = torch.frombuffer(bad, dtype=torch.uint8).to(dev)
    from fourier_b200 import FourierError
    be.dev_combine_commit_open(d_bad, nparts, d_out)  # asynchronous: the call itself cannot know
    with pytest.raises(FourierError) as ei:
        be.sync()                                     # ... fb_ctx_sync reports the sticky flag, once
    assert "G1 point" in str(ei.value)
    be.sync()
    if nparts == 1:  # the same for a scalar >= r through the device-resident decoder
        d_in = torch.frombuffer(bytearray(R.to_bytes(32, "big") + (5).to_bytes(32, "big")), dtype=torch.uint8).to(dev)
        d_fr = torch.zeros(64, dtype=torch.uint8, device=dev)
        be.dev_fr_from_be(d_in, d_fr, 2)
        with pytest.raises(FourierError) as ei:
            be.sync()
        assert "Invalid scalar" in str(ei.value)
        be.sync()
    be.close()


@pytest.mark.parametrize("n,length,shards", [(12, 1 << 12, 2), (12, 1 << 12, 3), (14, 1 << 14, 4), (12, 1 << 10, 4), (12, 2, 4), (18, 1 << 18, 8)])
def test_sharded_field_half_matches_the_unsharded_calls(gpu, n, length, shards):
    """fb_worker_commit_open_shard_begin / _finish: every shard uploads and works on its slice only; the 128-byte states are
    exchanged; the partials summed (host g1_sum and the device combine) equal worker_commit / worker_open of the whole row
    and the known-tau oracle.  alpha outside the domain, on the domain (the index lands in exactly one shard's slice), and
    rows shorter than the resident width (some shards hold no element of the row)."""
    import torch
    from fourier_b200 import PianoBackend, shard
    T = 1 << n
    rng = random.Random(n * 131 + shards + length)
    coeffs = [rng.randrange(R) for _ in range(length)]
    cb = fr_list_to_bytes(coeffs)
    full = PianoBackend(n, 0)
    full.setup(SECRET_X, SECRET_Y)
    bes = []
    for k in range(shards):
        col0, ncols = shard.split_range(T, shards, k)
        b = PianoBackend(n, 0)
        b.set_col_shard(col0, ncols)
        b.setup(SECRET_X, SECRET_Y)
        bes.append((b, col0, ncols))
    w = bls.root_of_unity(length.bit_length() - 1)
    alphas = [rng.randrange(R), pow(w, (length * 5) // 7, R), 1]
    if n <= 14:
        o = known_tau_oracle(n, 0)
    dev = torch.device("cuda", 0)
    for alpha in alphas:
        want_c = full.worker_commit(0, cb)
        want_y, want_p = full.worker_open(0, cb, alpha)
        if n <= 14 and length == T:
            assert want_c == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
            oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
            assert (want_y, want_p) == (oy.to_bytes(32, "big"), bls.g1_compress(opi))
        states = []
        for b, col0, ncols in bes:
            lo, hi = min(col0, length), min(col0 + ncols, length)
            states.append(b.commit_open_shard_begin(0, cb[32 * lo:32 * hi], length, alpha))
        gathered = b"".join(reversed(states))  # any order
        parts = []
        for b, col0, ncols in bes:
            c96, y, p96 = b.commit_open_shard_finish(0, gathered, alpha)
            assert y == want_y
            parts.append((c96, y, p96))
        assert full.g1_sum(b"".join(c for c, _, _ in parts), fmt=96) == want_c
        assert full.g1_sum(b"".join(p for _, _, p in parts), fmt=96) == want_p
        # the device combine of the gathered 224-byte records
        d_parts = torch.frombuffer(bytearray(b"".join(c + y + p for c, y, p in parts)), dtype=torch.uint8).to(dev)
        d_out = torch.zeros(128, dtype=torch.uint8, device=dev)
        full.dev_combine_commit_open(d_parts, shards, d_out)
        full.sync()
        assert bytes(d_out.cpu().numpy()) == want_c + want_y + want_p
    from fourier_b200 import FourierError
    with pytest.raises(FourierError):  # finish without begin
        bes[0][0].commit_open_shard_finish(0, bytes(128 * shards), 5)
    with pytest.raises(FourierError):  # a slice of the wrong length
        bes[0][0].commit_open_shard_begin(0, cb[:32], length * 2 if length * 2 <= T else length // 2, 5)
    if length >= shards:
        bad = bytearray(cb)
        bad[0:32] = R.to_bytes(32, "big")  # shard 0 owns coefficient 0
        with pytest.raises(FourierError) as ei:
            lo, hi = 0, min(bes[0][2], length)
            bes[0][0].commit_open_shard_begin(0, bytes(bad[32 * lo:32 * hi]), length, 5)
        assert "Invalid scalar" in str(ei.value)
    for b, _, _ in bes:
        b.close()
    full.close()


def test_column_shard_rejects_bad_scalar_in_its_slice(gpu):
    """fb_worker_commit_partial uploads only its slice, so the shard that owns an out-of-range scalar
    is the one that says "Invalid scalar" (some shard always does); the open uploads the full row."""
    from fourier_b200 import FourierError, PianoBackend, shard
    n, shards = 10, 4
    T = 1 << n
    coeffs = [5] * T
    bad_at = 3 * T // 4 + 7  # in the last shard's slice
    coeffs[bad_at] = R
    cb = fr_list_to_bytes(coeffs)
    rejected = []
    for k in range(shards):
        col0, ncols = shard.split_range(T, shards, k)
        b = PianoBackend(n, 0)
        b.set_col_shard(col0, ncols)
        b.setup(SECRET_X, SECRET_Y)
        try:
            b.worker_commit_partial(0, cb)
            rejected.append(False)
        except FourierError as e:
            assert "Invalid scalar" in str(e)
            rejected.append(True)
        with pytest.raises(FourierError):
            b.worker_open_partial(0, cb, 3)
        b.close()
    assert rejected == [False, False, False, True]


# ------------------------------------------------------------------ pianist_test sweep
@pytest.mark.parametrize("n", list(range(2, 9)))
def test_pianist_sweep(gpu, n):
    """piano.rs:1604-1722 runs the whole protocol for every n in 2..8 and m in 1..n-1."""
    from fourier_b200 import PianoBackend
    for m in range(1, n):
        o = opiano.PianoOracle(n, m)
        o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=(n <= 5))
        b = PianoBackend(n, m)
        b.setup(SECRET_X, SECRET_Y)
        rng = random.Random(100 * n + m)
        rows = [[rng.randrange(R) for _ in range(o.T)] for _ in range(o.M)]
        alpha, beta = rng.randrange(R), rng.randrange(R)
        commits, ys, pis = [], [], []
        for i in range(o.M):
            c = b.worker_commit(i, rows[i])
            y, pi = b.worker_open(i, rows[i], alpha)
            if n <= 5:
                assert c == bls.g1_compress(o.worker_commit(i, rows[i]))
                oy, opi = o.worker_open(i, rows[i], alpha)
            else:
                assert c == bls.g1_compress(o.worker_commit_known_tau(i, rows[i]))
                oy, opi = o.worker_open_known_tau(i, rows[i], alpha)
            assert y == oy.to_bytes(32, "big") and pi == bls.g1_compress(opi)
            assert (c, y, pi) == b.worker_commit_open(i, rows[i], alpha)
            assert b.worker_verify(i, c, alpha, y, pi)
            commits.append(c); ys.append(y); pis.append(pi)
        com = b.master_commit(commits)
        z, (pi0, pi1) = b.master_open(ys, pis, beta)
        # z = f(beta, alpha) (piano.rs:1478-1563) through the oracle's own master_open_scalars
        oz, _ = o.master_open_scalars([int.from_bytes(y, "big") for y in ys], beta)
        assert z == oz.to_bytes(32, "big")
        assert pi0 == bls.g1_compress(o.master_commit([bls.g1_decompress(p) for p in pis]))
        assert b.master_verify(com, beta, alpha, z, (pi0, pi1))
        assert not b.master_verify(com, beta, alpha, (oz + 1) % R, (pi0, pi1))
        b.close()


# ------------------------------------------------------------------ literal opening
@pytest.mark.parametrize("log_n", [12, 16, 20])
def test_literal_open_path_vs_c_oracle(gpu, log_n):
    """fb_ctx_set_open_mode(1): iNTT -> Horner/quotient scan -> NTT (piano.rs:193-236) at worker sizes.
    y equals the C oracle's, and the proof equals the MSM of the oracle's NTT(q || 0) over the same
    row (worker_commit IS that MSM), and both equal the evaluation-form default."""
    from fourier_b200 import PianoBackend, synth
    from oracle import cref
    T = 1 << log_n
    b = PianoBackend(log_n, 0)
    b.setup(SECRET_X, SECRET_Y)
    coeffs = synth.fr_bytes(60 + log_n, T)
    alpha = synth.fr_bytes(61, 1).tobytes()
    oy, oq = cref.open_scalars(coeffs, alpha)
    want_proof = b.worker_commit(0, oq)
    y0, p0 = b.worker_open(0, coeffs, alpha)  # evaluation form
    b.set_open_mode(True)
    y1, p1 = b.worker_open(0, coeffs, alpha)  # literal
    c2, y2, p2 = b.worker_commit_open(0, coeffs, alpha)
    b.set_open_mode(False)
    assert y0 == y1 == y2 == oy
    assert p0 == p1 == p2 == want_proof
    assert c2 == b.worker_commit(0, coeffs)
    if log_n <= 16:
        # alpha on the domain, literal path
        w = bls.root_of_unity(log_n)
        a = pow(w, 12345 % T, R).to_bytes(32, "big")
        oy, oq = cref.open_scalars(coeffs, a)
        b.set_open_mode(True)
        y1, p1 = b.worker_open(0, coeffs, a)
        b.set_open_mode(False)
        y0, p0 = b.worker_open(0, coeffs, a)
        assert y0 == y1 == oy == bytes(coeffs[12345 % T])
        assert p0 == p1 == b.worker_commit(0, oq)
    b.close()


# ------------------------------------------------------------------ C1: 2^12 through the RPC incl. verify
def test_c1_degree_2_12_through_rpc(gpu):
    from fourier_b200 import PianoBackend, rpc, synth
    n = 12
    be = PianoBackend(n, 0)
    be.setup(SECRET_X, SECRET_Y)
    o = known_tau_oracle(n, 0)
    srv, _ = rpc.serve_in_thread(be)
    url = "http://127.0.0.1:%d" % srv.server_address[1]

    def enc(bv: bytes) -> str:
        return base64.b64encode(bv).decode().rstrip("=")

    def dec(s: str) -> bytes:
        return base64.b64decode(s + "=" * (-len(s) % 4))

    def post(method, params=None):
        body = json.dumps({"id": 0, "method": method, "params": params, "jsonrpc": "2.0"}).encode()
        with urllib.request.urlopen(urllib.request.Request(url, data=body), timeout=120) as resp:
            return json.loads(resp.read())

    try:
        coeffs = synth.fr_ints(31, 1 << n)
        alpha = synth.fr_ints(33, 1)[0]
        row = [enc(v.to_bytes(32, "big")) for v in coeffs]
        a64 = enc(alpha.to_bytes(32, "big"))
        c = post("workerCommit", {"i": 0, "poly": row})["commitment"]
        op = post("workerOpen", {"i": 0, "poly": row, "x": a64})
        assert dec(c) == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
        oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
        assert dec(op["eval"]) == oy.to_bytes(32, "big") and dec(op["proof"]) == bls.g1_compress(opi)
        ok = post("workerVerify", {"i": 0, "alpha": a64, "proof": op["proof"], "eval": op["eval"], "commitment": c})
        assert ok == {"valid": True}
        bad = enc(((oy + 1) % R).to_bytes(32, "big"))
        assert post("workerVerify", {"i": 0, "alpha": a64, "proof": op["proof"], "eval": bad, "commitment": c}) == {"valid": False}
        # M = 1: the master step is the identity on one row, and must verify too
        com = post("masterCommit", {"commitments": [c]})["commitment"]
        assert com == c
        beta = enc((7).to_bytes(32, "big"))
        mo = post("masterOpen", {"evals": [op["eval"]], "proofs": [op["proof"]], "beta": beta})
        assert post("masterVerify", {"commitment": com, "beta": beta, "alpha": a64, **mo}) == {"valid": True}
    finally:
        srv.shutdown()
        srv.server_close()
        be.close()


# ------------------------------------------------------------------ C3: NTT 2^24
@pytest.mark.timeout(900)
def test_c3_ntt_2_24_vs_c_oracle(gpu):
    from fourier_b200 import PianoBackend, synth
    from oracle import cref
    n = 1 << 24
    b = PianoBackend(24, 0)
    data = synth.fr_bytes(64, n)
    fwd = b.ntt(data, inverse=False)
    assert fwd == cref.ntt(data, False)
    inv = b.ntt(data, inverse=True)
    assert inv == cref.ntt(data, True)
    assert b.ntt(fwd, inverse=True) == data.tobytes()
    b.close()


# ------------------------------------------------------------------ C4: MSM 2^24
def _dlog_sum(seed_s, seed_k, n):
    """sum_i s_i * k_i mod r over the synthetic vectors, in chunks (python big ints)"""
    from fourier_b200 import synth
    acc = 0
    step = 1 << 18
    for lo in range(0, n, step):
        sw = synth.fr_words(seed_s, min(step, n - lo), lo)
        kw = synth.fr_words(seed_k, min(step, n - lo), lo)
        sb = sw.astype("<u8").tobytes()
        kb = kw.astype("<u8").tobytes()
        acc += sum(int.from_bytes(sb[32 * i:32 * i + 32], "little") * int.from_bytes(kb[32 * i:32 * i + 32], "little")
                   for i in range(min(step, n - lo)))
        acc %= R
    return acc


@pytest.mark.timeout(1200)
def test_c4_msm_2_24_known_dlog(gpu):
    """P_k = s_k G, so sum k_k P_k = (sum k_k s_k) G: one scalar multiplication in the oracle.  Tabulated
    (the bench configuration: 2^24 x 12 windows approaches the 31-bit entry index) and untabulated."""
    from fourier_b200 import PianoBackend, synth
    n = 1 << 24
    b = PianoBackend(10, 0)
    bases = b.bases_from_scalars(synth.fr_bytes(21, n))
    ks = synth.fr_bytes(22, n)
    want = bls.g1_compress(bls.g1_mul(bls.G1_GEN, _dlog_sum(21, 22, n)))
    assert b.msm(bases, ks) == want  # untabulated: per-window bucket sets + doubling ladder
    bases.precompute()
    assert b.msm(bases, ks) == want
    # shorter scalar vector than the base set, and the partial (uncompressed) form
    m = n - 12345
    want_m = bls.g1_compress(bls.g1_mul(bls.G1_GEN, _dlog_sum(21, 22, m)))
    part = b.msm_partial(bases, ks[:m])
    assert b.g1_sum(part, fmt=96) == want_m
    bases.free()
    b.close()


# ------------------------------------------------------------------ C5: 2^10 rows x 2^12
@pytest.mark.timeout(1200)
def test_c5_bivariate_rows_full_size(gpu):
    from fourier_b200 import PianoBackend, synth
    n, m = 22, 10
    M, T = 1 << m, 1 << (n - m)
    b = PianoBackend(n, m)
    b.setup(SECRET_X, SECRET_Y)
    o = known_tau_oracle(n, m)
    coeffs = synth.fr_bytes(51, M * T)
    alpha_b = synth.fr_bytes(53, 1).tobytes()
    beta_b = synth.fr_bytes(54, 1).tobytes()
    alpha, beta = int.from_bytes(alpha_b, "big"), int.from_bytes(beta_b, "big")
    commits = b.commit_rows(0, M, coeffs)
    ys, pis = b.open_rows(0, M, coeffs, alpha_b)
    assert b.commit_open_rows(0, M, coeffs, alpha_b) == (commits, ys, pis)
    sample = sorted({0, 1, 2, 3, 17, 255, 256, 511, 512, 513, 700, 767, 1000, 1021, 1022, 1023})
    for i in sample:
        row = synth.fr_ints(51, T, i * T)
        assert commits[i] == bls.g1_compress(o.worker_commit_known_tau(i, row)), i
        oy, opi = o.worker_open_known_tau(i, row, alpha)
        assert ys[i] == oy.to_bytes(32, "big") and pis[i] == bls.g1_compress(opi), i
    # per-row calls give the rows of the batched call
    for i in (0, 513, 1023):
        row_b = coeffs[i * T:(i + 1) * T]
        assert b.worker_commit(i, row_b) == commits[i]
        assert b.worker_open(i, row_b, alpha_b) == (ys[i], pis[i])
        assert b.worker_verify(i, commits[i], alpha_b, ys[i], pis[i])
    com = b.master_commit(commits)
    z, (pi0, pi1) = b.master_open(ys, pis, beta_b)
    oz, _ = o.master_open_scalars([int.from_bytes(y, "big") for y in ys], beta)
    assert z == oz.to_bytes(32, "big")
    assert b.master_verify(com, beta_b, alpha_b, z, (pi0, pi1))
    assert not b.master_verify(com, beta_b, alpha_b, (oz + 1) % R, (pi0, pi1))
    assert not b.master_verify(commits[0], beta_b, alpha_b, z, (pi0, pi1))
    # the master commitment is G * f(tau_Y, tau_X) = G * sum_i R_i(tau_Y) sum_j c_ij L_j(tau_X)
    s = o.settings
    words = synth.fr_words(51, M * T).astype("<u8")
    acc = 0
    for i in range(M):
        rb = words[i * T:(i + 1) * T].tobytes()
        inner = sum(int.from_bytes(rb[32 * j:32 * j + 32], "little") * s.l_at_tau[j] for j in range(T)) % R
        acc = (acc + inner * s.r_at_tau[i]) % R
    assert com == bls.g1_compress(bls.g1_mul(bls.G1_GEN, acc))
    b.close()


# ------------------------------------------------------------------ lanes, fused call, concurrent setup
def test_lanes_overlap_device_calls_same_bytes(gpu):
    """fb_ctx_set_lanes(2): dev commit and dev open are enqueued on two internal streams forked from the
    caller's stream; after join the caller's stream sees both results.  Same bytes as the host calls,
    as the fused call, and as lanes = 1."""
    import torch
    from fourier_b200 import PianoBackend, synth
    n = 14
    T = 1 << n
    b = PianoBackend(n, 0)
    b.setup(SECRET_X, SECRET_Y)
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    b.set_stream(st.cuda_stream)
    coeffs = synth.fr_bytes(71, T)
    alpha_b = synth.fr_bytes(72, 1).tobytes()
    want_c = b.worker_commit(0, coeffs)
    want_y, want_p = b.worker_open(0, coeffs, alpha_b)
    with torch.cuda.stream(st):
        d_be = torch.from_numpy(coeffs.reshape(-1)).to(dev)
        d_a_be = torch.frombuffer(bytearray(alpha_b), dtype=torch.uint8).to(dev)
        d_fr = torch.empty_like(d_be)
        d_alpha = torch.empty_like(d_a_be)
        for lanes in (1, 2, 3):
            b.set_lanes(lanes)
            d_c = torch.zeros(48, dtype=torch.uint8, device=dev)
            d_o = torch.zeros(80, dtype=torch.uint8, device=dev)
            d_f = torch.zeros(128, dtype=torch.uint8, device=dev)
            for _ in range(3):
                b.dev_fr_from_be(d_be, d_fr, T)       # utilities run on the caller's stream
                b.dev_fr_from_be(d_a_be, d_alpha, 1)
                b.dev_worker_commit(0, d_fr, T, d_c)   # lane k
                b.dev_worker_open(0, d_fr, T, d_alpha, d_o)  # lane k+1
                b.dev_worker_commit_open(0, d_fr, T, d_alpha, d_f)
                b.join()
                got = (bytes(d_c.cpu().numpy()), bytes(d_o.cpu().numpy()), bytes(d_f.cpu().numpy()))
                assert got[0] == want_c and got[1] == want_y + want_p
                assert got[2] == want_c + want_y + want_p
    b.set_lanes(1)
    assert b.worker_commit_open(0, coeffs, alpha_b) == (want_c, want_y, want_p)
    b.close()


@pytest.mark.parametrize("n,m", [(19, 0), (14, 8)])
def test_split_msm_same_bytes(gpu, n, m):
    """fb_ctx_set_msm_split: one MSM as 2..4 sub-MSMs on sub-streams - over point ranges for a single row (and the
    fused pair), over row ranges for a batch of rows - gives the bytes of the unsplit run."""
    from fourier_b200 import PianoBackend, synth
    M, T = 1 << m, 1 << (n - m)
    b = PianoBackend(n, m)
    b.setup(SECRET_X, SECRET_Y)
    coeffs = synth.fr_bytes(91, M * T)
    alpha_b = synth.fr_bytes(92, 1).tobytes()
    want = (b.commit_rows(0, M, coeffs), b.open_rows(0, M, coeffs, alpha_b))
    for mode in (0, 1):  # the fused rows call: commit MSMs on a sub-stream / everything on one stream
        b.set_pair_mode(mode)
        cs, ys, ps = b.commit_open_rows(0, M, coeffs, alpha_b)
        assert (cs, (ys, ps)) == want
    b.set_pair_mode(0)
    want_fused = b.worker_commit_open(M - 1, coeffs[(M - 1) * T:], alpha_b)
    assert want_fused == (want[0][M - 1], want[1][0][M - 1], want[1][1][M - 1])
    for parts in (2, 3, 4):
        b.set_msm_split(parts)
        for _ in range(2):
            assert (b.commit_rows(0, M, coeffs), b.open_rows(0, M, coeffs, alpha_b)) == want
            assert b.worker_commit_open(M - 1, coeffs[(M - 1) * T:], alpha_b) == want_fused
    b.set_msm_split(1)
    for mode in (1, 2, 0):  # the fused pair as one batch of two, as two MSMs on two streams, by size
        b.set_pair_mode(mode)
        for _ in range(2):
            assert b.worker_commit_open(M - 1, coeffs[(M - 1) * T:], alpha_b) == want_fused
    if m == 0:
        o = known_tau_oracle(n, 0)
        assert want[0][0] == bls.g1_compress(o.worker_commit_known_tau(0, synth.fr_ints(91, T)))
    b.close()


def test_pageable_and_pinned_host_inputs_agree(gpu):
    """A Rust Vec<u8> / Python bytes is pageable: it goes through the lane's pinned double buffer (several 4 MiB
    chunks here); a pinned tensor is copied directly; forcing the driver's own pageable path gives the same bytes."""
    import torch
    from fourier_b200 import PianoBackend, synth
    n = 19  # 16 MiB of coefficients
    b = PianoBackend(n, 0)
    b.setup(SECRET_X, SECRET_Y)
    arr = synth.fr_bytes(81, 1 << n)
    as_bytes = arr.tobytes()
    pinned = torch.from_numpy(arr.reshape(-1).copy()).pin_memory()
    r_ring = b.worker_commit(0, as_bytes)
    r_pin = b.worker_commit(0, pinned)
    b.set_staging(True)
    r_drv = b.worker_commit(0, as_bytes)
    b.set_staging(False)
    assert r_ring == r_pin == r_drv
    o = known_tau_oracle(n, 0)
    assert r_ring == bls.g1_compress(o.worker_commit_known_tau(0, synth.fr_ints(81, 1 << n)))
    b.close()


def test_concurrent_setup_on_several_contexts(gpu):
    """Two contexts running setup() at the same time from two threads (same device; and two devices when the
    box has them) produce the SRS of a serial run: the generator comb table is per device and published
    only once built."""
    import torch
    from fourier_b200 import PianoBackend
    n, m = 8, 2
    ref = PianoBackend(n, m)
    ref.setup(SECRET_X, SECRET_Y)
    want = ref.setup_to_bytes(True)
    ref.close()
    devices = [0, 0, 0, 0]
    if torch.cuda.device_count() > 1:
        devices += [1, 1]
    for _ in range(2):
        out, errs = [None] * len(devices), []

        def work(k):
            try:
                b = PianoBackend(n, m, device=devices[k])
                b.setup(SECRET_X, SECRET_Y)
                out[k] = b.setup_to_bytes(True)
                b.close()
            except Exception as e:  # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=work, args=(k,)) for k in range(len(devices))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errs, errs
        assert all(o == want for o in out)


def test_last_error_is_per_thread(gpu):
    """fb_last_error is thread-local: a failing call on one thread cannot change what another thread reads."""
    from fourier_b200 import FourierError, PianoBackend
    b = PianoBackend(6, 1)
    b.setup(SECRET_X, SECRET_Y)
    msgs = {}

    def bad_scalar():
        for _ in range(20):
            try:
                b.worker_commit(0, [R] + [0] * 31)
            except FourierError as e:
                msgs.setdefault("scalar", set()).add(str(e))

    def bad_index():
        for _ in range(20):
            try:
                b.worker_commit(5, [1] * 32)
            except FourierError as e:
                msgs.setdefault("index", set()).add(str(e))

    th = [threading.Thread(target=bad_scalar), threading.Thread(target=bad_index)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert msgs["scalar"] == {"Invalid scalar"}
    assert msgs["index"] == {"machine index out of range"}
    b.close()


def test_concurrent_calls_overlap_on_the_gpu(gpu):
    """rpc.rs:414-416, 502: one tokio task per connection on an immutable backend, so calls overlap.  Here every
    host call takes its own lane (stream + scratch): eight threads must finish the same jobs clearly faster than
    one thread does, with identical results."""
    import time
    from fourier_b200 import PianoBackend
    n, m = 14, 2
    b = PianoBackend(n, m)
    b.setup(SECRET_X, SECRET_Y)
    T, M = 1 << (n - m), 1 << m
    rng = random.Random(5)
    jobs = [(k % M, fr_list_to_bytes([rng.randrange(R) for _ in range(T)]), rng.randrange(R)) for k in range(16)]

    def run(job):
        i, coeffs, alpha = job
        return b.worker_commit(i, coeffs), b.worker_open(i, coeffs, alpha)

    serial = [run(j) for j in jobs]  # warm-up (scratch of lane 0)
    out = [None] * len(jobs)

    def work(ids):
        for k in ids:
            out[k] = run(jobs[k])

    def timed(nthreads):
        th = [threading.Thread(target=work, args=(range(t, len(jobs), nthreads),)) for t in range(nthreads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        return time.perf_counter() - t0

    timed(8)  # warm-up of the other lanes' scratch
    assert out == serial
    t1 = min(timed(1) for _ in range(3))
    t8 = min(timed(8) for _ in range(3))
    assert out == serial
    print("serial %.2f ms, 8 threads %.2f ms, speed-up %.2fx" % (t1 * 1e3, t8 * 1e3, t1 / t8))
    assert t1 / t8 > 1.5
    b.close()
