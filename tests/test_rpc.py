"""The RPC shim (fourier_b200/rpc.py) against the reference's wire shapes
(src/rpc.rs:18-143, 403-451; tests rpc.rs:620-1000; client fourier/fourier.py).
CPU part: codec, request decoding, response shapes and error behaviour with a
backend built from the oracle (test infrastructure).  GPU part: the reference
client's test_routine (fourier.py:472-533) over HTTP against the real backend."""
import base64
import json
import random
import urllib.request

import pytest

from util import R, SECRET_X, SECRET_Y, bls, opiano

from fourier_b200 import rpc


def enc(b: bytes) -> str:
    return base64.b64encode(b).decode().rstrip("=")


def fr(v: int) -> str:
    return enc(v.to_bytes(32, "big"))


class OracleBackend:
    """PianoBackend-shaped object over the oracle, for the host-logic tests."""

    def __init__(self, n, m):
        self.n, self.m = n, m
        self.o = opiano.PianoOracle(n, m)
        self.o.generate_trusted_setup(SECRET_X, SECRET_Y)

    @staticmethod
    def _ints(raw):
        return [int.from_bytes(raw[i:i + 32], "big") for i in range(0, len(raw), 32)]

    def _check(self, vals):
        if any(v >= R for v in vals):
            raise ValueError("Invalid scalar")
        return vals

    def worker_commit(self, i, coeffs):
        if i >= self.o.M:
            raise ValueError("machine index out of range")
        return bls.g1_compress(self.o.worker_commit(i, self._check(self._ints(coeffs))))

    def worker_open(self, i, coeffs, alpha):
        y, pi = self.o.worker_open(i, self._check(self._ints(coeffs)), int.from_bytes(alpha, "big"))
        return y.to_bytes(32, "big"), bls.g1_compress(pi)

    def worker_verify(self, i, c, alpha, y, pi):
        return self.o.worker_verify(i, bls.g1_decompress(c), int.from_bytes(alpha, "big"), int.from_bytes(y, "big"),
                                    bls.g1_decompress(pi))

    def master_commit(self, cs):
        return bls.g1_compress(self.o.master_commit([bls.g1_decompress(c) for c in cs]))

    def master_open(self, evals, proofs, beta):
        z, (p0, p1) = self.o.master_open(self._ints(evals), [bls.g1_decompress(p) for p in proofs],
                                         int.from_bytes(beta, "big"))
        return z.to_bytes(32, "big"), (bls.g1_compress(p0), bls.g1_compress(p1))

    def master_verify(self, c, beta, alpha, z, pi_f):
        return self.o.master_verify(bls.g1_decompress(c), int.from_bytes(beta, "big"), int.from_bytes(alpha, "big"),
                                    int.from_bytes(z, "big"), tuple(bls.g1_decompress(p) for p in pi_f))

    def evaluate(self, poly, x):
        return opiano.poly_eval(self._ints(poly), int.from_bytes(x, "big")).to_bytes(32, "big")

    def fft(self, data, inverse=False, left=True):
        f = self.o.fft_left if left else self.o.fft_right
        return b"".join(v.to_bytes(32, "big") for v in f(self._ints(data), inverse))


def test_base64_codec_roundtrip_and_strictness():
    rng = random.Random(3)
    vals = [rng.randrange(R) for _ in range(257)] + [0, 1, R - 1]
    raw = b"".join(v.to_bytes(32, "big") for v in vals)
    strs = rpc.encode_fr_list(raw)
    assert strs == [fr(v) for v in vals]  # same text as base64 STANDARD_NO_PAD
    assert rpc.decode_fr_list(strs) == raw
    assert rpc.decode_fr_list([]) == b"" and rpc.encode_fr_list(b"") == []
    with pytest.raises(rpc.RpcError):
        rpc.decode_fr_list([fr(1)[:-1] + "B"])  # non-zero trailing bits
    with pytest.raises(rpc.RpcError):
        rpc.decode_fr_list([fr(1)[:-1] + "*"])
    with pytest.raises(rpc.RpcError) as ei:
        rpc.decode_fr_list([fr(1), enc(b"\x01" * 31)])
    assert "Expected 32, got 31" in str(ei.value)
    with pytest.raises(rpc.RpcError):
        rpc.decode_fr(fr(1) + "=")  # padded input is refused by the NO_PAD engine
    with pytest.raises(rpc.RpcError):
        rpc.decode_g1(fr(1))


@pytest.fixture(scope="module")
def handler():
    return rpc.RpcHandler(OracleBackend(4, 1))


def call(h, method, params=None, **extra):
    req = {"method": method, "params": params, "id": 7, "jsonrpc": "2.0", **extra}
    return json.loads(h.handle_bytes(json.dumps(req).encode()))


def test_shapes_and_protocol_on_oracle_backend(handler):
    h = handler
    assert call(h, "ping") is None  # Pong -> null, params: null accepted, id/jsonrpc ignored
    poly = call(h, "randomPoly")["poly"]
    assert len(poly) == 2 and all(len(row) == 8 and all(len(s) == 43 for s in row) for row in poly)
    alpha, beta = call(h, "randomPoint")["point"], call(h, "randomPoint")["point"]
    commits, evals, proofs = [], [], []
    for i, row in enumerate(poly):
        c = call(h, "workerCommit", {"i": i, "poly": row})
        assert set(c) == {"commitment"}
        o = call(h, "workerOpen", {"i": i, "poly": row, "x": alpha})
        assert set(o) == {"proof", "eval"}
        v = call(h, "workerVerify", {"i": i, "alpha": alpha, "proof": o["proof"], "eval": o["eval"],
                                     "commitment": c["commitment"]})
        assert v == {"valid": True}
        commits.append(c["commitment"]); evals.append(o["eval"]); proofs.append(o["proof"])
    com = call(h, "masterCommit", {"commitments": commits})["commitment"]
    mo = call(h, "masterOpen", {"evals": evals, "proofs": proofs, "beta": beta})
    assert set(mo) == {"z", "pi_0", "pi_1"}
    assert call(h, "masterVerify", {"commitment": com, "beta": beta, "alpha": alpha, **mo}) == {"valid": True}
    assert call(h, "masterVerify", {"commitment": com, "beta": alpha, "alpha": beta, **mo}) == {"valid": False}
    # evaluate answers under "y" (rpc.rs:134-136); fft round trip
    y = call(h, "evaluate", {"poly": [fr(1), fr(2), fr(3)], "x": fr(2)})
    assert y == {"y": fr(17)}
    f = call(h, "fft", {"poly": poly[0], "left": True, "inverse": False})["poly"]
    assert call(h, "fft", {"poly": f, "left": True, "inverse": True})["poly"] == poly[0]


def test_errors_are_message_objects(handler):
    h = handler
    assert "unknown variant" in call(h, "prove", {"poly": []})["message"]  # fourier.py:345-348 calls a missing method
    assert "missing field `poly`" in call(h, "workerCommit", {"i": 0})["message"]
    assert "message" in call(h, "workerCommit", {"i": 0, "poly": [fr(R)] * 8})  # scalar >= r
    assert "message" in call(h, "workerCommit", {"i": 9, "poly": [fr(1)] * 8})
    assert "Expected 48" in call(h, "masterCommit", {"commitments": [fr(1)]})["message"]
    assert "message" in json.loads(h.handle_bytes(b"{not json"))
    assert "message" in json.loads(h.handle_bytes(b"[]"))


def test_http_server_any_method_and_path(handler):
    srv = rpc.make_server(handler, "127.0.0.1", 0)
    import threading
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    try:
        url = "http://127.0.0.1:%d" % srv.server_address[1]
        for method, path in (("POST", "/"), ("PUT", "/anything"), ("GET", "/")):
            req = urllib.request.Request(url + path, data=json.dumps({"method": "ping"}).encode(), method=method)
            with urllib.request.urlopen(req, timeout=10) as resp:
                assert resp.status == 200 and json.loads(resp.read()) is None
        req = urllib.request.Request(url, data=json.dumps({"method": "randomPoint", "params": None}).encode())
        with urllib.request.urlopen(req, timeout=10) as resp:
            assert len(json.loads(resp.read())["point"]) == 43
    finally:
        srv.shutdown()
        srv.server_close()


def test_cli_flags_parse_like_the_reference():
    p = rpc._parser()
    a = p.parse_args(["run", "--host", "127.0.0.1", "--port", "1400", "--scale", "6", "--setup-path", "s",
                      "--precompute-path", "p", "--machines-scale", "2", "--uncompressed"])
    assert (a.host, a.port, a.scale, a.machines_scale, a.setup_path, a.uncompressed) == ("127.0.0.1", 1400, 6, 2, "s", True)
    d = p.parse_args(["run"])
    assert (d.scale, d.machines_scale, d.host, d.port, d.setup_path) == (20, 1, "localhost", 1337, None)  # cli.rs:26-40
    s = p.parse_args(["setup", "--generate-setup", "--generate-precompute", "--overwrite", "--scale", "5"])
    assert s.setup_path == "data/setup" and s.precompute_path == "data/precompute" and s.generate_setup and s.overwrite


def test_setup_flag_combinations_are_refused_like_the_reference(tmp_path, capsys):
    """SetupArgs::can_proceed (cli.rs:89-121), checked ahead of backend creation (no GPU needed to be refused)."""
    existing = tmp_path / "setup"
    existing.write_bytes(b"x")
    base = ["setup", "--setup-path", str(existing), "--precompute-path", str(tmp_path / "pre"), "--scale", "4", "--machines-scale", "1"]
    cases = [
        (["--generate-setup"], "already exists, use --overwrite to overwrite"),
        (["--compress-existing", "--decompress-existing", "--uncompressed"], "Cannot compress and decompress at the same time, choose one"),
        (["--compress-existing"], "Cannot compress an already compressed file"),
        (["--decompress-existing", "--uncompressed"], "Cannot decompress an already decompressed file"),
    ]
    for flags, msg in cases:
        assert rpc.main(base + flags) == 1
        assert msg in capsys.readouterr().err
    pre = tmp_path / "pre"
    pre.write_bytes(b"x")
    assert rpc.main(["setup", "--setup-path", str(tmp_path / "new"), "--precompute-path", str(pre), "--generate-precompute"]) == 1
    assert "already exists" in capsys.readouterr().err


@pytest.mark.gpu
def test_client_routine_over_http_on_gpu(gpu, tmp_path):
    """fourier.py:472-533 test_routine: setup file -> server -> random poly -> every
    worker commits / opens / verifies -> master commits / opens / verifies."""
    n, m = 8, 2
    setup_path, pre_path = str(tmp_path / "setup"), str(tmp_path / "precompute")
    assert rpc.main(["setup", "--setup-path", setup_path, "--precompute-path", pre_path, "--scale", str(n),
                     "--machines-scale", str(m), "--generate-setup", "--generate-precompute"]) == 0
    assert rpc.main(["setup", "--setup-path", setup_path, "--precompute-path", pre_path, "--scale", str(n),
                     "--machines-scale", str(m), "--generate-setup"]) == 1  # exists, no --overwrite
    from fourier_b200 import PianoBackend
    be = PianoBackend(n, m)
    be.load_setup_from_file(setup_path, True)
    srv, _ = rpc.serve_in_thread(be)
    url = "http://127.0.0.1:%d" % srv.server_address[1]

    def post(method, params=None):
        body = json.dumps({"id": 0, "method": method, "params": params, "jsonrpc": "2.0"}).encode()
        with urllib.request.urlopen(urllib.request.Request(url, data=body), timeout=120) as resp:
            return json.loads(resp.read())

    try:
        assert post("ping") is None
        poly = post("randomPoly")["poly"]
        assert len(poly) == 1 << m and len(poly[0]) == 1 << (n - m)
        alpha, beta = post("randomPoint")["point"], post("randomPoint")["point"]
        commits, evals, proofs = [], [], []
        for i, row in enumerate(poly):
            c = post("workerCommit", {"i": i, "poly": row})["commitment"]
            o = post("workerOpen", {"i": i, "poly": row, "x": alpha})
            assert post("workerVerify", {"i": i, "alpha": alpha, "proof": o["proof"], "eval": o["eval"],
                                         "commitment": c}) == {"valid": True}
            commits.append(c); evals.append(o["eval"]); proofs.append(o["proof"])
        com = post("masterCommit", {"commitments": commits})["commitment"]
        mo = post("masterOpen", {"evals": evals, "proofs": proofs, "beta": beta})
        assert post("masterVerify", {"commitment": com, "beta": beta, "alpha": alpha, **mo}) == {"valid": True}
        bad = dict(mo, z=fr((int.from_bytes(base64.b64decode(mo["z"] + "="), "big") + 1) % R))
        assert post("masterVerify", {"commitment": com, "beta": beta, "alpha": alpha, **bad}) == {"valid": False}
        assert "Invalid scalar" in post("workerCommit", {"i": 0, "poly": [fr(R)] * len(poly[0])})["message"]
        # same bytes as the oracle run on the file the server loaded
        o = opiano.PianoOracle(n, m)
        o.settings = opiano.setup_from_bytes(open(setup_path, "rb").read(), True)
        row0 = [int.from_bytes(base64.b64decode(s + "="), "big") for s in poly[0]]
        assert commits[0] == enc(bls.g1_compress(o.worker_commit(0, row0)))
    finally:
        srv.shutdown()
        srv.server_close()
        be.close()
