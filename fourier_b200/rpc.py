"""The reference's RPC surface over the GPU backend.

Serves the request / response shapes of the reference server (src/rpc.rs:18-143,
403-451) so that its Python client (fourier/fourier.py) talks to this process
unchanged:

  * request  {"method": "<camelCase>", "params": {...} | null}; extra keys such as
    "id" / "jsonrpc" are ignored (serde's default), any HTTP method and path;
  * response the flat result object (`#[serde(untagged)]`, rpc.rs:98-143), `null`
    for ping (unit variant `Pong`), {"message": "..."} for every error, always
    HTTP 200 (rpc.rs:427-438);
  * scalars / points are base64 STANDARD_NO_PAD (src/utils.rs) of FsFr::to_bytes
    (32 B big-endian) and FsG1::to_bytes (48 B compressed).

Command line (src/cli.rs:17-88), e.g. through `bin/fourier`:
    fourier run   [--host H] [--port P] [--scale N] [--machines-scale M]
                  [--setup-path F] [--precompute-path F] [--uncompressed]
    fourier setup [--setup-path F] [--precompute-path F] [--scale N] [--machines-scale M]
                  [--overwrite] [--generate-setup] [--generate-precompute] [--uncompressed]
                  [--compress-existing] [--decompress-existing]
The precompute file of the reference is an opaque dump of its BGMW tables
(piano.rs:506-585); here the window tables are rebuilt in HBM whenever a setup is
loaded, so the path is accepted and, on `setup --generate-precompute`, an empty
placeholder is written so that existence checks keep working.

With 2^20 coefficients a request is ~46 MiB of JSON, so the scalar lists go through
a vectorised base64 codec (numpy) instead of one b64decode per element.
"""
from __future__ import annotations

import argparse
import base64
import json
import os
import secrets
import sys
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict, List, Optional

import numpy as np

R_MODULUS = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
FR_B64 = 43  # ceil(32 * 4 / 3) characters without padding
G1_B64 = 64

_ALPHABET = b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/"
_DEC = np.full(256, 255, dtype=np.uint8)
_DEC[np.frombuffer(_ALPHABET, dtype=np.uint8)] = np.arange(64, dtype=np.uint8)
_ENC = np.frombuffer(_ALPHABET, dtype=np.uint8)


class RpcError(Exception):
    """Err(String) of a handler: becomes {"message": ...} (rpc.rs:427-432)."""


# ------------------------------------------------------------------ base64, STANDARD_NO_PAD
def b64_encode(raw: bytes) -> str:
    return base64.b64encode(raw).decode("ascii").rstrip("=")


def b64_decode(s: str) -> bytes:
    if not isinstance(s, str):
        raise RpcError("invalid type: expected a string")
    if "=" in s:
        raise RpcError("Invalid padding")  # the NO_PAD engine refuses padded input
    try:
        return base64.b64decode(s + "=" * (-len(s) % 4), validate=True)
    except Exception as e:  # binascii.Error
        raise RpcError("Invalid byte or length in base64 input: %s" % e)


def decode_fr_list(items: List[str]) -> bytes:
    """List of base64 scalars -> n x 32 bytes (the C ABI's input)."""
    n = len(items)
    if n == 0:
        return b""
    if not all(isinstance(s, str) and len(s) == FR_B64 for s in items):
        out = []
        for s in items:  # slow path: report the first offending element as the reference would
            raw = b64_decode(s)
            if len(raw) != 32:
                raise RpcError("Invalid byte length. Expected 32, got %d" % len(raw))
            out.append(raw)
        return b"".join(out)
    try:
        chars = np.frombuffer("".join(items).encode("ascii"), dtype=np.uint8).reshape(n, FR_B64)
    except (UnicodeEncodeError, ValueError):
        raise RpcError("Invalid byte in base64 input")
    v = _DEC[chars]
    if (v == 255).any():
        raise RpcError("Invalid byte in base64 input")
    if (v[:, FR_B64 - 1] & 3).any():
        raise RpcError("Invalid last symbol in base64 input")  # non-zero trailing bits
    q = np.zeros((n, 44), dtype=np.uint8)
    q[:, :FR_B64] = v
    q = q.reshape(n, 11, 4)
    out = np.empty((n, 11, 3), dtype=np.uint8)
    out[:, :, 0] = (q[:, :, 0] << 2) | (q[:, :, 1] >> 4)
    out[:, :, 1] = (q[:, :, 1] << 4) | (q[:, :, 2] >> 2)
    out[:, :, 2] = (q[:, :, 2] << 6) | q[:, :, 3]
    return out.reshape(n, 33)[:, :32].tobytes()


def encode_fr_list(raw: bytes) -> List[str]:
    """n x 32 bytes -> list of base64 scalars."""
    n = len(raw) // 32
    if n == 0:
        return []
    b = np.zeros((n, 33), dtype=np.uint8)
    b[:, :32] = np.frombuffer(raw, dtype=np.uint8).reshape(n, 32)
    b = b.reshape(n, 11, 3)
    q = np.empty((n, 11, 4), dtype=np.uint8)
    q[:, :, 0] = b[:, :, 0] >> 2
    q[:, :, 1] = ((b[:, :, 0] & 3) << 4) | (b[:, :, 1] >> 4)
    q[:, :, 2] = ((b[:, :, 1] & 15) << 2) | (b[:, :, 2] >> 6)
    q[:, :, 3] = b[:, :, 2] & 63
    text = _ENC[q.reshape(n, 44)[:, :FR_B64]].tobytes().decode("ascii")
    return [text[i * FR_B64:(i + 1) * FR_B64] for i in range(n)]


def decode_fr(s: str) -> bytes:
    raw = b64_decode(s)
    if len(raw) != 32:
        raise RpcError("Invalid byte length. Expected 32, got %d" % len(raw))
    return raw


def decode_g1(s: str) -> bytes:
    raw = b64_decode(s)
    if len(raw) != 48:
        raise RpcError("Invalid byte length. Expected 48, got %d" % len(raw))
    return raw


def random_fr() -> bytes:
    """FsFr::rand(): uniform in [0, r)."""
    return (int.from_bytes(secrets.token_bytes(48), "big") % R_MODULUS).to_bytes(32, "big")


# ------------------------------------------------------------------ handlers (rpc.rs:181-400)
def _need(params, *names):
    if not isinstance(params, dict):
        raise RpcError("invalid type: expected struct variant with params")
    out = []
    for k in names:
        if k not in params:
            raise RpcError("missing field `%s`" % k)
        out.append(params[k])
    return out


def _index(i) -> int:
    if isinstance(i, bool) or not isinstance(i, int) or i < 0:
        raise RpcError("invalid type: expected usize for `i`")
    return i


class RpcHandler:
    """Maps a decoded request onto a backend with the PianoBackend method names
    (fourier_b200.engine.PianoBackend, or any object of the same shape)."""

    def __init__(self, backend):
        self.backend = backend
        self._methods: Dict[str, Callable] = {
            "workerCommit": self.worker_commit, "workerOpen": self.worker_open, "workerVerify": self.worker_verify,
            "masterCommit": self.master_commit, "masterOpen": self.master_open, "masterVerify": self.master_verify,
            "randomPoly": self.random_poly, "randomPoint": self.random_point, "evaluate": self.evaluate,
            "fft": self.fft, "ping": self.ping,
        }

    # worker
    def worker_commit(self, p):
        i, poly = _need(p, "i", "poly")
        c = self.backend.worker_commit(_index(i), decode_fr_list(poly))
        return {"commitment": b64_encode(c)}

    def worker_open(self, p):
        i, poly, x = _need(p, "i", "poly", "x")
        y, proof = self.backend.worker_open(_index(i), decode_fr_list(poly), decode_fr(x))
        return {"proof": b64_encode(proof), "eval": b64_encode(y)}

    def worker_verify(self, p):
        i, alpha, proof, ev, com = _need(p, "i", "alpha", "proof", "eval", "commitment")
        ok = self.backend.worker_verify(_index(i), decode_g1(com), decode_fr(alpha), decode_fr(ev), decode_g1(proof))
        return {"valid": bool(ok)}

    # master
    def master_commit(self, p):
        (coms,) = _need(p, "commitments")
        return {"commitment": b64_encode(self.backend.master_commit([decode_g1(c) for c in coms]))}

    def master_open(self, p):
        evals, proofs, beta = _need(p, "evals", "proofs", "beta")
        z, (pi0, pi1) = self.backend.master_open(decode_fr_list(evals), [decode_g1(q) for q in proofs], decode_fr(beta))
        return {"z": b64_encode(z), "pi_0": b64_encode(pi0), "pi_1": b64_encode(pi1)}

    def master_verify(self, p):
        com, beta, alpha, z, pi0, pi1 = _need(p, "commitment", "beta", "alpha", "z", "pi_0", "pi_1")
        ok = self.backend.master_verify(decode_g1(com), decode_fr(beta), decode_fr(alpha), decode_fr(z),
                                        (decode_g1(pi0), decode_g1(pi1)))
        return {"valid": bool(ok)}

    # utils
    def random_poly(self, p):
        rows, cols = 1 << self.backend.m, 1 << (self.backend.n - self.backend.m)  # piano.rs:38-48
        return {"poly": [encode_fr_list(b"".join(random_fr() for _ in range(cols))) for _ in range(rows)]}

    def random_point(self, p):
        return {"point": b64_encode(random_fr())}

    def evaluate(self, p):
        poly, x = _need(p, "poly", "x")
        return {"y": b64_encode(self.backend.evaluate(decode_fr_list(poly), decode_fr(x)))}

    def fft(self, p):
        poly, left, inverse = _need(p, "poly", "left", "inverse")
        if not isinstance(left, bool) or not isinstance(inverse, bool):
            raise RpcError("invalid type: expected a boolean")
        return {"poly": encode_fr_list(self.backend.fft(decode_fr_list(poly), inverse=inverse, left=left))}

    def ping(self, p):
        return None  # RpcResult::Pong is a unit variant: untagged serde writes null

    def handle(self, req) -> Optional[dict]:
        """One decoded JSON request -> the JSON-able result (errors included)."""
        try:
            if not isinstance(req, dict) or "method" not in req:
                raise RpcError("missing field `method`")
            fn = self._methods.get(req["method"])
            if fn is None:
                raise RpcError("unknown variant `%s`, expected one of %s"
                               % (req["method"], ", ".join("`%s`" % k for k in self._methods)))
            return fn(req.get("params"))
        except RpcError as e:
            return {"message": str(e)}
        except Exception as e:  # backend errors: Result<_, String>::Err (piano.rs:143, 187, 250)
            return {"message": str(e)}

    def handle_bytes(self, body: bytes) -> bytes:
        try:
            req = json.loads(body)
        except Exception as e:
            return json.dumps({"message": str(e)}).encode()
        return json.dumps(self.handle(req), separators=(",", ":")).encode()


# ------------------------------------------------------------------ HTTP (rpc.rs:403-451, 489-512)
def make_server(handler: RpcHandler, host: str, port: int) -> ThreadingHTTPServer:
    class _H(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def _serve(self):
            n = int(self.headers.get("Content-Length") or 0)
            body = self.rfile.read(n) if n else b""
            out = handler.handle_bytes(body)
            self.send_response(200)
            self.send_header("Content-Length", str(len(out)))
            self.end_headers()
            self.wfile.write(out)

        do_POST = do_GET = do_PUT = do_PATCH = do_DELETE = _serve

        def log_message(self, fmt, *args):  # the reference logs through tracing at info level
            if os.environ.get("FB_RPC_LOG"):
                sys.stderr.write("rpc: " + fmt % args + "\n")

    return ThreadingHTTPServer((host, port), _H)


def serve_in_thread(backend, host: str = "127.0.0.1", port: int = 0):
    """Start the server on a daemon thread; returns (server, thread).  port 0 picks a free one."""
    srv = make_server(RpcHandler(backend), host, port)
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    return srv, th


# ------------------------------------------------------------------ command line (cli.rs)
def _parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="fourier-rpc", description="Fourier RPC server (B200 backend)")
    sub = ap.add_subparsers(dest="subcmd", required=True)
    run = sub.add_parser("run")
    run.add_argument("--setup-path")
    run.add_argument("--precompute-path")
    run.add_argument("--scale", type=int, default=20)
    run.add_argument("--machines-scale", type=int, default=1)
    run.add_argument("--host", default="localhost")
    run.add_argument("--port", type=int, default=1337)
    run.add_argument("--uncompressed", action="store_true")
    run.add_argument("--device", type=int, default=0)
    st = sub.add_parser("setup")
    st.add_argument("--setup-path", default="data/setup")
    st.add_argument("--precompute-path", default="data/precompute")
    st.add_argument("--scale", type=int, default=20)
    st.add_argument("--machines-scale", type=int, default=1)
    for flag in ("--overwrite", "--generate-setup", "--generate-precompute", "--uncompressed",
                 "--decompress-existing", "--compress-existing"):
        st.add_argument(flag, action="store_true")
    st.add_argument("--device", type=int, default=0)
    return ap


def _backend(args):
    from .engine import PianoBackend
    return PianoBackend(args.scale, args.machines_scale, device=args.device)


def cmd_run(args, backend_factory=None) -> int:
    # bind first: the reference client pings one second after spawning the process (fourier.py:113-122, 262-267);
    # a connection made while the setup is still loading waits in the listen queue instead of being refused
    handler = RpcHandler(None)
    srv = make_server(handler, args.host, args.port)
    be = (backend_factory or _backend)(args)
    if args.setup_path:  # omitted: generate a fresh setup (cli.rs:19-21, piano.rs:93-97)
        be.load_setup_from_file(args.setup_path, not args.uncompressed)
    else:
        be.setup(secrets.token_bytes(32), secrets.token_bytes(32))
    handler.backend = be
    print("fourier_b200 rpc listening on %s:%d (scale %d, machines_scale %d)"
          % (args.host, args.port, args.scale, args.machines_scale), flush=True)
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        srv.server_close()
        be.close()
    return 0


def can_proceed(args) -> bool:
    """SetupArgs::can_proceed (cli.rs:89-121): same checks, same order, same messages."""
    def error(msg):
        print(msg, file=sys.stderr)
        return False

    if os.path.exists(args.setup_path) and args.generate_setup and not args.overwrite:
        return error("File %s already exists, use --overwrite to overwrite" % args.setup_path)
    if os.path.exists(args.precompute_path) and args.generate_precompute and not args.overwrite:
        return error("File %s already exists, use --overwrite to overwrite" % args.precompute_path)
    if args.compress_existing and args.decompress_existing:
        return error("Cannot compress and decompress at the same time, choose one")
    if args.compress_existing and not args.uncompressed:
        return error("Cannot compress an already compressed file")
    if args.decompress_existing and args.uncompressed:
        return error("Cannot decompress an already decompressed file")
    return True


def cmd_setup(args) -> int:
    if not can_proceed(args):  # main.rs:13 asserts it: a refused combination ends the process with a failure
        return 1
    be = _backend(args)
    compressed = not args.uncompressed
    if args.compress_existing or args.decompress_existing:  # re-encode an existing file
        be.load_setup_from_file(args.setup_path, not args.compress_existing)
        be.save_setup_to_file(args.setup_path, bool(args.compress_existing))
    elif args.generate_setup:
        be.setup(secrets.token_bytes(32), secrets.token_bytes(32))
        os.makedirs(os.path.dirname(os.path.abspath(args.setup_path)), exist_ok=True)
        be.save_setup_to_file(args.setup_path, compressed)
    else:
        be.load_setup_from_file(args.setup_path, compressed)
    if args.generate_precompute:
        os.makedirs(os.path.dirname(os.path.abspath(args.precompute_path)), exist_ok=True)
        open(args.precompute_path, "wb").close()  # tables live in HBM; see the module docstring
    be.close()
    return 0


def main(argv=None, backend_factory=None) -> int:
    args = _parser().parse_args(argv)
    return cmd_run(args, backend_factory) if args.subcmd == "run" else cmd_setup(args)


if __name__ == "__main__":
    sys.exit(main())
