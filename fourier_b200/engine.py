"""Host-side mirror of fourier's PianoBackend (reference src/engine/piano.rs:26-465)
over the C ABI.  Method names, argument meaning and error behaviour follow the
reference; the arithmetic runs in libfourier_b200.so on the GPU.

Values cross this layer in the reference's wire encodings: Fr = 32-byte
big-endian ``bytes`` (FsFr::to_bytes), G1 = 48-byte compressed ``bytes``
(FsG1::to_bytes).  Vectors of scalars may be given as a list of ``bytes`` / ``int``
or as one contiguous buffer (bytes, numpy uint8, pinned torch uint8 tensor).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence, Tuple, Union

from . import _lib

FR_BYTES = 32
G1_BYTES = 48

ERR_NAMES = {1: "CUDA", 2: "argument", 3: "Invalid scalar", 4: "invalid G1 point", 5: "state", 6: "I/O"}


class FourierError(RuntimeError):
    """Result<_, String>::Err of the reference (piano.rs:143, 187, 250)."""

    def __init__(self, code: int, message: str):
        super().__init__(message or ERR_NAMES.get(code, "error %d" % code))
        self.code = code


ScalarVec = Union[bytes, bytearray, memoryview, Sequence[Union[int, bytes]]]


def pack_scalars(v: ScalarVec):
    """-> (buffer, count).  Lists are packed; contiguous buffers pass through."""
    if isinstance(v, (bytes, bytearray, memoryview)):
        if len(v) % FR_BYTES:
            raise FourierError(2, "Invalid byte length. Expected a multiple of 32, got %d" % len(v))
        return v, len(v) // FR_BYTES
    if hasattr(v, "data_ptr") or hasattr(v, "ctypes"):
        nbytes = v.numel() * v.element_size() if hasattr(v, "numel") else v.nbytes
        if nbytes % FR_BYTES:
            raise FourierError(2, "Invalid byte length. Expected a multiple of 32, got %d" % nbytes)
        return v, nbytes // FR_BYTES
    out = bytearray()
    for x in v:
        if isinstance(x, int):
            if x < 0 or x >> 256:
                raise FourierError(3, "Invalid scalar")
            out += x.to_bytes(FR_BYTES, "big")
        else:
            if len(x) != FR_BYTES:
                raise FourierError(2, "Invalid byte length. Expected 32, got %d" % len(x))
            out += x
    return bytes(out), len(out) // FR_BYTES


def pack_scalar(x: Union[int, bytes]) -> bytes:
    buf, n = pack_scalars([x])
    return bytes(buf)


def pack_points(v) -> Tuple[bytes, int]:
    if isinstance(v, (bytes, bytearray, memoryview)):
        if len(v) % G1_BYTES:
            raise FourierError(2, "Invalid byte length. Expected a multiple of 48, got %d" % len(v))
        return bytes(v), len(v) // G1_BYTES
    for p in v:
        if len(p) != G1_BYTES:
            raise FourierError(2, "Invalid byte length. Expected 48, got %d" % len(p))
    return b"".join(v), len(v)


class Bases:
    """A resident array of G1 points (optionally with window tables)."""

    def __init__(self, backend: "PianoBackend", handle: int, count: int):
        self._b, self._h, self.count = backend, handle, count

    def precompute(self, window_bits: int = 0) -> "Bases":
        self._b._chk(self._b._lib.fb_bases_precompute(self._b._ctx, self._h, window_bits))
        return self

    def window(self) -> Tuple[int, int]:
        """(window bits, number of windows) of the table as built; (0, 0) without one."""
        c, w = C.c_int(0), C.c_int(0)
        self._b._chk(self._b._lib.fb_bases_get_window(self._b._ctx, self._h, C.byref(c), C.byref(w)))
        return c.value, w.value

    def point(self, index: int) -> bytes:
        out = C.create_string_buffer(G1_BYTES)
        self._b._chk(self._b._lib.fb_bases_get(self._b._ctx, self._h, index, C.addressof(out)))
        return out.raw

    def free(self):
        if self._h:
            self._b._lib.fb_bases_free(self._b._ctx, self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PianoBackend:
    """PianoBackend::new(cfg) (piano.rs:82-85): scale n, machines_scale m."""

    def __init__(self, scale: int, machines_scale: int = 0, device: int = 0, window_bits: int = 0):
        self._lib = _lib.load()
        if machines_scale > scale:
            raise FourierError(2, "m must be less than or equal to n")  # piano.rs:1062-1064
        ctx = C.c_void_p()
        rc = self._lib.fb_ctx_create(scale, machines_scale, device, C.byref(ctx))
        if rc:
            raise FourierError(rc, "fb_ctx_create failed (no usable CUDA device %d?)" % device)
        self._ctx = ctx
        self.n, self.m, self.t = scale, machines_scale, scale - machines_scale
        self.T, self.M = 1 << self.t, 1 << self.m
        if window_bits:
            self._chk(self._lib.fb_ctx_set_window(self._ctx, window_bits))

    # -- plumbing -------------------------------------------------------------
    def _chk(self, rc: int):
        if rc:
            msg = self._lib.fb_last_error(self._ctx)
            raise FourierError(rc, msg.decode() if msg else "")

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.fb_ctx_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        self._chk(self._lib.fb_ctx_set_stream(self._ctx, cuda_stream))

    def sync(self):
        self._chk(self._lib.fb_ctx_sync(self._ctx))

    def set_lanes(self, lanes: int):
        """lanes >= 2: the heavy dev_* calls rotate over that many internal streams (overlap); see join()."""
        self._chk(self._lib.fb_ctx_set_lanes(self._ctx, lanes))

    def set_msm_split(self, parts: int):
        """parts >= 2: every large MSM runs as that many sub-MSMs on as many streams of its lane."""
        self._chk(self._lib.fb_ctx_set_msm_split(self._ctx, parts))

    def set_pair_mode(self, mode: int):
        """commit + open of one row: 0 by size, 1 one batch-of-two MSM, 2 two MSMs on two streams."""
        self._chk(self._lib.fb_ctx_set_pair_mode(self._ctx, mode))

    def join(self):
        """The caller's stream waits (on the device) for every dev_* call in flight on the lanes."""
        self._chk(self._lib.fb_ctx_join(self._ctx))

    def set_window(self, window_bits: int):
        """Window bits of the tables built by the next setup / load (0: by row length, -1: no tables)."""
        self._chk(self._lib.fb_ctx_set_window(self._ctx, window_bits))

    def window(self) -> Tuple[int, int]:
        """(window bits, number of windows) of the resident row tables as built; (0, 0) without tables."""
        c, w = C.c_int(0), C.c_int(0)
        self._chk(self._lib.fb_ctx_get_window(self._ctx, C.byref(c), C.byref(w)))
        return c.value, w.value

    def set_open_mode(self, literal: bool):
        """False (default): evaluation-form opening; True: the reference's literal iNTT -> eval -> divide -> NTT."""
        self._chk(self._lib.fb_ctx_set_open_mode(self._ctx, int(bool(literal))))

    def set_staging(self, direct: bool):
        """Pageable host inputs: False (default) through the library's pinned ring, True straight to cudaMemcpyAsync."""
        self._chk(self._lib.fb_ctx_set_staging(self._ctx, int(bool(direct))))

    def launch_count(self) -> int:
        return self._lib.fb_launch_count(self._ctx)

    def profile(self, on: bool = True):
        self._chk(self._lib.fb_prof_enable(self._ctx, int(on)))

    def profile_read(self) -> dict:
        """{phase: (total_ms, count)} measured with CUDA events on the launching stream."""
        import json
        buf = C.create_string_buffer(1 << 16)
        self._chk(self._lib.fb_prof_read(self._ctx, buf, len(buf)))
        return {k: tuple(v) for k, v in json.loads(buf.value.decode()).items()}

    def msm_stats(self) -> dict:
        """Work counters of the MSMs run while profiling was on (reset by the read)."""
        out = (C.c_uint64 * 8)()
        self._chk(self._lib.fb_msm_stats(self._ctx, out))
        return {"msms": out[0], "entries": out[1], "round_slots": out[2], "survivors": out[3], "buckets": out[4]}

    # -- setup (piano.rs:87-122, 883-1037, 649-846) -----------------------------
    def setup(self, secret_x: bytes, secret_y: bytes, row0: int = 0, nrows: int = 0):
        """generate_trusted_setup(secrets); rows [row0, row0+nrows) stay resident."""
        assert len(secret_x) == 32 and len(secret_y) == 32
        self._chk(self._lib.fb_srs_generate(self._ctx, _lib.addr(secret_x), _lib.addr(secret_y), row0, nrows))

    def load_setup_from_file(self, path: str, compressed: bool = True, row0: int = 0, nrows: int = 0):
        self._chk(self._lib.fb_srs_load_file(self._ctx, path.encode(), int(compressed), row0, nrows))

    def load_setup_from_bytes(self, data: bytes, compressed: bool = True, row0: int = 0, nrows: int = 0):
        self._chk(self._lib.fb_srs_load_bytes(self._ctx, _lib.addr(data), len(data), int(compressed), row0, nrows))

    def save_setup_to_file(self, path: str, compressed: bool = True):
        self._chk(self._lib.fb_srs_save_file(self._ctx, path.encode(), int(compressed)))

    def setup_to_bytes(self, compressed: bool = True) -> bytes:
        n = C.c_size_t()
        self._chk(self._lib.fb_srs_to_bytes(self._ctx, int(compressed), None, 0, C.byref(n)))
        out = C.create_string_buffer(n.value)
        self._chk(self._lib.fb_srs_to_bytes(self._ctx, int(compressed), C.addressof(out), n.value, C.byref(n)))
        return out.raw

    # -- worker side ---------------------------------------------------------------
    def worker_commit(self, i: int, coeffs: ScalarVec) -> bytes:
        """piano.rs:143-161."""
        buf, n = pack_scalars(coeffs)
        out = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_worker_commit(self._ctx, i, _lib.addr(buf), n, C.addressof(out)))
        return out.raw

    def worker_open(self, i: int, coeffs: ScalarVec, alpha: Union[int, bytes]) -> Tuple[bytes, bytes]:
        """piano.rs:187-238 -> (y, proof)."""
        buf, n = pack_scalars(coeffs)
        a = pack_scalar(alpha)
        y = C.create_string_buffer(FR_BYTES)
        pi = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_worker_open(self._ctx, i, _lib.addr(buf), n, _lib.addr(a), C.addressof(y), C.addressof(pi)))
        return y.raw, pi.raw

    def worker_commit_open(self, i: int, coeffs: ScalarVec, alpha) -> Tuple[bytes, bytes, bytes]:
        """worker_commit + worker_open of the same row in one pass -> (commitment, y, proof)."""
        buf, n = pack_scalars(coeffs)
        a = pack_scalar(alpha)
        cm, y, pi = C.create_string_buffer(G1_BYTES), C.create_string_buffer(FR_BYTES), C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_worker_commit_open(self._ctx, i, _lib.addr(buf), n, _lib.addr(a), C.addressof(cm),
                                                  C.addressof(y), C.addressof(pi)))
        return cm.raw, y.raw, pi.raw

    def worker_commit_open_partial(self, i: int, coeffs: ScalarVec, alpha) -> Tuple[bytes, bytes, bytes]:
        buf, n = pack_scalars(coeffs)
        a = pack_scalar(alpha)
        cm, y, pi = C.create_string_buffer(96), C.create_string_buffer(FR_BYTES), C.create_string_buffer(96)
        self._chk(self._lib.fb_worker_commit_open_partial(self._ctx, i, _lib.addr(buf), n, _lib.addr(a), C.addressof(cm),
                                                          C.addressof(y), C.addressof(pi)))
        return cm.raw, y.raw, pi.raw

    def commit_rows(self, row0: int, nrows: int, coeffs: ScalarVec) -> List[bytes]:
        buf, n = pack_scalars(coeffs)
        if nrows == 0 or n % nrows:
            raise FourierError(2, "coefficient count is not a multiple of the row count")
        out = C.create_string_buffer(G1_BYTES * nrows)
        self._chk(self._lib.fb_commit_rows(self._ctx, row0, nrows, _lib.addr(buf), n // nrows, C.addressof(out)))
        return [out.raw[k * G1_BYTES:(k + 1) * G1_BYTES] for k in range(nrows)]

    def open_rows(self, row0: int, nrows: int, coeffs: ScalarVec, alpha) -> Tuple[List[bytes], List[bytes]]:
        buf, n = pack_scalars(coeffs)
        if nrows == 0 or n % nrows:
            raise FourierError(2, "coefficient count is not a multiple of the row count")
        a = pack_scalar(alpha)
        ys = C.create_string_buffer(FR_BYTES * nrows)
        pis = C.create_string_buffer(G1_BYTES * nrows)
        self._chk(self._lib.fb_open_rows(self._ctx, row0, nrows, _lib.addr(buf), n // nrows, _lib.addr(a),
                                         C.addressof(ys), C.addressof(pis)))
        return ([ys.raw[k * 32:(k + 1) * 32] for k in range(nrows)],
                [pis.raw[k * 48:(k + 1) * 48] for k in range(nrows)])

    def commit_open_rows(self, row0: int, nrows: int, coeffs: ScalarVec, alpha) -> Tuple[List[bytes], List[bytes], List[bytes]]:
        """commit_rows + open_rows in one call (one upload) -> (commitments, evaluations, proofs)."""
        buf, n = pack_scalars(coeffs)
        if nrows == 0 or n % nrows:
            raise FourierError(2, "coefficient count is not a multiple of the row count")
        a = pack_scalar(alpha)
        cs = C.create_string_buffer(G1_BYTES * nrows)
        ys = C.create_string_buffer(FR_BYTES * nrows)
        pis = C.create_string_buffer(G1_BYTES * nrows)
        self._chk(self._lib.fb_commit_open_rows(self._ctx, row0, nrows, _lib.addr(buf), n // nrows, _lib.addr(a),
                                                C.addressof(cs), C.addressof(ys), C.addressof(pis)))
        return ([cs.raw[k * 48:(k + 1) * 48] for k in range(nrows)], [ys.raw[k * 32:(k + 1) * 32] for k in range(nrows)],
                [pis.raw[k * 48:(k + 1) * 48] for k in range(nrows)])

    def dev_commit_open_rows(self, row0: int, nrows: int, d_coeffs, length: int, d_alpha, d_out48s, d_out_ys, d_out_proofs):
        self._chk(self._lib.fb_dev_commit_open_rows(self._ctx, row0, nrows, _lib.addr(d_coeffs), length, _lib.addr(d_alpha),
                                                    _lib.addr(d_out48s), _lib.addr(d_out_ys), _lib.addr(d_out_proofs)))

    # -- column shards (one MSM over several GPUs) ----------------------------------
    def set_col_shard(self, col0: int, ncols: int):
        """Keep only columns [col0, col0+ncols) resident (applies at the next setup/load)."""
        self._chk(self._lib.fb_ctx_set_col_shard(self._ctx, col0, ncols))

    def worker_commit_partial(self, i: int, coeffs: ScalarVec) -> bytes:
        buf, n = pack_scalars(coeffs)
        out = C.create_string_buffer(96)
        self._chk(self._lib.fb_worker_commit_partial(self._ctx, i, _lib.addr(buf), n, C.addressof(out)))
        return out.raw

    def worker_open_partial(self, i: int, coeffs: ScalarVec, alpha) -> Tuple[bytes, bytes]:
        buf, n = pack_scalars(coeffs)
        a = pack_scalar(alpha)
        y = C.create_string_buffer(FR_BYTES)
        pi = C.create_string_buffer(96)
        self._chk(self._lib.fb_worker_open_partial(self._ctx, i, _lib.addr(buf), n, _lib.addr(a), C.addressof(y),
                                                   C.addressof(pi)))
        return y.raw, pi.raw

    def commit_open_shard_begin(self, i: int, coeffs_slice: ScalarVec, length: int, alpha) -> bytes:
        """First half of a commit + open whose field half is sharded too: the slice of the row this context holds -> 128
        opaque bytes of state, to be gathered over the shards."""
        buf, n = pack_scalars(coeffs_slice)
        a = pack_scalar(alpha)
        st = C.create_string_buffer(128)
        self._chk(self._lib.fb_worker_commit_open_shard_begin(self._ctx, i, _lib.addr(buf), n, length, _lib.addr(a), C.addressof(st)))
        return st.raw

    def commit_open_shard_finish(self, i: int, states: bytes, alpha) -> Tuple[bytes, bytes, bytes]:
        """Second half: the states of all shards -> (partial commitment 96 B, y, partial proof 96 B)."""
        a = pack_scalar(alpha)
        cm, y, pi = C.create_string_buffer(96), C.create_string_buffer(FR_BYTES), C.create_string_buffer(96)
        self._chk(self._lib.fb_worker_commit_open_shard_finish(self._ctx, i, _lib.addr(states), len(states) // 128, _lib.addr(a),
                                                               C.addressof(cm), C.addressof(y), C.addressof(pi)))
        return cm.raw, y.raw, pi.raw

    def dev_commit_open_shard_begin(self, i: int, d_slice, length: int, d_alpha, d_state128):
        self._chk(self._lib.fb_dev_commit_open_shard_begin(self._ctx, i, _lib.addr(d_slice), length, _lib.addr(d_alpha),
                                                           _lib.addr(d_state128)))

    def dev_commit_open_shard_finish(self, i: int, d_slice, length: int, d_alpha, d_states, nparts: int, d_out224):
        self._chk(self._lib.fb_dev_commit_open_shard_finish(self._ctx, i, _lib.addr(d_slice), length, _lib.addr(d_alpha),
                                                            _lib.addr(d_states), nparts, _lib.addr(d_out224)))

    def dev_combine_commit_open(self, d_parts224, nparts: int, d_out128):
        """Adds the shards' partials on the device: nparts x (96 | 32 | 96) -> 48 | 32 | 48."""
        self._chk(self._lib.fb_dev_combine_commit_open(self._ctx, _lib.addr(d_parts224), nparts, _lib.addr(d_out128)))

    def dev_worker_commit_partial(self, i: int, d_coeffs, count: int, d_out96):
        self._chk(self._lib.fb_dev_worker_commit_partial(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_out96)))

    def dev_worker_open_partial(self, i: int, d_coeffs, count: int, d_alpha, d_out128):
        self._chk(self._lib.fb_dev_worker_open_partial(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_alpha),
                                                       _lib.addr(d_out128)))

    # -- master side ---------------------------------------------------------------
    def master_commit(self, commitments) -> bytes:
        """piano.rs:164-175."""
        buf, n = pack_points(commitments)
        out = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_master_commit(self._ctx, _lib.addr(buf), n, C.addressof(out)))
        return out.raw

    def master_open(self, evals: ScalarVec, proofs, beta) -> Tuple[bytes, Tuple[bytes, bytes]]:
        """piano.rs:250-298 -> (z, (pi_0, pi_1))."""
        ebuf, n = pack_scalars(evals)
        pbuf, np_ = pack_points(proofs)
        if n != np_:
            raise FourierError(2, "evals and proofs differ in length")
        b = pack_scalar(beta)
        z = C.create_string_buffer(FR_BYTES)
        p0 = C.create_string_buffer(G1_BYTES)
        p1 = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_master_open(self._ctx, _lib.addr(ebuf), _lib.addr(pbuf), n, _lib.addr(b),
                                           C.addressof(z), C.addressof(p0), C.addressof(p1)))
        return z.raw, (p0.raw, p1.raw)

    # -- verification ---------------------------------------------------------------------
    def worker_verify(self, i: int, commitment: bytes, alpha, y, pi: bytes) -> bool:
        """piano.rs:309-356: e(C - y g^{R_i(tau_Y)}, g2) == e(pi, g2^{tau_X - alpha})."""
        cb, _ = pack_points([commitment])
        pb, _ = pack_points([pi])
        a, yb = pack_scalar(alpha), pack_scalar(y)
        ok = C.c_int(0)
        self._chk(self._lib.fb_worker_verify(self._ctx, i, _lib.addr(cb), _lib.addr(a), _lib.addr(yb), _lib.addr(pb),
                                             C.byref(ok)))
        return bool(ok.value)

    def master_verify(self, commitment: bytes, beta, alpha, z, pi_f: Tuple[bytes, bytes]) -> bool:
        """piano.rs:393-420: e(C - z g, g2) == e(pi_0, g2^{tau_X - alpha}) e(pi_1, g2^{tau_Y - beta})."""
        cb, _ = pack_points([commitment])
        p0, _ = pack_points([pi_f[0]])
        p1, _ = pack_points([pi_f[1]])
        b, a, zb = pack_scalar(beta), pack_scalar(alpha), pack_scalar(z)
        ok = C.c_int(0)
        self._chk(self._lib.fb_master_verify(self._ctx, _lib.addr(cb), _lib.addr(b), _lib.addr(a), _lib.addr(zb),
                                             _lib.addr(p0), _lib.addr(p1), C.byref(ok)))
        return bool(ok.value)

    # -- field-side RPC helpers -------------------------------------------------------
    def fft(self, data: ScalarVec, inverse: bool = False, left: bool = True) -> bytes:
        """fft_left / fft_right (piano.rs:1096-1103) as served by the `fft` RPC."""
        buf, n = pack_scalars(data)
        out = C.create_string_buffer(FR_BYTES * max(n, 1))
        self._chk(self._lib.fb_fft(self._ctx, int(left), int(inverse), _lib.addr(buf), n, C.addressof(out)))
        return out.raw[: FR_BYTES * n]

    def ntt(self, data: ScalarVec, inverse: bool = False, out=None):
        """Natural-order NTT of 32-byte big-endian scalars on host buffers.  `out`: a caller-owned host buffer
        of 32*n bytes (numpy / torch, e.g. pinned) to receive the result in place of a returned bytes object."""
        buf, n = pack_scalars(data)
        if out is not None:
            self._chk(self._lib.fb_ntt(self._ctx, int(inverse), _lib.addr(buf), n, _lib.addr(out)))
            return None
        res = C.create_string_buffer(FR_BYTES * max(n, 1))
        self._chk(self._lib.fb_ntt(self._ctx, int(inverse), _lib.addr(buf), n, C.addressof(res)))
        return res.raw[: FR_BYTES * n]

    def evaluate(self, poly: ScalarVec, x) -> bytes:
        """FsPoly::eval as served by the `evaluate` RPC (rpc.rs:60-64)."""
        buf, n = pack_scalars(poly)
        xb = pack_scalar(x)
        out = C.create_string_buffer(FR_BYTES)
        self._chk(self._lib.fb_evaluate(self._ctx, _lib.addr(buf), n, _lib.addr(xb), C.addressof(out)))
        return out.raw

    # -- raw kernels ---------------------------------------------------------------------
    def bases_from_points(self, points, fmt: int = 48) -> Bases:
        if isinstance(points, (list, tuple)):
            points = b"".join(points)
        h = C.c_void_p()
        n = len(points) // fmt
        self._chk(self._lib.fb_bases_from_points(self._ctx, _lib.addr(points), n, fmt, C.byref(h)))
        return Bases(self, h, n)

    def bases_from_scalars(self, scalars: ScalarVec) -> Bases:
        buf, n = pack_scalars(scalars)
        h = C.c_void_p()
        self._chk(self._lib.fb_bases_from_scalars(self._ctx, _lib.addr(buf), n, C.byref(h)))
        return Bases(self, h, n)

    def msm(self, bases: Bases, scalars: ScalarVec) -> bytes:
        buf, n = pack_scalars(scalars)
        out = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_msm(self._ctx, bases._h, _lib.addr(buf), n, C.addressof(out)))
        return out.raw

    def msm_partial(self, bases: Bases, scalars: ScalarVec) -> bytes:
        buf, n = pack_scalars(scalars)
        out = C.create_string_buffer(96)
        self._chk(self._lib.fb_msm_partial(self._ctx, bases._h, _lib.addr(buf), n, C.addressof(out)))
        return out.raw

    def g1_sum(self, points, fmt: int = 48) -> bytes:
        if isinstance(points, (list, tuple)):
            points = b"".join(points)
        out = C.create_string_buffer(G1_BYTES)
        self._chk(self._lib.fb_g1_sum(self._ctx, _lib.addr(points), len(points) // fmt, fmt, C.addressof(out)))
        return out.raw

    # -- device-resident (addresses of device buffers, e.g. torch tensors) ---------------
    def dev_fr_from_be(self, d_in, d_out, count: int):
        self._chk(self._lib.fb_dev_fr_from_be(self._ctx, _lib.addr(d_in), _lib.addr(d_out), count))

    def dev_fr_to_be(self, d_in, d_out, count: int):
        self._chk(self._lib.fb_dev_fr_to_be(self._ctx, _lib.addr(d_in), _lib.addr(d_out), count))

    def dev_ntt(self, d_in, d_out, count: int, inverse: bool = False):
        self._chk(self._lib.fb_dev_ntt(self._ctx, int(inverse), _lib.addr(d_in), _lib.addr(d_out), count))

    def dev_msm(self, bases: Bases, d_scalars, count: int, d_out48):
        self._chk(self._lib.fb_dev_msm(self._ctx, bases._h, _lib.addr(d_scalars), count, _lib.addr(d_out48)))

    def dev_msm_partial(self, bases: Bases, d_scalars, count: int, d_out96):
        self._chk(self._lib.fb_dev_msm_partial(self._ctx, bases._h, _lib.addr(d_scalars), count, _lib.addr(d_out96)))

    def dev_commit_rows(self, row0: int, nrows: int, d_coeffs, length: int, d_out48s):
        self._chk(self._lib.fb_dev_commit_rows(self._ctx, row0, nrows, _lib.addr(d_coeffs), length, _lib.addr(d_out48s)))

    def dev_open_rows(self, row0: int, nrows: int, d_coeffs, length: int, d_alpha, d_out_ys, d_out_proofs):
        self._chk(self._lib.fb_dev_open_rows(self._ctx, row0, nrows, _lib.addr(d_coeffs), length, _lib.addr(d_alpha),
                                             _lib.addr(d_out_ys), _lib.addr(d_out_proofs)))

    def dev_worker_commit(self, i: int, d_coeffs, count: int, d_out48):
        self._chk(self._lib.fb_dev_worker_commit(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_out48)))

    def dev_worker_commit_open(self, i: int, d_coeffs, count: int, d_alpha, d_out128):
        self._chk(self._lib.fb_dev_worker_commit_open(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_alpha),
                                                      _lib.addr(d_out128)))

    def dev_worker_commit_open_partial(self, i: int, d_coeffs, count: int, d_alpha, d_out224):
        self._chk(self._lib.fb_dev_worker_commit_open_partial(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_alpha),
                                                              _lib.addr(d_out224)))

    def dev_worker_open(self, i: int, d_coeffs, count: int, d_alpha, d_out80):
        self._chk(self._lib.fb_dev_worker_open(self._ctx, i, _lib.addr(d_coeffs), count, _lib.addr(d_alpha),
                                               _lib.addr(d_out80)))


def pairing_check(g1_points96, g2_points192) -> bool:
    """prod_k e(g1[k], g2[k]) == 1 on serialized points (pairings_verify, piano.rs:422-464).
    Host arithmetic inside the library: needs neither a context nor a GPU."""
    lib = _lib.load()
    g1, g2 = b"".join(g1_points96), b"".join(g2_points192)
    n = len(g1) // 96
    if len(g1) != 96 * n or len(g2) != 192 * n:
        raise FourierError(2, "expected n x 96 bytes of G1 and n x 192 bytes of G2")
    ok = C.c_int(0)
    rc = lib.fb_pairing_check(_lib.addr(g1), _lib.addr(g2), n, C.byref(ok))
    if rc:
        raise FourierError(rc, "malformed point")
    return bool(ok.value)


def pairing_value(g1_points96, g2_points192) -> bytes:
    """(prod_k e(g1[k], g2[k]))^3 as 12 big-endian Fp coefficients (see fb_pairing_value)."""
    lib = _lib.load()
    g1, g2 = b"".join(g1_points96), b"".join(g2_points192)
    n = len(g1) // 96
    out = C.create_string_buffer(576)
    rc = lib.fb_pairing_value(_lib.addr(g1), _lib.addr(g2), n, C.addressof(out))
    if rc:
        raise FourierError(rc, "malformed point")
    return out.raw
