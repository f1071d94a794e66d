"""ctypes wrapper of the C oracle (oracle/c/oracle.c -> liboracle.so).

TEST INFRASTRUCTURE ONLY - see the header of oracle/c/oracle.c.  Values cross in
the reference's wire encodings (Fr 32-byte BE, G1 48-byte compressed / 96-byte
uncompressed), so results compare byte-for-byte with the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "c", "oracle.c")
LIB = os.path.join(_HERE, "c", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
        subprocess.check_call([gcc, "-O3", "-mbmi2", "-madx", "-fopenmp", "-shared", "-fPIC", "-o", LIB, SRC])
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        l = C.CDLL(LIB)
        l.or_bench_ntt.restype = C.c_double
        l.or_bench_msm.restype = C.c_double
        for name in ("or_ntt", "or_eval", "or_open_scalars", "or_lagrange_at", "or_g1_mul_gen", "or_msm", "or_g1_sum",
                     "or_g1_recode", "or_bench_commit_open", "or_threads"):
            getattr(l, name).restype = C.c_int
        _lib = l
    return _lib


def _p(b):
    if hasattr(b, "ctypes"):
        return C.c_void_p(b.ctypes.data)
    if isinstance(b, (bytes,)):
        return C.cast(C.c_char_p(b), C.c_void_p)
    return C.cast(b, C.c_void_p)


def _chk(rc, what):
    if rc:
        raise ValueError("%s failed: %s" % (what, {1: "Invalid scalar", 2: "bad length", 4: "bad point"}.get(rc, rc)))


def threads() -> int:
    return lib().or_threads()


def set_threads(n: int):
    lib().or_set_threads(C.c_int(n))


def ntt(data_be, inverse: bool = False) -> bytes:
    n = len(data_be) // 32 if isinstance(data_be, (bytes, bytearray)) else data_be.size // 32
    out = C.create_string_buffer(32 * n)
    _chk(lib().or_ntt(_p(data_be), C.c_uint64(n), C.c_int(int(inverse)), out), "or_ntt")
    return out.raw


def evaluate(poly_be, x_be: bytes) -> bytes:
    n = len(poly_be) // 32 if isinstance(poly_be, (bytes, bytearray)) else poly_be.size // 32
    out = C.create_string_buffer(32)
    _chk(lib().or_eval(_p(poly_be), C.c_uint64(n), _p(x_be), out), "or_eval")
    return out.raw


def open_scalars(coeffs_be, alpha_be: bytes):
    """-> (y, NTT(q || 0)) as bytes: the field half of worker_open (piano.rs:193-233)."""
    n = len(coeffs_be) // 32 if isinstance(coeffs_be, (bytes, bytearray)) else coeffs_be.size // 32
    y = C.create_string_buffer(32)
    q = C.create_string_buffer(32 * n)
    _chk(lib().or_open_scalars(_p(coeffs_be), C.c_uint64(n), _p(alpha_be), y, q), "or_open_scalars")
    return y.raw, q.raw


def lagrange_at(tau_be: bytes, log_n: int) -> bytes:
    out = C.create_string_buffer(32 << log_n)
    _chk(lib().or_lagrange_at(_p(tau_be), C.c_int(log_n), out), "or_lagrange_at")
    return out.raw


def g1_mul_gen(scalars_be) -> bytes:
    """96-byte uncompressed s_k * G for every scalar."""
    n = len(scalars_be) // 32 if isinstance(scalars_be, (bytes, bytearray)) else scalars_be.size // 32
    out = C.create_string_buffer(96 * n)
    _chk(lib().or_g1_mul_gen(_p(scalars_be), C.c_uint64(n), out), "or_g1_mul_gen")
    return out.raw


def msm(points96, scalars_be) -> bytes:
    n = len(scalars_be) // 32 if isinstance(scalars_be, (bytes, bytearray)) else scalars_be.size // 32
    out = C.create_string_buffer(48)
    _chk(lib().or_msm(_p(points96), _p(scalars_be), C.c_uint64(n), out), "or_msm")
    return out.raw


def g1_sum(points48: bytes) -> bytes:
    out = C.create_string_buffer(48)
    _chk(lib().or_g1_sum(_p(points48), C.c_uint64(len(points48) // 48), out), "or_g1_sum")
    return out.raw


def g1_recode(pt: bytes, out_fmt: int) -> bytes:
    out = C.create_string_buffer(out_fmt)
    _chk(lib().or_g1_recode(_p(pt), C.c_int(len(pt)), out, C.c_int(out_fmt)), "or_g1_recode")
    return out.raw


def bench_ntt(data_be, inverse: bool, reps: int) -> float:
    n = len(data_be) // 32 if isinstance(data_be, (bytes, bytearray)) else data_be.size // 32
    return lib().or_bench_ntt(_p(data_be), C.c_uint64(n), C.c_int(int(inverse)), C.c_int(reps))


def bench_msm(points96, scalars_be, reps: int):
    n = len(scalars_be) // 32 if isinstance(scalars_be, (bytes, bytearray)) else scalars_be.size // 32
    out = C.create_string_buffer(48)
    t = lib().or_bench_msm(_p(points96), _p(scalars_be), C.c_uint64(n), C.c_int(reps), out)
    return t, out.raw


def bench_commit_open(points96, coeffs_be, alpha_be: bytes, reps: int):
    """-> (commit_s, open_s, commit48, y32, proof48)"""
    n = len(coeffs_be) // 32 if isinstance(coeffs_be, (bytes, bytearray)) else coeffs_be.size // 32
    c = C.create_string_buffer(48)
    y = C.create_string_buffer(32)
    p = C.create_string_buffer(48)
    times = (C.c_double * 2)()
    _chk(lib().or_bench_commit_open(_p(points96), _p(coeffs_be), C.c_uint64(n), _p(alpha_be), C.c_int(reps), c, y, p,
                                    times), "or_bench_commit_open")
    return times[0], times[1], c.raw, y.raw, p.raw
