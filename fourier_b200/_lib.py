"""ctypes binding of libfourier_b200.so (include/fourier_b200.h).

There is no fallback: if the shared library is missing or a CUDA device is not
usable, importing / creating a context raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FB_LIB_PATH: an alternative build of the same library (A/B runs of compile-time variants); never a fallback
LIB_PATH = os.environ.get("FB_LIB_PATH") or os.path.join(_HERE, "libfourier_b200.so")

# every symbol include/fourier_b200.h declares: (name, restype, argtypes)
_u8p = C.c_void_p  # byte buffers are passed as addresses (bytes, bytearray, numpy, torch all work)
_ctx = C.c_void_p
_bases = C.c_void_p
SYMBOLS = [
    ("fb_ctx_create", C.c_int, [C.c_uint32, C.c_uint32, C.c_int, C.POINTER(_ctx)]),
    ("fb_ctx_destroy", None, [_ctx]),
    ("fb_last_error", C.c_char_p, [_ctx]),
    ("fb_version", C.c_char_p, []),
    ("fb_ctx_set_stream", C.c_int, [_ctx, C.c_void_p]),
    ("fb_ctx_sync", C.c_int, [_ctx]),
    ("fb_ctx_set_window", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_get_window", C.c_int, [_ctx, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("fb_ctx_set_open_mode", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_set_staging", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_set_msm_split", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_set_pair_mode", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_set_lanes", C.c_int, [_ctx, C.c_int]),
    ("fb_ctx_join", C.c_int, [_ctx]),
    ("fb_srs_generate", C.c_int, [_ctx, _u8p, _u8p, C.c_uint64, C.c_uint64]),
    ("fb_srs_load_file", C.c_int, [_ctx, C.c_char_p, C.c_int, C.c_uint64, C.c_uint64]),
    ("fb_srs_load_bytes", C.c_int, [_ctx, _u8p, C.c_size_t, C.c_int, C.c_uint64, C.c_uint64]),
    ("fb_srs_save_file", C.c_int, [_ctx, C.c_char_p, C.c_int]),
    ("fb_srs_to_bytes", C.c_int, [_ctx, C.c_int, _u8p, C.c_size_t, C.POINTER(C.c_size_t)]),
    ("fb_worker_commit", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p]),
    ("fb_worker_open", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p]),
    ("fb_commit_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, _u8p, C.c_uint64, _u8p]),
    ("fb_open_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p]),
    ("fb_master_commit", C.c_int, [_ctx, _u8p, C.c_uint64, _u8p]),
    ("fb_master_open", C.c_int, [_ctx, _u8p, _u8p, C.c_uint64, _u8p, _u8p, _u8p, _u8p]),
    ("fb_worker_verify", C.c_int, [_ctx, C.c_uint64, _u8p, _u8p, _u8p, _u8p, C.POINTER(C.c_int)]),
    ("fb_master_verify", C.c_int, [_ctx, _u8p, _u8p, _u8p, _u8p, _u8p, _u8p, C.POINTER(C.c_int)]),
    ("fb_pairing_check", C.c_int, [_u8p, _u8p, C.c_uint64, C.POINTER(C.c_int)]),
    ("fb_pairing_value", C.c_int, [_u8p, _u8p, C.c_uint64, _u8p]),
    ("fb_fft", C.c_int, [_ctx, C.c_int, C.c_int, _u8p, C.c_uint64, _u8p]),
    ("fb_evaluate", C.c_int, [_ctx, _u8p, C.c_uint64, _u8p, _u8p]),
    ("fb_bases_from_points", C.c_int, [_ctx, _u8p, C.c_uint64, C.c_int, C.POINTER(_bases)]),
    ("fb_bases_from_scalars", C.c_int, [_ctx, _u8p, C.c_uint64, C.POINTER(_bases)]),
    ("fb_bases_precompute", C.c_int, [_ctx, _bases, C.c_int]),
    ("fb_bases_get", C.c_int, [_ctx, _bases, C.c_uint64, _u8p]),
    ("fb_bases_get_window", C.c_int, [_ctx, _bases, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("fb_bases_free", None, [_ctx, _bases]),
    ("fb_msm", C.c_int, [_ctx, _bases, _u8p, C.c_uint64, _u8p]),
    ("fb_msm_partial", C.c_int, [_ctx, _bases, _u8p, C.c_uint64, _u8p]),
    ("fb_g1_sum", C.c_int, [_ctx, _u8p, C.c_uint64, C.c_int, _u8p]),
    ("fb_ntt", C.c_int, [_ctx, C.c_int, _u8p, C.c_uint64, _u8p]),
    ("fb_dev_fr_from_be", C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_uint64]),
    ("fb_dev_fr_to_be", C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_uint64]),
    ("fb_dev_ntt", C.c_int, [_ctx, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]),
    ("fb_dev_msm", C.c_int, [_ctx, _bases, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_msm_partial", C.c_int, [_ctx, _bases, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_commit_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_open_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fb_dev_worker_commit", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_worker_open", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    ("fb_ctx_set_col_shard", C.c_int, [_ctx, C.c_uint64, C.c_uint64]),
    ("fb_worker_commit_partial", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p]),
    ("fb_worker_open_partial", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p]),
    ("fb_dev_worker_commit_partial", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_worker_open_partial", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    ("fb_worker_commit_open", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p, _u8p]),
    ("fb_worker_commit_open_partial", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p, _u8p]),
    ("fb_dev_worker_commit_open", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    ("fb_dev_worker_commit_open_partial", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    ("fb_worker_commit_open_shard_begin", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, C.c_uint64, _u8p, _u8p]),
    ("fb_worker_commit_open_shard_finish", C.c_int, [_ctx, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p, _u8p]),
    ("fb_dev_commit_open_shard_begin", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    ("fb_dev_commit_open_shard_finish", C.c_int, [_ctx, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_dev_combine_commit_open", C.c_int, [_ctx, C.c_void_p, C.c_uint64, C.c_void_p]),
    ("fb_commit_open_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, _u8p, C.c_uint64, _u8p, _u8p, _u8p, _u8p]),
    ("fb_dev_commit_open_rows", C.c_int, [_ctx, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fb_dev_evaluate_noalloc", C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fb_launch_count", C.c_uint64, [_ctx]),
    ("fb_prof_enable", C.c_int, [_ctx, C.c_int]),
    ("fb_prof_read", C.c_int, [_ctx, C.c_char_p, C.c_size_t]),
    ("fb_msm_stats", C.c_int, [_ctx, C.POINTER(C.c_uint64)]),
]

_lib = None


def load() -> C.CDLL:
    """Load libfourier_b200.so and type every exported symbol.  Raises if the
    library has not been built (python __graft_entry__.py build / make)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build the CUDA extension first (make, or "
            "python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the header and the library drifted apart
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def addr(buf) -> int:
    """Address of a bytes-like / numpy / torch buffer (host or device)."""
    if buf is None:
        return 0
    if isinstance(buf, int):
        return buf
    if hasattr(buf, "data_ptr"):  # torch tensor
        return buf.data_ptr()
    if hasattr(buf, "ctypes"):  # numpy array
        return buf.ctypes.data
    if isinstance(buf, bytes):
        return C.cast(C.c_char_p(buf), C.c_void_p).value
    if isinstance(buf, (bytearray, memoryview)):
        return C.addressof((C.c_char * len(buf)).from_buffer(buf))
    raise TypeError(f"cannot take the address of {type(buf)}")
