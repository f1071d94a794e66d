"""fb_dev_evaluate_noalloc: the allocation-free evaluation entry point (static __device__ scratch; no counterpart in the
reference) against the existing entry point fb_evaluate, through the C ABI.
Exact integer arithmetic, so the tolerance is the one every parity test here uses: identical bytes.  Collected last
(test_zz_*)."""
import ctypes as C
import random

import pytest

from util import R, fr_list_to_bytes

pytestmark = pytest.mark.gpu


def test_noalloc_evaluate_matches_fb_evaluate_on_consecutive_calls(gpu):
    import torch
    from fourier_b200 import PianoBackend, _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    b = PianoBackend(4, 0)  # the existing entry points (fb_evaluate, the codecs) need a context; the new one does not
    st = torch.cuda.Stream(device=dev)
    b.set_stream(st.cuda_stream)  # codecs and the new call are ordered on one stream
    rng = random.Random(2024)
    x = rng.randrange(R)
    xb = x.to_bytes(32, "big")
    # a long call first, then shorter ones: the shared scratch keeps the long call's contents and must not leak;
    # 1 << 20 is the stated maximum, 33 / 32 / 1 straddle the chunk size of the blocked scan
    sizes = [1 << 20, 1000, 1 << 20, 33, 32, 1, 70001, 70001]
    with torch.cuda.stream(st):
        d_x_be = torch.frombuffer(bytearray(xb), dtype=torch.uint8).to(dev)
        d_x = torch.empty(32, dtype=torch.uint8, device=dev)
        b.dev_fr_from_be(d_x_be, d_x, 1)
        for n in sizes:
            if n >= 1 << 16:
                from fourier_b200 import synth
                poly = synth.fr_bytes(7 + n % 5, n).tobytes()
            else:
                poly = fr_list_to_bytes([rng.randrange(R) for _ in range(n)])
            want = b.evaluate(poly, xb)  # existing entry point (host buffers, allocating scratch)
            d_be = torch.frombuffer(bytearray(poly), dtype=torch.uint8).to(dev)
            d_fr = torch.empty_like(d_be)
            b.dev_fr_from_be(d_be, d_fr, n)
            d_y = torch.zeros(32, dtype=torch.uint8, device=dev)
            d_out = torch.zeros(32, dtype=torch.uint8, device=dev)
            for _ in range(2):  # two calls in a row on the same inputs
                d_y.fill_(0xAB)
                rc = lib.fb_dev_evaluate_noalloc(0, d_fr.data_ptr(), n, d_x.data_ptr(), d_y.data_ptr(), st.cuda_stream)
                assert rc == 0, lib.fb_last_error(None)
                b.dev_fr_to_be(d_y, d_out, 1)
                st.synchronize()
                assert bytes(d_out.cpu().numpy()) == want, n
        # the empty polynomial evaluates to zero
        d_y = torch.full((32,), 0xAB, dtype=torch.uint8, device=dev)
        assert lib.fb_dev_evaluate_noalloc(0, None, 0, d_x.data_ptr(), d_y.data_ptr(), st.cuda_stream) == 0
        st.synchronize()
        assert bytes(d_y.cpu().numpy()) == bytes(32)
        # beyond the stated maximum: a clear error, nothing launched
        l0 = lib.fb_launch_count(None)
        rc = lib.fb_dev_evaluate_noalloc(0, d_fr.data_ptr(), (1 << 20) + 1, d_x.data_ptr(), d_y.data_ptr(), st.cuda_stream)
        assert rc == 2 and b"2^20" in lib.fb_last_error(None)
        assert lib.fb_dev_evaluate_noalloc(0, d_fr.data_ptr(), 8, None, d_y.data_ptr(), st.cuda_stream) == 2
        assert lib.fb_dev_evaluate_noalloc(99, d_fr.data_ptr(), 8, d_x.data_ptr(), d_y.data_ptr(), st.cuda_stream) == 2
        assert lib.fb_launch_count(None) == l0
    b.close()
