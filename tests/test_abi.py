"""CPU tests of the boundary: the shared library loads, exports every symbol the
header declares, and fails loudly (no fallback) when there is no CUDA device."""
import os
import re

import pytest

from util import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "fourier_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fb_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    from fourier_b200 import _lib
    names = sorted(n for n, _, _ in _lib.SYMBOLS)
    assert names == header_symbols()


def test_library_exports_every_symbol():
    from fourier_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        pytest.fail("libfourier_b200.so is not built: run `make` (or __graft_entry__.build())")
    lib = _lib.load()
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.fb_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from fourier_b200 import FourierError, PianoBackend
    with pytest.raises(FourierError):
        PianoBackend(4, 0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "fourier_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("no oracle", ""), os.path.join(dirpath, f)


def test_synth_is_deterministic_and_reduced():
    from fourier_b200 import synth
    a = synth.fr_ints(7, 3000)
    assert all(0 <= x < synth.R for x in a)
    assert synth.fr_ints(7, 10, start=5) == a[5:15]
    assert synth.fr_bytes(7, 4).tobytes() == b"".join(x.to_bytes(32, "big") for x in a[:4])
    # re-derive two entries with plain Python integers
    M = (1 << 64) - 1

    def sm(x):
        x = (x + 0x9E3779B97F4A7C15) & M
        z = x
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    for idx in (0, 1234):
        base = ((7 & 0xFFFFFF) << 40) + 4 * idx
        v = sum(sm((base + k) & M) << (64 * k) for k in range(4)) & ((1 << 255) - 1)
        if v >= synth.R:
            v -= synth.R
        assert a[idx] == v
