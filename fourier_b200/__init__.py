"""fourier_b200 - B200-native KZG / Pianist hot path (G1 MSM, Fr NTT, opening
quotient, row-wise bivariate commits) behind fourier's PianoBackend API.

The arithmetic lives in libfourier_b200.so (hand-written sm_100a CUDA behind the C
ABI of include/fourier_b200.h); this package is the thin host mirror of the
reference's engine interface.  There is no CPU fallback.
"""
from .engine import Bases, FourierError, PianoBackend  # noqa: F401

__all__ = ["PianoBackend", "Bases", "FourierError"]
