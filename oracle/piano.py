"""CPU ORACLE for fourier's Pianist/KZG engine (reference src/engine/piano.rs).

TEST INFRASTRUCTURE ONLY - see the header of oracle/bls12_381.py.  Each
function cites the reference lines it restates.  Field elements are Python
ints in [0, r); G1 points are Jacobian int triples; wire encodings are the
reference's (Fr 32-byte BE, G1 48-byte compressed).

Parity status: the protocol math below is a line-by-line restatement of
piano.rs, checked against the algebraic properties the reference's own tests
pin (tests/test_oracle_*.py).  The reference ships no golden G1 byte string,
so compressed outputs are pinned only through the known-tau identities
(SURVEY.md §8c): "parity unpinned" against blst bytes themselves.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

from . import bls12_381 as bls
from .bls12_381 import R


# ----------------------------------------------------------------------------
# FsFFTSettings::fft_fr  (called through PianoFFTSettings::fft_left/right,
# piano.rs:1096-1103).  [recalled] rust-kzg: natural order in/out, roots
# omega^i with omega = 7^((r-1)/2^scale); shorter power-of-two inputs use
# stride max_width/len; inverse uses the reversed roots then scales by 1/len.
# ----------------------------------------------------------------------------
def _ntt_inplace(a: List[int], omega: int) -> None:
    n = len(a)
    j = 0
    for i in range(1, n):
        bit = n >> 1
        while j & bit:
            j ^= bit
            bit >>= 1
        j |= bit
        if i < j:
            a[i], a[j] = a[j], a[i]
    length = 2
    while length <= n:
        w_len = pow(omega, n // length, R)
        half = length >> 1
        tw = [1] * half
        for k in range(1, half):
            tw[k] = tw[k - 1] * w_len % R
        for start in range(0, n, length):
            for k in range(half):
                u = a[start + k]
                v = a[start + k + half] * tw[k] % R
                a[start + k] = (u + v) % R
                a[start + k + half] = (u - v) % R
        length <<= 1


def fft_fr(data: Sequence[int], inverse: bool, max_scale: int) -> List[int]:
    n = len(data)
    max_width = 1 << max_scale
    if n > max_width:
        raise ValueError("Supplied list is longer than the available max width")
    if n == 0 or n & (n - 1):
        raise ValueError("A list with power-of-two length is expected")
    log_n = n.bit_length() - 1
    omega = bls.root_of_unity(log_n)  # == root_of_unity(max_scale)^(max_width/n)
    out = [x % R for x in data]
    if inverse:
        _ntt_inplace(out, bls.fr_inv(omega))
        inv_n = bls.fr_inv(n)
        out = [x * inv_n % R for x in out]
    else:
        _ntt_inplace(out, omega)
    return out


def fft_fr_slow(data: Sequence[int], inverse: bool) -> List[int]:
    """O(n^2) definition, used to pin fft_fr on small sizes."""
    n = len(data)
    omega = bls.root_of_unity(n.bit_length() - 1)
    if inverse:
        omega = bls.fr_inv(omega)
    out = [sum(data[j] * pow(omega, j * k, R) for j in range(n)) % R for k in range(n)]
    if inverse:
        inv_n = bls.fr_inv(n)
        out = [x * inv_n % R for x in out]
    return out


# ----------------------------------------------------------------------------
# FsPoly::eval / FsPoly::div by a monic linear factor (piano.rs:202, 215, 271,
# 282) [recalled: Horner; long_div with a len-2 synthetic-division path].
# ----------------------------------------------------------------------------
def poly_eval(coeffs: Sequence[int], x: int) -> int:
    if len(coeffs) == 0:
        return 0
    if x == 0:
        return coeffs[0] % R
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * x + c) % R
    return acc


def poly_div_linear(numer: Sequence[int], alpha: int) -> List[int]:
    """numer / (X - alpha): quotient of length len-1, remainder dropped."""
    n = len(numer)
    if n < 2:
        return []
    q = [0] * (n - 1)
    q[n - 2] = numer[n - 1] % R
    for k in range(n - 2, 0, -1):
        q[k - 1] = (numer[k] + alpha * q[k]) % R
    return q


def bipoly_eval(parts: Sequence[Sequence[int]], x: int, y: int) -> int:
    """BivariateFsPolynomial::eval (bipoly.rs:42-55): f = sum_i x^i parts[i](y)."""
    maxlen = max(len(p) for p in parts)
    res = [0] * maxlen
    pw = 1
    for f in parts:
        for j, c in enumerate(f):
            res[j] = (res[j] + c * pw) % R
        pw = pw * x % R
    return poly_eval(res, y)


# ----------------------------------------------------------------------------
# settings
# ----------------------------------------------------------------------------
@dataclass
class PianoSettings:
    """piano.rs:587-599."""
    g: tuple
    g_tau_x: List[tuple]
    g_tau_y: List[tuple]
    u: List[List[tuple]]
    g2: tuple
    g2_tau_x: tuple
    g2_tau_y: tuple
    # kept only by the oracle (the reference forgets them) so the tests can
    # use the known-tau identities of SURVEY.md §8c
    tau_x: Optional[int] = None
    tau_y: Optional[int] = None
    r_at_tau: List[int] = field(default_factory=list)
    l_at_tau: List[int] = field(default_factory=list)


def lagrange_at(tau: int, log_n: int) -> List[int]:
    """[L_j(tau)]_j for the size-2^log_n domain, closed form
    L_j(X) = omega^j (X^n - 1) / (n (X - omega^j))  (piano.rs:1117-1135 doc
    comments).  Same field elements as the reference's iFFT(e_j) + Horner
    (piano.rs:956-988), without the O(n^2 log n) cost."""
    n = 1 << log_n
    omega = bls.root_of_unity(log_n)
    roots = [1] * n
    for j in range(1, n):
        roots[j] = roots[j - 1] * omega % R
    zn = (pow(tau, n, R) - 1) % R
    if zn == 0:  # tau is in the domain
        return [1 if roots[j] == tau % R else 0 for j in range(n)]
    inv_n = bls.fr_inv(n)
    # batch inversion of (tau - omega^j)
    d = [(tau - w) % R for w in roots]
    pref = [1] * (n + 1)
    for j in range(n):
        pref[j + 1] = pref[j] * d[j] % R
    inv_all = bls.fr_inv(pref[n])
    out = [0] * n
    for j in range(n - 1, -1, -1):
        dinv = inv_all * pref[j] % R
        inv_all = inv_all * d[j] % R
        out[j] = roots[j] * zn % R * inv_n % R * dinv % R
    return out


def lagrange_at_literal(tau: int, log_n: int) -> List[int]:
    """The reference's literal algorithm (piano.rs:968, 985, 1120-1135):
    iFFT of the unit vector e_j, then Horner at tau.  Small sizes only."""
    n = 1 << log_n
    out = []
    for j in range(n):
        e = [0] * n
        e[j] = 1
        out.append(poly_eval(fft_fr(e, True, log_n), tau))
    return out


class PianoOracle:
    """PianoBackend (piano.rs:26-465) restated."""

    def __init__(self, n: int, m: int):
        # PianoFFTSettings::new, piano.rs:1061-1077
        if m > n:
            raise ValueError("m must be less than or equal to n")
        self.n, self.m, self.t = n, m, n - m
        self.T = 1 << self.t
        self.M = 1 << self.m
        self.settings: Optional[PianoSettings] = None

    # -- FFT settings -------------------------------------------------------
    def fft_left(self, data, inverse):  # piano.rs:1096-1098
        return fft_fr(data, inverse, self.t)

    def fft_right(self, data, inverse):  # piano.rs:1101-1103
        return fft_fr(data, inverse, self.m)

    def left_lagrange_poly(self, j):  # piano.rs:1120-1125
        e = [0] * self.T
        e[j] = 1
        return self.fft_left(e, True)

    def right_lagrange_poly(self, i):  # piano.rs:1130-1135
        e = [0] * self.M
        e[i] = 1
        return self.fft_right(e, True)

    # -- setup ----------------------------------------------------------------
    def generate_trusted_setup(self, secret_x: bytes, secret_y: bytes,
                               with_points: bool = True, literal: bool = False):
        """generate_trusted_setup, piano.rs:883-1037.
        with_points=False keeps only the scalars (tau, R_i, L_j): enough for the
        known-tau checks at sizes where 2^20 Python scalar muls are too slow."""
        tau_x = bls.hash_to_bls_field(secret_x)  # :890
        tau_y = bls.hash_to_bls_field(secret_y)  # :891
        lag = lagrange_at_literal if literal else lagrange_at
        r_vals = lag(tau_y, self.m)  # :956-971
        l_vals = lag(tau_x, self.t)  # :973-988
        g = bls.G1_GEN
        if with_points:
            g_tau_x, pw = [], 1
            for _ in range(self.T):  # :916-931
                g_tau_x.append(bls.g1_mul(g, pw))
                pw = pw * tau_x % R
            g_tau_y, pw = [], 1
            for _ in range(self.M):  # :934-949
                g_tau_y.append(bls.g1_mul(g, pw))
                pw = pw * tau_y % R
            u = [[bls.g1_mul(g, ri * lj % R) for lj in l_vals] for ri in r_vals]  # :995-1007
        else:
            g_tau_x, g_tau_y, u = [], [], []
        self.settings = PianoSettings(
            g=g, g_tau_x=g_tau_x, g_tau_y=g_tau_y, u=u,
            g2=bls.G2_GEN,
            g2_tau_x=bls.g2_mul(bls.G2_GEN, tau_x),  # :1019
            g2_tau_y=bls.g2_mul(bls.G2_GEN, tau_y),  # :1021
            tau_x=tau_x, tau_y=tau_y, r_at_tau=r_vals, l_at_tau=l_vals,
        )
        return self.settings

    # -- commit / open ------------------------------------------------------
    def worker_commit(self, i: int, coeffs: Sequence[int]):
        """piano.rs:143-161: C_i = sum_j coeffs[j] * U[i][j]."""
        s = self.settings
        return bls.g1_msm([s.u[i][j] for j in range(len(coeffs))], list(coeffs))

    def master_commit(self, commitments):
        """piano.rs:164-175: sum of the commitments (MSM with all-ones)."""
        acc = bls.G1_INF
        for c in commitments:
            acc = bls.g1_add(acc, c)
        return acc

    def open_scalars(self, coeffs: Sequence[int], alpha: int) -> Tuple[int, List[int]]:
        """The field half of worker_open, piano.rs:193-233: returns
        (y, NTT(q || 0)) - the scalars the final MSM consumes."""
        poly = self.fft_left(coeffs, True)  # :194
        y = poly_eval(poly, alpha)  # :202
        numer = list(poly)
        numer[0] = (numer[0] - y) % R  # :207-208
        q = poly_div_linear(numer, alpha)  # :212-215
        q_cob = [0] * len(poly)  # :224-225
        q_cob[: len(q)] = q
        return y, self.fft_left(q_cob, False)  # :233

    def worker_open(self, i: int, coeffs: Sequence[int], alpha: int):
        """piano.rs:187-238 -> (y, pi)."""
        y, q_hat = self.open_scalars(coeffs, alpha)
        s = self.settings
        pi = bls.g1_msm([s.u[i][j] for j in range(len(q_hat))], q_hat)  # :228-236
        return y, pi

    def master_open_scalars(self, evals: Sequence[int], beta: int) -> Tuple[int, List[int]]:
        poly = self.fft_right(evals, True)  # :268
        z = poly_eval(poly, beta)  # :271
        numer = list(poly)
        numer[0] = (numer[0] - z) % R  # :275-276
        return z, poly_div_linear(numer, beta)  # :280-282

    def master_open(self, evals, proofs, beta):
        """piano.rs:250-298 -> (z, (pi0, pi1))."""
        pi0 = self.master_commit(proofs)  # :258-265
        z, q = self.master_open_scalars(evals, beta)
        s = self.settings
        pi1 = bls.g1_msm([s.g_tau_y[j] for j in range(len(q))], q)  # :286-295
        return z, (pi0, pi1)

    # -- known-tau shortcuts (oracle only) -----------------
    def worker_commit_known_tau(self, i: int, coeffs: Sequence[int]):
        """G * (R_i(tau_Y) * sum_j c_j L_j(tau_X)): the same group element as
        worker_commit, O(T) field ops + one scalar multiplication."""
        s = self.settings
        acc = 0
        for c, l in zip(coeffs, s.l_at_tau):
            acc += c * l
        return bls.g1_mul(bls.G1_GEN, acc % R * s.r_at_tau[i] % R)

    def worker_open_known_tau(self, i: int, coeffs: Sequence[int], alpha: int):
        y, q_hat = self.open_scalars(coeffs, alpha)
        return y, self.worker_commit_known_tau(i, q_hat)

    # -- verify ---------------------------------------------------------------
    def worker_verify(self, i, commitment, alpha, y, pi) -> bool:
        """piano.rs:309-356: e(C - g^{R_i(tau_Y) y}, g2) == e(pi, g2^{tau_X - alpha})."""
        s = self.settings
        r_poly = self.right_lagrange_poly(i)  # :321
        if s.g_tau_y:
            r_i_tau_y = bls.g1_msm([s.g_tau_y[j] for j in range(len(r_poly))], r_poly)  # :326-335
        else:
            r_i_tau_y = bls.g1_mul(bls.G1_GEN, s.r_at_tau[i])
        g_y_prime = bls.g1_mul(r_i_tau_y, y)  # :338
        lhs = bls.g1_add(commitment, bls.g1_neg(g_y_prime))  # :341
        g2_x_minus_alpha = bls.g2_sub(s.g2_tau_x, bls.g2_mul(s.g2, alpha))  # :344-347
        # pairings_verify_single, :358-388: e(-lhs, g2) * e(pi, g2^{tau_X-alpha}) == 1
        return bls.pairing_product_is_one([(bls.g1_neg(lhs), s.g2), (pi, g2_x_minus_alpha)])

    def master_verify(self, commitment, beta, alpha, z, pi_f) -> bool:
        """piano.rs:393-420 (+ :422-464)."""
        s = self.settings
        lhs = bls.g1_add(commitment, bls.g1_neg(bls.g1_mul(s.g, z)))  # :402
        g2_x = bls.g2_sub(s.g2_tau_x, bls.g2_mul(s.g2, alpha))  # :403-406
        g2_y = bls.g2_sub(s.g2_tau_y, bls.g2_mul(s.g2, beta))  # :407-410
        return bls.pairing_product_is_one(
            [(bls.g1_neg(lhs), s.g2), (pi_f[0], g2_x), (pi_f[1], g2_y)])


# ----------------------------------------------------------------------------
# setup file (piano.rs:649-706 save, :711-846 load).  No header, no flag.
#   G1 g | u64le |g_tau_x| | g_tau_x[] | u64le |g_tau_y| | g_tau_y[] |
#   u64le rows | u64le cols | u[rows][cols] | G2 g2 | G2 g2^tauX | G2 g2^tauY
# ----------------------------------------------------------------------------
def setup_to_bytes(s: PianoSettings, compressed: bool) -> bytes:
    w1 = bls.g1_compress if compressed else bls.g1_serialize
    w2 = bls.g2_compress if compressed else bls.g2_serialize
    out = [w1(s.g)]
    out.append(struct.pack("<Q", len(s.g_tau_x)))
    out += [w1(p) for p in s.g_tau_x]
    out.append(struct.pack("<Q", len(s.g_tau_y)))
    out += [w1(p) for p in s.g_tau_y]
    out.append(struct.pack("<QQ", len(s.u), len(s.u[0])))
    for row in s.u:
        out += [w1(p) for p in row]
    out += [w2(s.g2), w2(s.g2_tau_x), w2(s.g2_tau_y)]
    return b"".join(out)


def setup_from_bytes(b: bytes, compressed: bool) -> PianoSettings:
    n1 = 48 if compressed else 96
    n2 = 96 if compressed else 192
    r1 = bls.g1_decompress if compressed else bls.g1_deserialize
    r2 = bls.g2_decompress if compressed else bls.g2_deserialize
    pos = 0

    def take(k):
        nonlocal pos
        chunk = b[pos:pos + k]
        if len(chunk) != k:
            raise ValueError("failed to fill whole buffer")
        pos += k
        return chunk

    g = r1(take(n1))
    (cx,) = struct.unpack("<Q", take(8))
    gx = [r1(take(n1)) for _ in range(cx)]
    (cy,) = struct.unpack("<Q", take(8))
    gy = [r1(take(n1)) for _ in range(cy)]
    rows, cols = struct.unpack("<QQ", take(16))
    u = [[r1(take(n1)) for _ in range(cols)] for _ in range(rows)]
    g2, g2x, g2y = r2(take(n2)), r2(take(n2)), r2(take(n2))
    return PianoSettings(g=g, g_tau_x=gx, g_tau_y=gy, u=u, g2=g2, g2_tau_x=g2x, g2_tau_y=g2y)


# ----------------------------------------------------------------------------
# deterministic synthetic inputs: SplitMix64 -> Fr
# ----------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def synth_fr(seed: int, index: int) -> int:
    """u256 = 4 x SplitMix64(seed*2^32 + 4*index + k) words, reduced mod r."""
    base = ((seed & 0xFFFFFFFF) << 40) + 4 * index
    v = 0
    for k in range(4):
        v |= splitmix64((base + k) & _M64) << (64 * k)
    return v % R


def synth_fr_vec(seed: int, n: int) -> List[int]:
    return [synth_fr(seed, i) for i in range(n)]
