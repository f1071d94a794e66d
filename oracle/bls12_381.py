"""BLS12-381 big-int arithmetic: the bottom layer of the CPU ORACLE.

TEST INFRASTRUCTURE ONLY.  Nothing under ``fourier_b200/`` may import this
module; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker.

What it restates.  The reference (apollozkp/fourier) holds no arithmetic of
its own: every field / curve / pairing operation is a call into three crates
that are NOT vendored under /root/reference:

  * ``kzg`` and ``rust-kzg-blst`` from github.com/apollozkp/rust-kzg, branch
    ``piano``, pinned at abb71933dd00adeec6654e263c1487c154cd7f9a
    (reference Cargo.toml:11-12, Cargo.lock:767-769, 1247-1249)
  * ``blst`` 0.3.11 (reference Cargo.toml:29, Cargo.lock:139-142)

so this file restates the *published* BLS12-381 definitions those crates
implement (curve y^2 = x^3 + 4 over Fp, the ZCash serialisation, the optimal
ate pairing) with Python integers.  Every output of the path is a canonical
encoding of a uniquely defined field/group element, hence any correct
implementation is bit-identical.

Pinning (see tests/test_oracle_kat.py):
  * p and r are re-derived from the BLS parameter z and checked,
  * the generator is on the curve, has order r, and compresses to the
    well-known ``97f1d3a7...c6bb`` string,
  * Fr byte order / modulus are pinned by the only numeric known-answer
    vector in the reference (src/engine/blst.rs:304-324),
  * the pairing is checked for bilinearity.
Compressed G1 *outputs* have no golden vector anywhere in the reference
 - "parity unpinned" for those beyond the checks above.
"""

# ----------------------------------------------------------------------------
# parameters
# ----------------------------------------------------------------------------
Z_PARAM = -0xD201000000010000  # BLS parameter (negative)
P = 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB
R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
B_COEFF = 4

G1_X = 0x17F1D3A73197D7942695638C4FA9AC0FC3688C4F9774B905A14E3A3F171BAC586C55E83FF97A1AEFFB3AF00ADB22C6BB
G1_Y = 0x08B3F481E3AAA0F1A09E30ED741D8AE4FCF5E095D5D00AF600DB18CB2C04B3EDD03CC744A2888AE40CAA232946C5E7E1

G2_X = (
    0x024AA2B2F08F0A91260805272DC51051C6E47AD4FA403B02B4510B647AE3D1770BAC0326A805BBEFD48056C8C121BDB8,
    0x13E02B6052719F607DACD3A088274F65596BD0D09920B61AB5DA61BBDC7F5049334CF11213945D57E5AC7D055D042B7E,
)
G2_Y = (
    0x0CE5D527727D6E118CC9CDC6DA2E351AADFD9BAA8CBDD3A76D429A695160D12C923AC9CC3BACA289E193548608B82801,
    0x0606C4A02EA734CC32ACD2B02BC28B99CB3E287E85A763AF267492AB572E99AB3F370D275CEC1DA1AAA9075FF05F79BE,
)

# Fr multiplicative generator used by c-kzg / rust-kzg for the roots of unity:
# omega_{2^k} = 7^((r-1)/2^k)   [recalled; SURVEY.md §8 a3 checked the scale-2
# constant 0x8d51ccce...0000 numerically]
FR_GENERATOR = 7
FR_TWO_ADICITY = 32


def root_of_unity(log_n: int) -> int:
    """Primitive 2^log_n-th root of unity in Fr (rust-kzg SCALE2_ROOT_OF_UNITY[log_n])."""
    assert 0 <= log_n <= FR_TWO_ADICITY
    return pow(FR_GENERATOR, (R - 1) >> log_n, R)


# ----------------------------------------------------------------------------
# Fr / Fp helpers
# ----------------------------------------------------------------------------
def fr_inv(a: int) -> int:
    return pow(a % R, R - 2, R)


def fp_inv(a: int) -> int:
    return pow(a % P, P - 2, P)


def fr_from_bytes(b: bytes) -> int:
    """FsFr::from_bytes: 32-byte big-endian, canonical (< r) else error
    (used at reference piano.rs:62, 70; BE pinned by blst.rs:304-324)."""
    if len(b) != 32:
        raise ValueError("Invalid byte length. Expected 32, got %d" % len(b))
    v = int.from_bytes(b, "big")
    if v >= R:
        raise ValueError("Invalid scalar")
    return v


def fr_to_bytes(v: int) -> bytes:
    return (v % R).to_bytes(32, "big")


def hash_to_bls_field(secret: bytes) -> int:
    """kzg::eip_4844::hash_to_bls_field as used at piano.rs:890-891: the 32
    bytes read as a big-endian integer, reduced mod r [recalled]."""
    assert len(secret) == 32
    return int.from_bytes(secret, "big") % R


# ----------------------------------------------------------------------------
# G1: Jacobian (X, Y, Z) tuples of ints, Z == 0 is the identity
# ----------------------------------------------------------------------------
G1_INF = (0, 1, 0)
G1_GEN = (G1_X, G1_Y, 1)


def g1_is_inf(pt):
    return pt[2] == 0


def g1_double(pt):
    X, Y, Z = pt
    if Z == 0 or Y == 0:
        return G1_INF
    A = X * X % P
    Bq = Y * Y % P
    C = Bq * Bq % P
    D = 2 * ((X + Bq) * (X + Bq) - A - C) % P
    E = 3 * A % P
    F = E * E % P
    X3 = (F - 2 * D) % P
    Y3 = (E * (D - X3) - 8 * C) % P
    Z3 = 2 * Y * Z % P
    return (X3, Y3, Z3)


def g1_add(p1, p2):
    X1, Y1, Z1 = p1
    X2, Y2, Z2 = p2
    if Z1 == 0:
        return p2
    if Z2 == 0:
        return p1
    Z1Z1 = Z1 * Z1 % P
    Z2Z2 = Z2 * Z2 % P
    U1 = X1 * Z2Z2 % P
    U2 = X2 * Z1Z1 % P
    S1 = Y1 * Z2 * Z2Z2 % P
    S2 = Y2 * Z1 * Z1Z1 % P
    if U1 == U2:
        if S1 == S2:
            return g1_double(p1)
        return G1_INF
    H = (U2 - U1) % P
    Rr = (S2 - S1) % P
    HH = H * H % P
    HHH = H * HH % P
    V = U1 * HH % P
    X3 = (Rr * Rr - HHH - 2 * V) % P
    Y3 = (Rr * (V - X3) - S1 * HHH) % P
    Z3 = Z1 * Z2 * H % P
    return (X3, Y3, Z3)


def g1_neg(pt):
    return (pt[0], (-pt[1]) % P, pt[2])


def g1_mul(pt, k: int):
    k %= R
    acc = G1_INF
    for bit in bin(k)[2:] if k else "":
        acc = g1_double(acc)
        if bit == "1":
            acc = g1_add(acc, pt)
    return acc


def g1_to_affine(pt):
    """-> (x, y) or None for the identity."""
    X, Y, Z = pt
    if Z == 0:
        return None
    zi = fp_inv(Z)
    zi2 = zi * zi % P
    return (X * zi2 % P, Y * zi2 * zi % P)


def g1_from_affine(a):
    return G1_INF if a is None else (a[0], a[1], 1)


def g1_eq(p1, p2):
    return g1_to_affine(p1) == g1_to_affine(p2)


def g1_on_curve_affine(a):
    if a is None:
        return True
    x, y = a
    return (y * y - x * x * x - B_COEFF) % P == 0


def g1_compress(pt) -> bytes:
    """FsG1::to_bytes (blst_p1_compress): ZCash 48-byte form.
    bit7 = compressed, bit6 = infinity, bit5 = y is the lexicographically
    larger root.  Used at reference rpc.rs:241, 268-269, 311, 330."""
    a = g1_to_affine(pt)
    if a is None:
        return bytes([0xC0]) + bytes(47)
    x, y = a
    out = bytearray(x.to_bytes(48, "big"))
    out[0] |= 0x80
    if y > (P - 1) // 2:
        out[0] |= 0x20
    return bytes(out)


def g1_decompress(b: bytes):
    """FsG1::from_bytes (blst_p1_uncompress): on-curve check, no subgroup check
    [recalled].  Used at reference piano.rs:57, 724, 759."""
    if len(b) != 48:
        raise ValueError("Invalid byte length. Expected 48, got %d" % len(b))
    flags = b[0]
    if not flags & 0x80:
        raise ValueError("Failed to uncompress: compression bit not set")
    x = int.from_bytes(b, "big") & ((1 << 381) - 1)
    if flags & 0x40:
        if x != 0 or flags & 0x20:
            raise ValueError("Failed to uncompress: bad infinity encoding")
        return G1_INF
    if x >= P:
        raise ValueError("Failed to uncompress: x >= p")
    rhs = (x * x * x + B_COEFF) % P
    y = pow(rhs, (P + 1) // 4, P)
    if y * y % P != rhs:
        raise ValueError("Failed to uncompress: not on curve")
    if (y > (P - 1) // 2) != bool(flags & 0x20):
        y = P - y
    return (x, y, 1)


def g1_serialize(pt) -> bytes:
    """FsG1::serialize (blst_p1_serialize): 96-byte x||y, bit6 = infinity.
    Used at reference piano.rs:655."""
    a = g1_to_affine(pt)
    if a is None:
        return bytes([0x40]) + bytes(95)
    return a[0].to_bytes(48, "big") + a[1].to_bytes(48, "big")


def g1_deserialize(b: bytes):
    """FsG1::deserialize (blst_p1_deserialize), reference piano.rs:728, 770."""
    if len(b) != 96:
        raise ValueError("Invalid byte length. Expected 96, got %d" % len(b))
    if b[0] & 0x80:
        raise ValueError("compressed flag set on uncompressed encoding")
    if b[0] & 0x40:
        if any(b[1:]) or b[0] & 0x3F:
            raise ValueError("bad infinity encoding")
        return G1_INF
    x = int.from_bytes(b[:48], "big")
    y = int.from_bytes(b[48:], "big")
    if x >= P or y >= P or not g1_on_curve_affine((x, y)):
        raise ValueError("not on curve")
    return (x, y, 1)


def g1_msm_naive(points, scalars):
    acc = G1_INF
    for pt, k in zip(points, scalars):
        acc = g1_add(acc, g1_mul(pt, k))
    return acc


def g1_msm(points, scalars, c: int = 8):
    """Windowed (Pippenger) MSM: same group element as the naive sum, used so
    the oracle finishes 2^12-point cases in seconds."""
    n = len(points)
    assert n == len(scalars)
    if n < 16:
        return g1_msm_naive(points, scalars)
    nwin = (255 + c - 1) // c
    total = G1_INF
    for w in reversed(range(nwin)):
        for _ in range(c):
            total = g1_double(total)
        buckets = [G1_INF] * (1 << c)
        for pt, k in zip(points, scalars):
            d = (k >> (w * c)) & ((1 << c) - 1)
            if d:
                buckets[d] = g1_add(buckets[d], pt)
        run = G1_INF
        acc = G1_INF
        for d in range((1 << c) - 1, 0, -1):
            run = g1_add(run, buckets[d])
            acc = g1_add(acc, run)
        total = g1_add(total, acc)
    return total


# ----------------------------------------------------------------------------
# Fp2 = Fp[u]/(u^2+1), elements are (c0, c1)
# ----------------------------------------------------------------------------
def f2_add(a, b):
    return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)


def f2_sub(a, b):
    return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)


def f2_neg(a):
    return ((-a[0]) % P, (-a[1]) % P)


def f2_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def f2_sqr(a):
    return f2_mul(a, a)


def f2_muli(a, k: int):
    return (a[0] * k % P, a[1] * k % P)


def f2_inv(a):
    d = fp_inv(a[0] * a[0] + a[1] * a[1])
    return (a[0] * d % P, (-a[1]) * d % P)


F2_ZERO = (0, 0)
F2_ONE = (1, 0)
B2_COEFF = (4, 4)  # twist: y^2 = x^3 + 4(1+u)


def f2_sqrt(a):
    """Square root in Fp2 (p = 3 mod 4), or None."""
    if a == F2_ZERO:
        return F2_ZERO
    # a^((p^2+7)/16)-style shortcuts are fiddly; use the norm method.
    c0, c1 = a
    if c1 == 0:
        s = pow(c0, (P + 1) // 4, P)
        if s * s % P == c0:
            return (s, 0)
        s = pow((-c0) % P, (P + 1) // 4, P)
        assert s * s % P == (-c0) % P
        return (0, s)
    norm = (c0 * c0 + c1 * c1) % P
    alpha = pow(norm, (P + 1) // 4, P)
    if alpha * alpha % P != norm:
        return None
    inv2 = fp_inv(2)
    delta = (c0 + alpha) * inv2 % P
    x0 = pow(delta, (P + 1) // 4, P)
    if x0 * x0 % P != delta:
        delta = (c0 - alpha) * inv2 % P
        x0 = pow(delta, (P + 1) // 4, P)
        if x0 * x0 % P != delta:
            return None
    x1 = c1 * fp_inv(2 * x0) % P
    res = (x0, x1)
    return res if f2_sqr(res) == (c0 % P, c1 % P) else None


# ----------------------------------------------------------------------------
# G2 on the twist, affine tuples (x, y) over Fp2 or None (host-side, tiny)
# ----------------------------------------------------------------------------
G2_GEN = (G2_X, G2_Y)


def g2_on_curve(q):
    if q is None:
        return True
    x, y = q
    return f2_sub(f2_sqr(y), f2_add(f2_mul(f2_sqr(x), x), B2_COEFF)) == F2_ZERO


def g2_neg(q):
    return None if q is None else (q[0], f2_neg(q[1]))


def g2_add(q1, q2):
    if q1 is None:
        return q2
    if q2 is None:
        return q1
    x1, y1 = q1
    x2, y2 = q2
    if x1 == x2:
        if y1 == y2 and y1 != F2_ZERO:
            lam = f2_mul(f2_muli(f2_sqr(x1), 3), f2_inv(f2_muli(y1, 2)))
        else:
            return None
    else:
        lam = f2_mul(f2_sub(y2, y1), f2_inv(f2_sub(x2, x1)))
    x3 = f2_sub(f2_sub(f2_sqr(lam), x1), x2)
    y3 = f2_sub(f2_mul(lam, f2_sub(x1, x3)), y1)
    return (x3, y3)


def g2_mul(q, k: int):
    k %= R
    acc = None
    for bit in bin(k)[2:] if k else "":
        acc = g2_add(acc, acc)
        if bit == "1":
            acc = g2_add(acc, q)
    return acc


def g2_sub(q1, q2):
    return g2_add(q1, g2_neg(q2))


def _f2_lex_larger(y):
    """ZCash sign rule for Fp2: compare c1 first, then c0."""
    if y[1] != 0:
        return y[1] > (P - 1) // 2
    return y[0] > (P - 1) // 2


def g2_compress(q) -> bytes:
    """FsG2::to_bytes (blst_p2_compress): 96 bytes, x.c1 || x.c0, flags in byte 0
    (reference piano.rs:662)."""
    if q is None:
        return bytes([0xC0]) + bytes(95)
    x, y = q
    out = bytearray(x[1].to_bytes(48, "big") + x[0].to_bytes(48, "big"))
    out[0] |= 0x80
    if _f2_lex_larger(y):
        out[0] |= 0x20
    return bytes(out)


def g2_decompress(b: bytes):
    if len(b) != 96:
        raise ValueError("Invalid byte length. Expected 96, got %d" % len(b))
    flags = b[0]
    if not flags & 0x80:
        raise ValueError("compression bit not set")
    c1 = int.from_bytes(b[:48], "big") & ((1 << 381) - 1)
    c0 = int.from_bytes(b[48:], "big")
    if flags & 0x40:
        if c0 or c1 or flags & 0x20:
            raise ValueError("bad infinity encoding")
        return None
    if c0 >= P or c1 >= P:
        raise ValueError("x >= p")
    x = (c0, c1)
    y = f2_sqrt(f2_add(f2_mul(f2_sqr(x), x), B2_COEFF))
    if y is None:
        raise ValueError("not on curve")
    if _f2_lex_larger(y) != bool(flags & 0x20):
        y = f2_neg(y)
    return (x, y)


def g2_serialize(q) -> bytes:
    """FsG2::serialize (blst_p2_serialize): 192 bytes x.c1||x.c0||y.c1||y.c0
    (reference piano.rs:665)."""
    if q is None:
        return bytes([0x40]) + bytes(191)
    x, y = q
    return b"".join(v.to_bytes(48, "big") for v in (x[1], x[0], y[1], y[0]))


def g2_deserialize(b: bytes):
    if len(b) != 192:
        raise ValueError("Invalid byte length. Expected 192, got %d" % len(b))
    if b[0] & 0x40:
        return None
    x = (int.from_bytes(b[48:96], "big"), int.from_bytes(b[:48], "big"))
    y = (int.from_bytes(b[144:192], "big"), int.from_bytes(b[96:144], "big"))
    if not g2_on_curve((x, y)):
        raise ValueError("not on curve")
    return (x, y)


# ----------------------------------------------------------------------------
# Fp12 as Fp[w]/(w^12 - 2 w^6 + 2)  (w^6 = 1 + u), lists of 12 ints.
# Slow but transparent; only used for the verify equations
# (reference piano.rs:358-388, 422-464: Pairing::raw_aggregate, final_exp,
# blst_fp12_is_one).
# ----------------------------------------------------------------------------
F12_ONE = [1] + [0] * 11


def f12_mul(a, b):
    t = [0] * 23
    for i, ai in enumerate(a):
        if ai:
            for j, bj in enumerate(b):
                if bj:
                    t[i + j] += ai * bj
    for k in range(22, 11, -1):
        c = t[k]
        if c:
            t[k - 6] += 2 * c
            t[k - 12] -= 2 * c
    return [v % P for v in t[:12]]


def f12_pow(a, e: int):
    res = F12_ONE
    for bit in bin(e)[2:]:
        res = f12_mul(res, res)
        if bit == "1":
            res = f12_mul(res, a)
    return res


def _f2_at(c, k):
    """Fp2 element c placed at w^k: (c0 + c1 u) w^k with u = w^6 - 1."""
    out = [0] * 12
    out[k] = (c[0] - c[1]) % P
    out[k + 6] = c[1] % P
    return out


def _line(lam, xt, yt, xp, yp):
    """Line through T (twist coords) with twist-slope lam, evaluated at the G1
    point (xp, yp), scaled by w^3 (a factor the final exponentiation kills):
        yp w^3 - lam xp w^2 + (lam xT - yT)."""
    c0 = f2_sub(f2_mul(lam, xt), yt)
    c2 = f2_muli(f2_neg(lam), xp)
    out = _f2_at(c0, 0)
    t2 = _f2_at(c2, 2)
    for k in range(12):
        out[k] = (out[k] + t2[k]) % P
    out[3] = (out[3] + yp) % P
    return out


def miller_loop(p_aff, q_aff):
    """f_{|z|,Q}(P) for P in G1 (affine ints), Q on the twist (affine Fp2).
    The sign of z is ignored (a common inversion of every factor leaves
    'product == 1' unchanged)."""
    if p_aff is None or q_aff is None:
        return F12_ONE
    xp, yp = p_aff
    f = F12_ONE
    T = q_aff
    for bit in bin(-Z_PARAM)[3:]:
        xt, yt = T
        lam = f2_mul(f2_muli(f2_sqr(xt), 3), f2_inv(f2_muli(yt, 2)))
        f = f12_mul(f12_mul(f, f), _line(lam, xt, yt, xp, yp))
        T = g2_add(T, T)
        if bit == "1":
            xt, yt = T
            lam = f2_mul(f2_sub(q_aff[1], yt), f2_inv(f2_sub(q_aff[0], xt)))
            f = f12_mul(f, _line(lam, xt, yt, xp, yp))
            T = g2_add(T, q_aff)
    return f


FINAL_EXP = (P**12 - 1) // R


def pairing_product_is_one(pairs) -> bool:
    """prod e(P_i, Q_i) == 1 for [(P_i jacobian/affine-None G1, Q_i affine G2)]."""
    f = F12_ONE
    for p1, q2 in pairs:
        pa = g1_to_affine(p1) if (p1 is not None and len(p1) == 3) else p1
        f = f12_mul(f, miller_loop(pa, q2))
    return f12_pow(f, FINAL_EXP) == F12_ONE
