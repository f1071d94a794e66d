// g1.cuh - BLS12-381 G1 (y^2 = x^3 + 4 over Fp) group law, host+device.
//
// Replaces blst's POINTonE1 arithmetic reached by the reference through
// rust_kzg_blst::kzg_proofs::g1_linear_combination (reference
// src/engine/piano.rs:150-158, 228-236, 258-265, 286-295).
//
// Resident points (SRS rows, window tables) are AFFINE Montgomery pairs, 96 B,
// with (0,0) standing for the identity ((0,0) is not on the curve).  Running sums
// are XYZZ (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2; ZZ == 0 is the identity): a mixed
// add is 8M+2S with no inversion, and the sums are only normalised once per
// output.  Every routine is complete: identity operands, P+P and P+(-P) are all
// handled, because the reference's tests feed identity commitments and
// duplicate points (piano.rs:1406-1411, 1466, 1754, 1769).
#pragma once
#include "field.cuh"

namespace fb {

struct G1Affine {
    Fp x, y;
    FB_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
    FB_HD static G1Affine inf() { return G1Affine{Fp::zero(), Fp::zero()}; }
};

struct G1Xyzz {
    Fp x, y, zz, zzz;
    FB_HD bool is_inf() const { return zz.is_zero(); }
    FB_HD static G1Xyzz inf() { return G1Xyzz{Fp::zero(), Fp::zero(), Fp::zero(), Fp::zero()}; }
    FB_HD static G1Xyzz from_affine(const G1Affine& p) {
        if (p.is_inf()) return inf();
        return G1Xyzz{p.x, p.y, Fp::one(), Fp::one()};
    }
};

FB_HD G1Affine g1_neg(const G1Affine& p) { return G1Affine{p.x, fe_neg<FpParams>(p.y)}; }
FB_HD G1Xyzz g1_neg(const G1Xyzz& p) { return G1Xyzz{p.x, fe_neg<FpParams>(p.y), p.zz, p.zzz}; }

// 2*P for an affine P (not the identity)
FB_HD G1Xyzz g1_dbl_affine(const G1Affine& p) {
    using F = FpParams;
    Fp u = fe_dbl<F>(p.y);
    Fp v = fe_sqr<F>(u);
    Fp w = fe_mul<F>(u, v);
    Fp s = fe_mul<F>(p.x, v);
    Fp xx = fe_sqr<F>(p.x);
    Fp m = fe_add<F>(fe_dbl<F>(xx), xx);
    G1Xyzz r;
    r.x = fe_sub<F>(fe_sqr<F>(m), fe_dbl<F>(s));
    r.y = fe_sub<F>(fe_mul<F>(m, fe_sub<F>(s, r.x)), fe_mul<F>(w, p.y));
    r.zz = v;
    r.zzz = w;
    return r;
}

FB_HD G1Xyzz g1_dbl(const G1Xyzz& p) {
    using F = FpParams;
    if (p.is_inf()) return p;
    Fp u = fe_dbl<F>(p.y);
    Fp v = fe_sqr<F>(u);
    Fp w = fe_mul<F>(u, v);
    Fp s = fe_mul<F>(p.x, v);
    Fp xx = fe_sqr<F>(p.x);
    Fp m = fe_add<F>(fe_dbl<F>(xx), xx);
    G1Xyzz r;
    r.x = fe_sub<F>(fe_sqr<F>(m), fe_dbl<F>(s));
    r.y = fe_sub<F>(fe_mul<F>(m, fe_sub<F>(s, r.x)), fe_mul<F>(w, p.y));
    r.zz = fe_mul<F>(v, p.zz);
    r.zzz = fe_mul<F>(w, p.zzz);
    return r;
}

// acc += q (q affine); the MSM inner loop.
FB_HD void g1_add_mixed(G1Xyzz& acc, const G1Affine& q) {
    using F = FpParams;
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = G1Xyzz{q.x, q.y, Fp::one(), Fp::one()};
        return;
    }
    Fp u2 = fe_mul<F>(q.x, acc.zz);
    Fp s2 = fe_mul<F>(q.y, acc.zzz);
    Fp p = fe_sub<F>(u2, acc.x);
    Fp r = fe_sub<F>(s2, acc.y);
    if (p.is_zero()) {
        acc = r.is_zero() ? g1_dbl_affine(q) : G1Xyzz::inf();
        return;
    }
    Fp pp = fe_sqr<F>(p);
    Fp ppp = fe_mul<F>(p, pp);
    Fp qq = fe_mul<F>(acc.x, pp);
    Fp x3 = fe_sub<F>(fe_sub<F>(fe_sqr<F>(r), ppp), fe_dbl<F>(qq));
    Fp y3 = fe_sub<F>(fe_mul<F>(r, fe_sub<F>(qq, x3)), fe_mul<F>(acc.y, ppp));
    acc.x = x3;
    acc.y = y3;
    acc.zz = fe_mul<F>(acc.zz, pp);
    acc.zzz = fe_mul<F>(acc.zzz, ppp);
}

// acc += q (both XYZZ); used by the bucket reduction and the final combines.
FB_HD void g1_add(G1Xyzz& acc, const G1Xyzz& q) {
    using F = FpParams;
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = q;
        return;
    }
    Fp u1 = fe_mul<F>(acc.x, q.zz);
    Fp u2 = fe_mul<F>(q.x, acc.zz);
    Fp s1 = fe_mul<F>(acc.y, q.zzz);
    Fp s2 = fe_mul<F>(q.y, acc.zzz);
    Fp p = fe_sub<F>(u2, u1);
    Fp r = fe_sub<F>(s2, s1);
    if (p.is_zero()) {
        acc = r.is_zero() ? g1_dbl(acc) : G1Xyzz::inf();
        return;
    }
    Fp pp = fe_sqr<F>(p);
    Fp ppp = fe_mul<F>(p, pp);
    Fp qq = fe_mul<F>(u1, pp);
    Fp x3 = fe_sub<F>(fe_sub<F>(fe_sqr<F>(r), ppp), fe_dbl<F>(qq));
    Fp y3 = fe_sub<F>(fe_mul<F>(r, fe_sub<F>(qq, x3)), fe_mul<F>(s1, ppp));
    acc.x = x3;
    acc.y = y3;
    acc.zz = fe_mul<F>(fe_mul<F>(acc.zz, q.zz), pp);
    acc.zzz = fe_mul<F>(fe_mul<F>(acc.zzz, q.zzz), ppp);
}

// XYZZ -> affine with one field inversion: 1/ZZZ gives 1/ZZ = ZZ^2 * (1/ZZZ)^... no:
// ZZ^3 = ZZZ^2, so 1/ZZ = ZZ^2/ZZZ^2 = (ZZ/ZZZ)^2 and 1/ZZZ is the inversion.
FB_HD G1Affine g1_to_affine(const G1Xyzz& p) {
    using F = FpParams;
    if (p.is_inf()) return G1Affine::inf();
    Fp izzz = fe_inv<F>(p.zzz);
    Fp t = fe_mul<F>(p.zz, izzz);
    Fp izz = fe_sqr<F>(t);
    return G1Affine{fe_mul<F>(p.x, izz), fe_mul<F>(p.y, izzz)};
}

// same, when the caller already holds 1/ZZZ (batch inversion)
FB_HD G1Affine g1_to_affine_with_inv(const G1Xyzz& p, const Fp& izzz) {
    using F = FpParams;
    if (p.is_inf()) return G1Affine::inf();
    Fp t = fe_mul<F>(p.zz, izzz);
    Fp izz = fe_sqr<F>(t);
    return G1Affine{fe_mul<F>(p.x, izz), fe_mul<F>(p.y, izzz)};
}

FB_HD bool g1_on_curve(const G1Affine& p) {
    using F = FpParams;
    if (p.is_inf()) return true;
    Fp four = Fp::one();
    four = fe_dbl<F>(fe_dbl<F>(four));
    Fp rhs = fe_add<F>(fe_mul<F>(fe_sqr<F>(p.x), p.x), four);
    return fe_sqr<F>(p.y) == rhs;
}

// ---------------------------------------------------------------- wire codecs
// FsG1::to_bytes = blst_p1_compress (reference src/rpc.rs:241, 268-269, 311,
// 330): 48 bytes big-endian x, bit7 = compressed, bit6 = infinity, bit5 = y is
// the larger of the two roots.
FB_HD bool fp_is_lex_larger(const Fp& y_canonical) {
    // y > (p-1)/2
    constexpr uint32_t half[12] = {0xffffd555u, 0xdcff7fffu, 0x58a9ffffu, 0x0f55ffffu, 0x7b587b12u, 0xb3986950u,
                                   0x79c2895fu, 0xb23ba5c2u, 0x21a5d66bu, 0x258dd3dbu, 0x1cbff34du, 0x0d0088f5u};
    for (int i = 11; i >= 0; i--) {
        if (y_canonical.v[i] > half[i]) return true;
        if (y_canonical.v[i] < half[i]) return false;
    }
    return false;
}

FB_HD void fp_to_be48(const Fp& canonical, uint8_t* out) {
    for (int i = 0; i < 12; i++) {
        uint32_t w = canonical.v[11 - i];
        out[4 * i + 0] = (uint8_t)(w >> 24);
        out[4 * i + 1] = (uint8_t)(w >> 16);
        out[4 * i + 2] = (uint8_t)(w >> 8);
        out[4 * i + 3] = (uint8_t)w;
    }
}

FB_HD Fp fp_from_be48(const uint8_t* in) {
    Fp r;
    for (int i = 0; i < 12; i++) {
        const uint8_t* b = in + 4 * (11 - i);
        r.v[i] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    return r;
}

FB_HD void g1_compress(const G1Affine& p, uint8_t* out48) {
    using F = FpParams;
    if (p.is_inf()) {
        out48[0] = 0xC0;
        for (int i = 1; i < 48; i++) out48[i] = 0;
        return;
    }
    fp_to_be48(fe_from_mont<F>(p.x), out48);
    out48[0] |= 0x80;
    if (fp_is_lex_larger(fe_from_mont<F>(p.y))) out48[0] |= 0x20;
}

FB_HD void g1_serialize(const G1Affine& p, uint8_t* out96) {  // blst_p1_serialize, piano.rs:655
    using F = FpParams;
    if (p.is_inf()) {
        out96[0] = 0x40;
        for (int i = 1; i < 96; i++) out96[i] = 0;
        return;
    }
    fp_to_be48(fe_from_mont<F>(p.x), out96);
    fp_to_be48(fe_from_mont<F>(p.y), out96 + 48);
}

// FsG1::from_bytes = blst_p1_uncompress (reference piano.rs:57, 724, 759):
// on-curve check, no subgroup check.  Returns 0 on success.
FB_HD int g1_decompress(const uint8_t* in48, G1Affine* out) {
    using F = FpParams;
    uint8_t flags = in48[0];
    if (!(flags & 0x80)) return 1;
    uint8_t tmp[48];
    for (int i = 0; i < 48; i++) tmp[i] = in48[i];
    tmp[0] &= 0x1f;
    Fp x = fp_from_be48(tmp);
    if (flags & 0x40) {
        if (!x.is_zero() || (flags & 0x20)) return 2;
        *out = G1Affine::inf();
        return 0;
    }
    Fp pm = Fp::modulus();
    if (fe_geq<F>(x.v, pm.v)) return 3;
    Fp xm = fe_to_mont<F>(x);
    Fp four = fe_dbl<F>(fe_dbl<F>(Fp::one()));
    Fp rhs = fe_add<F>(fe_mul<F>(fe_sqr<F>(xm), xm), four);
    constexpr uint32_t e[12] = {0xffffeaabu, 0xee7fbfffu, 0xac54ffffu, 0x07aaffffu, 0x3dac3d89u, 0xd9cc34a8u,
                                0x3ce144afu, 0xd91dd2e1u, 0x90d2eb35u, 0x92c6e9edu, 0x8e5ff9a6u, 0x0680447au};  // (p+1)/4
    Fp y = fe_pow<F, 12>(rhs, e);
    if (fe_sqr<F>(y) != rhs) return 4;
    bool larger = fp_is_lex_larger(fe_from_mont<F>(y));
    if (larger != ((flags & 0x20) != 0)) y = fe_neg<F>(y);
    out->x = xm;
    out->y = y;
    return 0;
}

FB_HD int g1_deserialize(const uint8_t* in96, G1Affine* out) {  // blst_p1_deserialize, piano.rs:728, 770
    using F = FpParams;
    if (in96[0] & 0x80) return 1;
    if (in96[0] & 0x40) {
        uint8_t o = in96[0] & 0x3f;
        for (int i = 1; i < 96; i++) o |= in96[i];
        if (o) return 2;
        *out = G1Affine::inf();
        return 0;
    }
    Fp x = fp_from_be48(in96), y = fp_from_be48(in96 + 48);
    Fp pm = Fp::modulus();
    if (fe_geq<F>(x.v, pm.v) || fe_geq<F>(y.v, pm.v)) return 3;
    G1Affine p{fe_to_mont<F>(x), fe_to_mont<F>(y)};
    if (!g1_on_curve(p)) return 4;
    *out = p;
    return 0;
}

// generator in Montgomery form
FB_HD G1Affine g1_generator() {
    constexpr uint32_t gx[12] = {0xfd530c16u, 0x5cb38790u, 0x9976fff5u, 0x7817fc67u, 0x143ba1c1u, 0x154f95c7u,
                                 0xf3d0e747u, 0xf0ae6acdu, 0x21dbf440u, 0xedce6eccu, 0x9e0bfb75u, 0x12017741u};
    constexpr uint32_t gy[12] = {0x0ce72271u, 0xbaac93d5u, 0x7918fd8eu, 0x8c22631au, 0x570725ceu, 0xdd595f13u,
                                 0x50405194u, 0x51ac5829u, 0xad0059c0u, 0x0e1c8c3fu, 0x5008a26au, 0x0bbc3efcu};
    G1Affine g;
    for (int i = 0; i < 12; i++) {
        g.x.v[i] = gx[i];
        g.y.v[i] = gy[i];
    }
    return g;
}

// k*P by double-and-add over a canonical (non-Montgomery) 8-limb scalar.
// Only for one-off results (table seeds, tests); the bulk paths are MSM /
// fixed-base tables.
FB_HD G1Xyzz g1_mul_scalar(const G1Affine& p, const uint32_t* k8) {
    G1Xyzz acc = G1Xyzz::inf();
    for (int i = 7; i >= 0; i--)
        for (int bit = 31; bit >= 0; bit--) {
            acc = g1_dbl(acc);
            if ((k8[i] >> bit) & 1) g1_add_mixed(acc, p);
        }
    return acc;
}

}  // namespace fb
