// g2host.cpp - the three G2 elements of the setup (g2, g2^tauX, g2^tauY), host side.
//
// They are not on the commit/open hot path: the reference only writes/reads them
// in the setup file (src/engine/piano.rs:662-665, 700-703, 828-830) and uses them
// in the pairing checks (piano.rs:344-347, 403-410).  Three scalar
// multiplications on the twist  y^2 = x^3 + 4(1+u)  over Fp2 = Fp[u]/(u^2+1) do
// not justify a kernel, so they run here on the same Montgomery field code as the
// device (field.cuh compiled for the host).
//
// The second half of the file is the pairing check behind worker_verify /
// master_verify (piano.rs:358-388, 422-464: blst Pairing::raw_aggregate, final_exp,
// blst_fp12_is_one): optimal-ate Miller loop on affine twist coordinates, tower
// Fp12 = Fp6[w]/(w^2-v), Fp6 = Fp2[v]/(v^3-(1+u)), final exponentiation with the
// (x-1)^2 (x+p)(x^2+p^2-1)+3 hard part.  A few pairings per verify call: host work.
#include <string.h>

#include <vector>

#include "g1.cuh"

namespace fb {

namespace {
using F = FpParams;

// Montgomery product on 6 x 64-bit limbs (the host has a 64x64->128 multiplier; the
// 32-bit carry-chain code in field_mul.cuh is written for the GPU's integer pipe).
// Same representation and the same fully reduced result as fe_mul<FpParams>.
Fp fmul(const Fp& a, const Fp& b) {
    typedef unsigned __int128 u128;
    static const Fp mod = Fp::modulus();
    static const uint64_t ninv = [] {
        uint64_t m0, inv = 1;
        memcpy(&m0, mod.v, 8);
        for (int i = 0; i < 6; i++) inv *= 2 - m0 * inv;
        return (uint64_t)0 - inv;
    }();
    uint64_t x[6], y[6], m[6], t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    memcpy(x, a.v, 48);
    memcpy(y, b.v, 48);
    memcpy(m, mod.v, 48);
#pragma GCC unroll 6
    for (int i = 0; i < 6; i++) {
        u128 carry = 0;
#pragma GCC unroll 6
        for (int j = 0; j < 6; j++) {
            u128 s = (u128)x[i] * y[j] + t[j] + carry;
            t[j] = (uint64_t)s;
            carry = s >> 64;
        }
        u128 s = (u128)t[6] + carry;
        t[6] = (uint64_t)s;
        t[7] = (uint64_t)(s >> 64);
        uint64_t q = t[0] * ninv;
        carry = ((u128)q * m[0] + t[0]) >> 64;
#pragma GCC unroll 6
        for (int j = 1; j < 6; j++) {
            s = (u128)q * m[j] + t[j] + carry;
            t[j - 1] = (uint64_t)s;
            carry = s >> 64;
        }
        s = (u128)t[6] + carry;
        t[5] = (uint64_t)s;
        t[6] = t[7] + (uint64_t)(s >> 64);
    }
    bool ge = t[6] != 0;
    if (!ge) {
        ge = true;
        for (int j = 5; j >= 0; j--)
            if (t[j] != m[j]) {
                ge = t[j] > m[j];
                break;
            }
    }
    if (ge) {
        uint64_t borrow = 0;
        for (int j = 0; j < 6; j++) {
            u128 d = (u128)t[j] - m[j] - borrow;
            t[j] = (uint64_t)d;
            borrow = (uint64_t)(d >> 64) & 1;
        }
    }
    Fp r;
    memcpy(r.v, t, 48);
    return r;
}

// a + b, a - b, -a mod p on 64-bit limbs (p < 2^381: the sum of two reduced values fits)
Fp fadd(const Fp& a, const Fp& b) {
    typedef unsigned __int128 u128;
    static const Fp mod = Fp::modulus();
    uint64_t x[6], y[6], m[6], t[6], d[6];
    memcpy(x, a.v, 48);
    memcpy(y, b.v, 48);
    memcpy(m, mod.v, 48);
    u128 c = 0;
#pragma GCC unroll 6
    for (int j = 0; j < 6; j++) {
        c += (u128)x[j] + y[j];
        t[j] = (uint64_t)c;
        c >>= 64;
    }
    uint64_t borrow = 0;
#pragma GCC unroll 6
    for (int j = 0; j < 6; j++) {
        u128 e = (u128)t[j] - m[j] - borrow;
        d[j] = (uint64_t)e;
        borrow = (uint64_t)(e >> 64) & 1;
    }
    Fp r;
    memcpy(r.v, borrow ? t : d, 48);
    return r;
}
Fp fsub(const Fp& a, const Fp& b) {
    typedef unsigned __int128 u128;
    static const Fp mod = Fp::modulus();
    uint64_t x[6], y[6], m[6], t[6];
    memcpy(x, a.v, 48);
    memcpy(y, b.v, 48);
    memcpy(m, mod.v, 48);
    uint64_t borrow = 0;
#pragma GCC unroll 6
    for (int j = 0; j < 6; j++) {
        u128 e = (u128)x[j] - y[j] - borrow;
        t[j] = (uint64_t)e;
        borrow = (uint64_t)(e >> 64) & 1;
    }
    if (borrow) {
        u128 c = 0;
        for (int j = 0; j < 6; j++) {
            c += (u128)t[j] + m[j];
            t[j] = (uint64_t)c;
            c >>= 64;
        }
    }
    Fp r;
    memcpy(r.v, t, 48);
    return r;
}
Fp fneg(const Fp& a) { return a.is_zero() ? a : fsub(Fp::zero(), a); }

struct Fp2 {
    Fp a, b;  // a + b*u
};
Fp2 f2_zero() { return Fp2{Fp::zero(), Fp::zero()}; }
Fp2 f2_one() { return Fp2{Fp::one(), Fp::zero()}; }
bool f2_is_zero(const Fp2& x) { return x.a.is_zero() && x.b.is_zero(); }
bool f2_eq(const Fp2& x, const Fp2& y) { return x.a == y.a && x.b == y.b; }
Fp2 f2_add(const Fp2& x, const Fp2& y) { return Fp2{fadd(x.a, y.a), fadd(x.b, y.b)}; }
Fp2 f2_sub(const Fp2& x, const Fp2& y) { return Fp2{fsub(x.a, y.a), fsub(x.b, y.b)}; }
Fp2 f2_neg(const Fp2& x) { return Fp2{fneg(x.a), fneg(x.b)}; }
Fp2 f2_mul(const Fp2& x, const Fp2& y) {
    Fp aa = fmul(x.a, y.a), bb = fmul(x.b, y.b);
    Fp cross = fmul(fadd(x.a, x.b), fadd(y.a, y.b));
    return Fp2{fsub(aa, bb), fsub(fsub(cross, aa), bb)};
}
Fp2 f2_sqr(const Fp2& x) {  // (a+b)(a-b) + 2ab u
    Fp ab = fmul(x.a, x.b);
    return Fp2{fmul(fadd(x.a, x.b), fsub(x.a, x.b)), fadd(ab, ab)};
}
Fp2 f2_dbl(const Fp2& x) { return f2_add(x, x); }
Fp2 f2_inv(const Fp2& x) {
    Fp nrm = fadd(fmul(x.a, x.a), fmul(x.b, x.b));
    Fp i = fe_inv_gcd<F>(nrm);
    return Fp2{fmul(x.a, i), fneg(fmul(x.b, i))};
}
Fp fp_sqrt_candidate(const Fp& x) {
    constexpr uint32_t e[12] = {0xffffeaabu, 0xee7fbfffu, 0xac54ffffu, 0x07aaffffu, 0x3dac3d89u, 0xd9cc34a8u,
                                0x3ce144afu, 0xd91dd2e1u, 0x90d2eb35u, 0x92c6e9edu, 0x8e5ff9a6u, 0x0680447au};
    return fe_pow<F, 12>(x, e);
}
bool f2_sqrt(const Fp2& x, Fp2* out) {
    if (f2_is_zero(x)) {
        *out = x;
        return true;
    }
    if (x.b.is_zero()) {
        Fp s = fp_sqrt_candidate(x.a);
        if (fe_sqr<F>(s) == x.a) {
            *out = Fp2{s, Fp::zero()};
            return true;
        }
        Fp na = fe_neg<F>(x.a);
        s = fp_sqrt_candidate(na);
        if (fe_sqr<F>(s) != na) return false;
        *out = Fp2{Fp::zero(), s};
        return true;
    }
    Fp nrm = fe_add<F>(fe_sqr<F>(x.a), fe_sqr<F>(x.b));
    Fp alpha = fp_sqrt_candidate(nrm);
    if (fe_sqr<F>(alpha) != nrm) return false;
    Fp two = fe_dbl<F>(Fp::one());
    Fp inv2 = fe_inv<F>(two);
    Fp delta = fe_mul<F>(fe_add<F>(x.a, alpha), inv2);
    Fp x0 = fp_sqrt_candidate(delta);
    if (fe_sqr<F>(x0) != delta) {
        delta = fe_mul<F>(fe_sub<F>(x.a, alpha), inv2);
        x0 = fp_sqrt_candidate(delta);
        if (fe_sqr<F>(x0) != delta) return false;
    }
    Fp x1 = fe_mul<F>(x.b, fe_inv<F>(fe_dbl<F>(x0)));
    Fp2 r{x0, x1};
    if (!f2_eq(f2_sqr(r), x)) return false;
    *out = r;
    return true;
}

struct G2Jac {
    Fp2 x, y, z;  // z == 0: identity
};
struct G2Aff {
    Fp2 x, y;
    bool inf;
};

Fp fp_from_hex_be(const char* hex) {  // 96 hex digits -> Montgomery
    Fp r = Fp::zero();
    for (int i = 0; i < 96; i++) {
        char ch = hex[i];
        uint32_t d = ch <= '9' ? ch - '0' : (ch | 32) - 'a' + 10;
        int bitpos = (95 - i) * 4;
        r.v[bitpos / 32] |= d << (bitpos % 32);
    }
    return fe_to_mont<F>(r);
}

G2Aff g2_generator() {
    G2Aff g;
    g.inf = false;
    g.x.a = fp_from_hex_be("024aa2b2f08f0a91260805272dc51051c6e47ad4fa403b02b4510b647ae3d1770bac0326a805bbefd48056c8c121bdb8");
    g.x.b = fp_from_hex_be("13e02b6052719f607dacd3a088274f65596bd0d09920b61ab5da61bbdc7f5049334cf11213945d57e5ac7d055d042b7e");
    g.y.a = fp_from_hex_be("0ce5d527727d6e118cc9cdc6da2e351aadfd9baa8cbdd3a76d429a695160d12c923ac9cc3baca289e193548608b82801");
    g.y.b = fp_from_hex_be("0606c4a02ea734cc32acd2b02bc28b99cb3e287e85a763af267492ab572e99ab3f370d275cec1da1aaa9075ff05f79be");
    return g;
}

G2Jac j_dbl(const G2Jac& p) {  // dbl-2009-l, a = 0
    if (f2_is_zero(p.z)) return p;
    Fp2 A = f2_sqr(p.x), B = f2_sqr(p.y), C = f2_sqr(B);
    Fp2 t = f2_sqr(f2_add(p.x, B));
    Fp2 D = f2_dbl(f2_sub(f2_sub(t, A), C));
    Fp2 E = f2_add(f2_dbl(A), A);
    Fp2 Fq = f2_sqr(E);
    G2Jac r;
    r.x = f2_sub(Fq, f2_dbl(D));
    Fp2 c8 = f2_dbl(f2_dbl(f2_dbl(C)));
    r.y = f2_sub(f2_mul(E, f2_sub(D, r.x)), c8);
    r.z = f2_dbl(f2_mul(p.y, p.z));
    return r;
}

G2Jac j_add_affine(const G2Jac& p, const G2Aff& q) {
    if (q.inf) return p;
    if (f2_is_zero(p.z)) return G2Jac{q.x, q.y, f2_one()};
    Fp2 z1z1 = f2_sqr(p.z);
    Fp2 u2 = f2_mul(q.x, z1z1);
    Fp2 s2 = f2_mul(f2_mul(q.y, p.z), z1z1);
    Fp2 h = f2_sub(u2, p.x), rr = f2_sub(s2, p.y);
    if (f2_is_zero(h)) {
        if (f2_is_zero(rr)) return j_dbl(p);
        return G2Jac{f2_zero(), f2_one(), f2_zero()};
    }
    Fp2 hh = f2_sqr(h), hhh = f2_mul(h, hh), v = f2_mul(p.x, hh);
    G2Jac r;
    r.x = f2_sub(f2_sub(f2_sqr(rr), hhh), f2_dbl(v));
    r.y = f2_sub(f2_mul(rr, f2_sub(v, r.x)), f2_mul(p.y, hhh));
    r.z = f2_mul(p.z, h);
    return r;
}

G2Aff j_to_affine(const G2Jac& p) {
    G2Aff r;
    if (f2_is_zero(p.z)) {
        r.inf = true;
        r.x = r.y = f2_zero();
        return r;
    }
    Fp2 zi = f2_inv(p.z), zi2 = f2_sqr(zi);
    r.inf = false;
    r.x = f2_mul(p.x, zi2);
    r.y = f2_mul(p.y, f2_mul(zi2, zi));
    return r;
}

G2Aff g2_mul(const G2Aff& p, const uint32_t k[8]) {
    G2Jac acc{f2_zero(), f2_one(), f2_zero()};
    for (int i = 7; i >= 0; i--)
        for (int bit = 31; bit >= 0; bit--) {
            acc = j_dbl(acc);
            if ((k[i] >> bit) & 1) acc = j_add_affine(acc, p);
        }
    return j_to_affine(acc);
}

bool f2_lex_larger(const Fp2& y) {  // ZCash rule: compare the u-coefficient first
    Fp b = fe_from_mont<F>(y.b);
    if (!b.is_zero()) return fp_is_lex_larger(b);
    return fp_is_lex_larger(fe_from_mont<F>(y.a));
}

// blst_p2_compress layout: x.c1 || x.c0 with flags in byte 0 (piano.rs:662)
void g2_compress(const G2Aff& p, uint8_t out[96]) {
    if (p.inf) {
        memset(out, 0, 96);
        out[0] = 0xC0;
        return;
    }
    fp_to_be48(fe_from_mont<F>(p.x.b), out);
    fp_to_be48(fe_from_mont<F>(p.x.a), out + 48);
    out[0] |= 0x80;
    if (f2_lex_larger(p.y)) out[0] |= 0x20;
}
void g2_serialize(const G2Aff& p, uint8_t out[192]) {  // blst_p2_serialize (piano.rs:665)
    if (p.inf) {
        memset(out, 0, 192);
        out[0] = 0x40;
        return;
    }
    fp_to_be48(fe_from_mont<F>(p.x.b), out);
    fp_to_be48(fe_from_mont<F>(p.x.a), out + 48);
    fp_to_be48(fe_from_mont<F>(p.y.b), out + 96);
    fp_to_be48(fe_from_mont<F>(p.y.a), out + 144);
}
bool fp_read(const uint8_t* in48, uint8_t mask0, Fp* out) {
    uint8_t tmp[48];
    memcpy(tmp, in48, 48);
    tmp[0] &= mask0;
    Fp x = fp_from_be48(tmp);
    Fp pm = Fp::modulus();
    if (fe_geq<F>(x.v, pm.v)) return false;
    *out = fe_to_mont<F>(x);
    return true;
}
bool g2_on_curve(const G2Aff& p) {
    if (p.inf) return true;
    Fp four = fe_dbl<F>(fe_dbl<F>(Fp::one()));
    Fp2 b{four, four};
    return f2_eq(f2_sqr(p.y), f2_add(f2_mul(f2_sqr(p.x), p.x), b));
}
bool g2_decompress(const uint8_t in[96], G2Aff* out) {
    uint8_t flags = in[0];
    if (!(flags & 0x80)) return false;
    Fp c1, c0;
    if (!fp_read(in, 0x1f, &c1) || !fp_read(in + 48, 0xff, &c0)) return false;
    if (flags & 0x40) {
        if (!c1.is_zero() || !c0.is_zero() || (flags & 0x20)) return false;
        out->inf = true;
        out->x = out->y = f2_zero();
        return true;
    }
    Fp four = fe_dbl<F>(fe_dbl<F>(Fp::one()));
    Fp2 x{c0, c1}, b{four, four}, y;
    if (!f2_sqrt(f2_add(f2_mul(f2_sqr(x), x), b), &y)) return false;
    if (f2_lex_larger(y) != ((flags & 0x20) != 0)) y = f2_neg(y);
    out->inf = false;
    out->x = x;
    out->y = y;
    return true;
}
bool g2_deserialize(const uint8_t in[192], G2Aff* out) {
    if (in[0] & 0x80) return false;
    if (in[0] & 0x40) {
        out->inf = true;
        out->x = out->y = f2_zero();
        return true;
    }
    G2Aff p;
    p.inf = false;
    if (!fp_read(in, 0xff, &p.x.b) || !fp_read(in + 48, 0xff, &p.x.a) || !fp_read(in + 96, 0xff, &p.y.b) ||
        !fp_read(in + 144, 0xff, &p.y.a))
        return false;
    if (!g2_on_curve(p)) return false;
    *out = p;
    return true;
}

void secret_to_scalar(const uint8_t s[32], uint32_t k[8]) {  // big-endian bytes mod r, canonical limbs
    Fr x;
    for (int i = 0; i < 8; i++) {
        const uint8_t* b = s + 4 * (7 - i);
        x.v[i] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    Fr m = Fr::modulus();
    for (int rep = 0; rep < 3; rep++)
        if (fe_geq<FrParams>(x.v, m.v)) {
            x.v[0] = sub_cc(x.v[0], m.v[0]);
            for (int i = 1; i < 7; i++) x.v[i] = subc_cc(x.v[i], m.v[i]);
            x.v[7] = subc(x.v[7], m.v[7]);
        }
    memcpy(k, x.v, 32);
}

// ================================================================= pairing
// Fp6 = Fp2[v]/(v^3 - xi), xi = 1 + u;  Fp12 = Fp6[w]/(w^2 - v).  Read as a
// polynomial in w (w^6 = xi) the coefficient of w^(2i) is a.c[i], of w^(2i+1) b.c[i].
struct Fp6 {
    Fp2 c[3];
};
struct Fp12 {
    Fp6 a, b;
};
Fp2 f2_conj(const Fp2& x) { return Fp2{x.a, fneg(x.b)}; }
Fp2 f2_mul_xi(const Fp2& x) { return Fp2{fsub(x.a, x.b), fadd(x.a, x.b)}; }  // (a+bu)(1+u)
Fp2 f2_mul_fp(const Fp2& x, const Fp& s) { return Fp2{fmul(x.a, s), fmul(x.b, s)}; }

Fp6 f6_zero() { return Fp6{{f2_zero(), f2_zero(), f2_zero()}}; }
Fp6 f6_one() { return Fp6{{f2_one(), f2_zero(), f2_zero()}}; }
Fp6 f6_add(const Fp6& x, const Fp6& y) { return Fp6{{f2_add(x.c[0], y.c[0]), f2_add(x.c[1], y.c[1]), f2_add(x.c[2], y.c[2])}}; }
Fp6 f6_sub(const Fp6& x, const Fp6& y) { return Fp6{{f2_sub(x.c[0], y.c[0]), f2_sub(x.c[1], y.c[1]), f2_sub(x.c[2], y.c[2])}}; }
Fp6 f6_neg(const Fp6& x) { return Fp6{{f2_neg(x.c[0]), f2_neg(x.c[1]), f2_neg(x.c[2])}}; }
Fp6 f6_mul_v(const Fp6& x) { return Fp6{{f2_mul_xi(x.c[2]), x.c[0], x.c[1]}}; }
bool f6_eq(const Fp6& x, const Fp6& y) { return f2_eq(x.c[0], y.c[0]) && f2_eq(x.c[1], y.c[1]) && f2_eq(x.c[2], y.c[2]); }
Fp6 f6_mul(const Fp6& x, const Fp6& y) {  // Karatsuba, 6 Fp2 products
    Fp2 v0 = f2_mul(x.c[0], y.c[0]), v1 = f2_mul(x.c[1], y.c[1]), v2 = f2_mul(x.c[2], y.c[2]);
    Fp2 t12 = f2_sub(f2_sub(f2_mul(f2_add(x.c[1], x.c[2]), f2_add(y.c[1], y.c[2])), v1), v2);
    Fp2 t01 = f2_sub(f2_sub(f2_mul(f2_add(x.c[0], x.c[1]), f2_add(y.c[0], y.c[1])), v0), v1);
    Fp2 t02 = f2_sub(f2_sub(f2_mul(f2_add(x.c[0], x.c[2]), f2_add(y.c[0], y.c[2])), v0), v2);
    return Fp6{{f2_add(v0, f2_mul_xi(t12)), f2_add(t01, f2_mul_xi(v2)), f2_add(t02, v1)}};
}
Fp6 f6_inv(const Fp6& x) {
    Fp2 c0 = f2_sub(f2_sqr(x.c[0]), f2_mul_xi(f2_mul(x.c[1], x.c[2])));
    Fp2 c1 = f2_sub(f2_mul_xi(f2_sqr(x.c[2])), f2_mul(x.c[0], x.c[1]));
    Fp2 c2 = f2_sub(f2_sqr(x.c[1]), f2_mul(x.c[0], x.c[2]));
    Fp2 t = f2_add(f2_mul(x.c[0], c0), f2_mul_xi(f2_add(f2_mul(x.c[2], c1), f2_mul(x.c[1], c2))));
    Fp2 ti = f2_inv(t);
    return Fp6{{f2_mul(c0, ti), f2_mul(c1, ti), f2_mul(c2, ti)}};
}

Fp12 f12_one() { return Fp12{f6_one(), f6_zero()}; }
bool f12_is_one(const Fp12& x) { return f6_eq(x.a, f6_one()) && f6_eq(x.b, f6_zero()); }
Fp12 f12_conj(const Fp12& x) { return Fp12{x.a, f6_neg(x.b)}; }  // x^(p^6): w -> -w
Fp12 f12_mul(const Fp12& x, const Fp12& y) {
    Fp6 aa = f6_mul(x.a, y.a), bb = f6_mul(x.b, y.b);
    Fp6 cross = f6_mul(f6_add(x.a, x.b), f6_add(y.a, y.b));
    return Fp12{f6_add(aa, f6_mul_v(bb)), f6_sub(f6_sub(cross, aa), bb)};
}
Fp12 f12_sqr(const Fp12& x) {  // (a + bw)^2 = (a+b)(a+vb) - ab - v ab + 2ab w
    Fp6 ab = f6_mul(x.a, x.b);
    Fp6 t = f6_mul(f6_add(x.a, x.b), f6_add(x.a, f6_mul_v(x.b)));
    return Fp12{f6_sub(f6_sub(t, ab), f6_mul_v(ab)), f6_add(ab, ab)};
}
Fp12 f12_inv(const Fp12& x) {
    Fp6 t = f6_inv(f6_sub(f6_mul(x.a, x.a), f6_mul_v(f6_mul(x.b, x.b))));
    return Fp12{f6_mul(x.a, t), f6_neg(f6_mul(x.b, t))};
}

// x^p: the coefficient c_k of w^k goes to conj(c_k) * gamma^k, gamma = xi^((p-1)/6) = w^(p-1)
struct FrobTable {
    Fp2 g[6];
    FrobTable() {
        constexpr uint32_t e[12] = {0xfffff1c7u, 0x49aa7fffu, 0x72e35555u, 0x051caaaau, 0xd3c82906u, 0xe688231au,
                                    0x7deb831fu, 0xe613e1ebu, 0xb5e1f223u, 0x0c849bf3u, 0x5eeaa66fu, 0x045582fcu};  // (p-1)/6
        Fp2 xi{Fp::one(), Fp::one()}, acc = f2_one();
        for (int i = 11; i >= 0; i--)
            for (int bit = 31; bit >= 0; bit--) {
                acc = f2_sqr(acc);
                if ((e[i] >> bit) & 1) acc = f2_mul(acc, xi);
            }
        g[0] = f2_one();
        for (int k = 1; k < 6; k++) g[k] = f2_mul(g[k - 1], acc);
    }
};
Fp12 f12_frob(const Fp12& x) {
    static const FrobTable tab;
    Fp12 r;
    for (int i = 0; i < 3; i++) {
        r.a.c[i] = f2_mul(f2_conj(x.a.c[i]), tab.g[2 * i]);
        r.b.c[i] = f2_mul(f2_conj(x.b.c[i]), tab.g[2 * i + 1]);
    }
    return r;
}

constexpr uint64_t kAbsX = 0xd201000000010000ull;  // the BLS parameter is -kAbsX

// Untwisting (x', y') -> (x'/w^2, y'/w^3), the line through T with twist slope lam
// evaluated at P = (xp, yp), times w^3 (in Fp4, erased by the final exponentiation):
//     (lam xT - yT) - lam xp w^2 + yp w^3
// T is kept in Jacobian coordinates and the line is scaled by the slope's
// denominator (an Fp2 factor, erased as well), so the loop has no inversion.
struct Line {
    Fp2 c0, c2, c3;  // coefficients of w^0, w^2, w^3
};

// f * (c0 + c2 w^2 + c3 w^3): as Fp6 pairs the line is (c0, c2, 0) + (0, c3, 0) w
Fp6 f6_mul_by_01(const Fp6& x, const Fp2& b0, const Fp2& b1) {
    Fp2 v0 = f2_mul(x.c[0], b0), v1 = f2_mul(x.c[1], b1);
    Fp2 t01 = f2_sub(f2_sub(f2_mul(f2_add(x.c[0], x.c[1]), f2_add(b0, b1)), v0), v1);
    return Fp6{{f2_add(v0, f2_mul_xi(f2_mul(x.c[2], b1))), t01, f2_add(f2_mul(x.c[2], b0), v1)}};
}
Fp6 f6_mul_by_1(const Fp6& x, const Fp2& b1) {
    return Fp6{{f2_mul_xi(f2_mul(x.c[2], b1)), f2_mul(x.c[0], b1), f2_mul(x.c[1], b1)}};
}
Fp12 f12_mul_line(const Fp12& f, const Line& l) {
    Fp6 aa = f6_mul_by_01(f.a, l.c0, l.c2), bb = f6_mul_by_1(f.b, l.c3);
    Fp6 cross = f6_mul_by_01(f6_add(f.a, f.b), l.c0, f2_add(l.c2, l.c3));
    return Fp12{f6_add(aa, f6_mul_v(bb)), f6_sub(f6_sub(cross, aa), bb)};
}

struct PairIn {
    Fp xp, yp;
    G2Aff q;
};

// T <- 2T, returning the tangent at T scaled by 2 Y Z^3:
//   lam = 3X^2 / (2YZ);  c0 = 3X^3 - 2Y^2,  c2 = -3X^2 Z^2 xp,  c3 = 2YZ^3 yp
Line double_step(G2Jac& t, const Fp& xp, const Fp& yp) {
    Fp2 A = f2_sqr(t.x), B = f2_sqr(t.y), C = f2_sqr(B), zz = f2_sqr(t.z);
    Fp2 D = f2_dbl(f2_sub(f2_sub(f2_sqr(f2_add(t.x, B)), A), C));
    Fp2 E = f2_add(f2_dbl(A), A);
    Fp2 z3 = f2_dbl(f2_mul(t.y, t.z));
    Line l;
    l.c0 = f2_sub(f2_mul(E, t.x), f2_dbl(B));
    l.c2 = f2_neg(f2_mul_fp(f2_mul(E, zz), xp));
    l.c3 = f2_mul_fp(f2_mul(z3, zz), yp);
    Fp2 x3 = f2_sub(f2_sqr(E), f2_dbl(D));
    t.y = f2_sub(f2_mul(E, f2_sub(D, x3)), f2_dbl(f2_dbl(f2_dbl(C))));
    t.x = x3;
    t.z = z3;
    return l;
}

// T <- T + Q (Q affine), returning the chord through Q scaled by Z3 = Z H:
//   lam = R / (Z H), H = xQ Z^2 - X, R = yQ Z^3 - Y;  c0 = R xQ - Z3 yQ,  c2 = -R xp,  c3 = Z3 yp
Line add_step(G2Jac& t, const G2Aff& q, const Fp& xp, const Fp& yp) {
    Fp2 zz = f2_sqr(t.z);
    Fp2 H = f2_sub(f2_mul(q.x, zz), t.x);
    Fp2 R = f2_sub(f2_mul(q.y, f2_mul(zz, t.z)), t.y);
    Fp2 z3 = f2_mul(t.z, H);
    Line l;
    l.c0 = f2_sub(f2_mul(R, q.x), f2_mul(z3, q.y));
    l.c2 = f2_neg(f2_mul_fp(R, xp));
    l.c3 = f2_mul_fp(z3, yp);
    Fp2 hh = f2_sqr(H), hhh = f2_mul(H, hh), v = f2_mul(t.x, hh);
    Fp2 x3 = f2_sub(f2_sub(f2_sqr(R), hhh), f2_dbl(v));
    t.y = f2_sub(f2_mul(R, f2_sub(v, x3)), f2_mul(t.y, hhh));
    t.x = x3;
    t.z = z3;
    return l;
}

// prod_i f_{x,Q_i}(P_i); pairs with an identity on either side are dropped before the
// call.  Q has order r and the loop scalar is far below r, so T never meets +-Q or O.
Fp12 miller_loop(const std::vector<PairIn>& in) {
    Fp12 f = f12_one();
    std::vector<G2Jac> t(in.size());
    for (size_t i = 0; i < in.size(); i++) t[i] = G2Jac{in[i].q.x, in[i].q.y, f2_one()};
    for (int bit = 62; bit >= 0; bit--) {
        f = f12_sqr(f);
        for (size_t i = 0; i < in.size(); i++) f = f12_mul_line(f, double_step(t[i], in[i].xp, in[i].yp));
        if ((kAbsX >> bit) & 1)
            for (size_t i = 0; i < in.size(); i++) f = f12_mul_line(f, add_step(t[i], in[i].q, in[i].xp, in[i].yp));
    }
    return f12_conj(f);  // negative parameter
}

// g^x for g in the cyclotomic subgroup (inverse = conjugate)
Fp12 cyc_exp_x(const Fp12& g) {
    Fp12 r = g;
    for (int bit = 62; bit >= 0; bit--) {
        r = f12_sqr(r);
        if ((kAbsX >> bit) & 1) r = f12_mul(r, g);
    }
    return f12_conj(r);
}
Fp12 cyc_exp_xm1(const Fp12& g) { return f12_mul(cyc_exp_x(g), f12_conj(g)); }

// f^(3 (p^12-1)/r): easy part (p^6-1)(p^2+1), then 3(p^4-p^2+1)/r = (x-1)^2 (x+p)(x^2+p^2-1) + 3.
// The factor 3 is coprime to r, so "== 1" is unchanged (blst_fp12_is_one, piano.rs:386, 462).
Fp12 final_exp3(const Fp12& f) {
    Fp12 e = f12_mul(f12_conj(f), f12_inv(f));
    e = f12_mul(f12_frob(f12_frob(e)), e);
    Fp12 a = cyc_exp_xm1(cyc_exp_xm1(e));
    Fp12 b = f12_mul(cyc_exp_x(a), f12_frob(a));
    Fp12 c = f12_mul(f12_mul(cyc_exp_x(cyc_exp_x(b)), f12_frob(f12_frob(b))), f12_conj(b));
    return f12_mul(c, f12_mul(f12_sqr(e), e));
}

bool g2_parse(const uint8_t* in, size_t len, G2Aff* out) {
    if (len == 96) return g2_decompress(in, out);
    if (len == 192) return g2_deserialize(in, out);
    return false;
}

bool gather_pairs(const G1Affine* g1, const uint8_t* g2_192, size_t n, std::vector<PairIn>& out) {
    for (size_t k = 0; k < n; k++) {
        G2Aff q;
        if (!g2_deserialize(g2_192 + 192 * k, &q)) return false;
        if (q.inf || (g1[k].x.is_zero() && g1[k].y.is_zero())) continue;  // e(O, Q) = e(P, O) = 1
        out.push_back(PairIn{g1[k].x, g1[k].y, q});
    }
    return true;
}

}  // namespace

bool g2_generate_host(const uint8_t sx[32], const uint8_t sy[32], std::vector<uint8_t> out[3]) {
    G2Aff g = g2_generator();
    uint32_t kx[8], ky[8];
    secret_to_scalar(sx, kx);
    secret_to_scalar(sy, ky);
    G2Aff pts[3] = {g, g2_mul(g, kx), g2_mul(g, ky)};  // piano.rs:1017-1021
    for (int i = 0; i < 3; i++) {
        out[i].resize(96);
        g2_compress(pts[i], out[i].data());
    }
    return true;
}

bool g2_recode_host(const std::vector<uint8_t>& in, bool to_compressed, std::vector<uint8_t>& out) {
    G2Aff p;
    if (to_compressed) {
        if (in.size() != 192 || !g2_deserialize(in.data(), &p)) return false;
        out.resize(96);
        g2_compress(p, out.data());
    } else {
        if (in.size() != 96 || !g2_decompress(in.data(), &p)) return false;
        out.resize(192);
        g2_serialize(p, out.data());
    }
    return true;
}

// blst_p1_deserialize layout (96 bytes, x || y big-endian, bit 6 of byte 0 = infinity) -> Montgomery affine
bool g1_deserialize_host(const uint8_t in[96], G1Affine* out) {
    if (in[0] & 0x80) return false;
    if (in[0] & 0x40) {
        out->x = out->y = Fp::zero();
        return true;
    }
    Fp x, y;
    if (!fp_read(in, 0xff, &x) || !fp_read(in + 48, 0xff, &y)) return false;
    Fp four = fe_dbl<F>(fe_dbl<F>(Fp::one()));
    if (fmul(y, y) != fe_add<F>(fmul(fmul(x, x), x), four)) return false;
    out->x = x;
    out->y = y;
    return true;
}

// out192 = base - k * g2  (g2^tauX - alpha * g2, piano.rs:344-347, 403-410); inputs in
// either setup-file encoding (96 compressed / 192 serialized), k canonical limbs
bool g2_sub_mul_host(const std::vector<uint8_t>& base, const std::vector<uint8_t>& g2, const uint32_t k[8], uint8_t out192[192]) {
    G2Aff b, g;
    if (!g2_parse(base.data(), base.size(), &b) || !g2_parse(g2.data(), g2.size(), &g)) return false;
    G2Aff kg = g2_mul(g, k);
    kg.y = f2_neg(kg.y);
    G2Jac acc = b.inf ? G2Jac{f2_zero(), f2_one(), f2_zero()} : G2Jac{b.x, b.y, f2_one()};
    g2_serialize(j_to_affine(j_add_affine(acc, kg)), out192);
    return true;
}

bool g2_to_serialized_host(const std::vector<uint8_t>& in, uint8_t out192[192]) {
    G2Aff p;
    if (!g2_parse(in.data(), in.size(), &p)) return false;
    g2_serialize(p, out192);
    return true;
}

// prod_k e(g1[k], g2[k]) == 1 ?  1 / 0, or -1 when a G2 element does not parse
// (pairings_verify_single / pairings_verify, piano.rs:358-388, 422-464)
int pairing_product_is_one_host(const G1Affine* g1, const uint8_t* g2_192, size_t n) {
    std::vector<PairIn> pairs;
    if (!gather_pairs(g1, g2_192, n, pairs)) return -1;
    if (pairs.empty()) return 1;
    return f12_is_one(final_exp3(miller_loop(pairs))) ? 1 : 0;
}

// The value the check above compares with one: (prod_k e(g1[k], g2[k]))^3, written as
// the 12 Fp coefficients (48-byte big-endian each) of w^0..w^5, real part first.
bool pairing_value_host(const G1Affine* g1, const uint8_t* g2_192, size_t n, uint8_t out576[576]) {
    std::vector<PairIn> pairs;
    if (!gather_pairs(g1, g2_192, n, pairs)) return false;
    Fp12 v = pairs.empty() ? f12_one() : final_exp3(miller_loop(pairs));
    for (int i = 0; i < 3; i++) {
        const Fp2* c[2] = {&v.a.c[i], &v.b.c[i]};
        for (int h = 0; h < 2; h++) {
            uint8_t* o = out576 + (size_t)(2 * i + h) * 96;
            fp_to_be48(fe_from_mont<F>(c[h]->a), o);
            fp_to_be48(fe_from_mont<F>(c[h]->b), o + 48);
        }
    }
    return true;
}

}  // namespace fb
