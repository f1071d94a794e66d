// g2host.cpp - the three G2 elements of the setup (g2, g2^tauX, g2^tauY), host side.
//
// They are not on the commit/open hot path: the reference only writes/reads them
// in the setup file (src/engine/piano.rs:662-665, 700-703, 828-830) and uses them
// in the pairing checks (piano.rs:344-347, 403-410).  Three scalar
// multiplications on the twist  y^2 = x^3 + 4(1+u)  over Fp2 = Fp[u]/(u^2+1) do
// not justify a kernel, so they run here on the same Montgomery field code as the
// device (field.cuh compiled for the host).
#include <string.h>

#include <vector>

#include "g1.cuh"

namespace fb {

namespace {
using F = FpParams;

struct Fp2 {
    Fp a, b;  // a + b*u
};
Fp2 f2_zero() { return Fp2{Fp::zero(), Fp::zero()}; }
Fp2 f2_one() { return Fp2{Fp::one(), Fp::zero()}; }
bool f2_is_zero(const Fp2& x) { return x.a.is_zero() && x.b.is_zero(); }
bool f2_eq(const Fp2& x, const Fp2& y) { return x.a == y.a && x.b == y.b; }
Fp2 f2_add(const Fp2& x, const Fp2& y) { return Fp2{fe_add<F>(x.a, y.a), fe_add<F>(x.b, y.b)}; }
Fp2 f2_sub(const Fp2& x, const Fp2& y) { return Fp2{fe_sub<F>(x.a, y.a), fe_sub<F>(x.b, y.b)}; }
Fp2 f2_neg(const Fp2& x) { return Fp2{fe_neg<F>(x.a), fe_neg<F>(x.b)}; }
Fp2 f2_mul(const Fp2& x, const Fp2& y) {
    Fp aa = fe_mul<F>(x.a, y.a), bb = fe_mul<F>(x.b, y.b);
    Fp cross = fe_mul<F>(fe_add<F>(x.a, x.b), fe_add<F>(y.a, y.b));
    return Fp2{fe_sub<F>(aa, bb), fe_sub<F>(fe_sub<F>(cross, aa), bb)};
}
Fp2 f2_sqr(const Fp2& x) { return f2_mul(x, x); }
Fp2 f2_dbl(const Fp2& x) { return f2_add(x, x); }
Fp2 f2_inv(const Fp2& x) {
    Fp nrm = fe_add<F>(fe_sqr<F>(x.a), fe_sqr<F>(x.b));
    Fp i = fe_inv<F>(nrm);
    return Fp2{fe_mul<F>(x.a, i), fe_neg<F>(fe_mul<F>(x.b, i))};
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

}  // namespace fb
