// field.cuh - Montgomery prime fields over 32-bit limbs: BLS12-381 Fr (8 limbs)
// and Fp (12 limbs).  Host+device; the device path is what ships, the host path
// exists so tests/ can check the very same routines without a GPU.
//
// Replaces (for the KZG hot path) blst's blst_fr / blst_fp Montgomery arithmetic
// that the reference reaches through rust-kzg-blst (reference Cargo.toml:11-12,
// 29; call sites src/engine/piano.rs:150-158, 194-236).
//
// Representation: little-endian uint32 limbs, value*2^(32N) mod m ("Montgomery
// form"), always fully reduced (< m).  mul is a CIOS-style row loop in which the
// products of even-indexed and odd-indexed limbs run as two independent
// mad.lo.cc/madc.hi.cc carry chains (each lo/hi pair is one IMAD.WIDE.U32.X),
// so a multiply costs 2*N*N wide MADs plus O(N) glue.
#pragma once
#include "limbs.cuh"

namespace fb {

struct FrParams {
    static constexpr int N = 8;
    static constexpr uint32_t INV = 0xffffffffu;  // -m^-1 mod 2^32
    FB_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t t[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u, 0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u};
        return t[i];
    }
    FB_HD static constexpr uint32_t one(int i) {  // 2^256 mod r
        constexpr uint32_t t[8] = {0xfffffffeu, 0x00000001u, 0x00034802u, 0x5884b7fau, 0xecbc4ff5u, 0x998c4fefu, 0xacc5056fu, 0x1824b159u};
        return t[i];
    }
    FB_HD static constexpr uint32_t r2(int i) {  // 2^512 mod r
        constexpr uint32_t t[8] = {0xf3f29c6du, 0xc999e990u, 0x87925c23u, 0x2b6cedcbu, 0x7254398fu, 0x05d31496u, 0x9f59ff11u, 0x0748d9d9u};
        return t[i];
    }
};

struct FpParams {
    static constexpr int N = 12;
    static constexpr uint32_t INV = 0xfffcfffdu;
    FB_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t t[12] = {0xffffaaabu, 0xb9feffffu, 0xb153ffffu, 0x1eabfffeu, 0xf6b0f624u, 0x6730d2a0u,
                                    0xf38512bfu, 0x64774b84u, 0x434bacd7u, 0x4b1ba7b6u, 0x397fe69au, 0x1a0111eau};
        return t[i];
    }
    FB_HD static constexpr uint32_t one(int i) {  // 2^384 mod p
        constexpr uint32_t t[12] = {0x0002fffdu, 0x76090000u, 0xc40c0002u, 0xebf4000bu, 0x53c758bau, 0x5f489857u,
                                    0x70525745u, 0x77ce5853u, 0xa256ec6du, 0x5c071a97u, 0xfa80e493u, 0x15f65ec3u};
        return t[i];
    }
    FB_HD static constexpr uint32_t r2(int i) {  // 2^768 mod p
        constexpr uint32_t t[12] = {0x1c341746u, 0xf4df1f34u, 0x09d104f1u, 0x0a76e6a6u, 0x4c95b6d5u, 0x8de5476cu,
                                    0x939d83c0u, 0x67eb88a9u, 0xb519952du, 0x9a793e85u, 0x92cae3aau, 0x11988fe5u};
        return t[i];
    }
};

template <class P>
struct alignas(16) Fe {
    static constexpr int N = P::N;
    uint32_t v[N];

    FB_HD static Fe zero() {
        Fe r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = 0;
        return r;
    }
    FB_HD static Fe one() {
        Fe r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = P::one(i);
        return r;
    }
    FB_HD static Fe modulus() {
        Fe r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = P::mod(i);
        return r;
    }
    FB_HD static Fe rsquared() {
        Fe r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = P::r2(i);
        return r;
    }
    FB_HD bool is_zero() const {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < N; i++) o |= v[i];
        return o == 0;
    }
    FB_HD bool operator==(const Fe& b) const {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < N; i++) o |= v[i] ^ b.v[i];
        return o == 0;
    }
    FB_HD bool operator!=(const Fe& b) const { return !(*this == b); }
};

// r = a - m if a >= m (a < 2m assumed), branch-free select.
template <class P>
FB_HD void fe_final_sub(uint32_t* r, const uint32_t* a, uint32_t top) {
    constexpr int N = P::N;
    uint32_t t[N];
    t[0] = sub_cc(a[0], P::mod(0));
#pragma unroll
    for (int i = 1; i < N; i++) t[i] = subc_cc(a[i], P::mod(i));
    uint32_t borrow = subc(top, 0);  // top - borrow: all-ones iff a < m (top is 0 or 1)
    bool keep = borrow != 0;
#pragma unroll
    for (int i = 0; i < N; i++) r[i] = keep ? a[i] : t[i];
}

template <class P>
FB_HD Fe<P> fe_add(const Fe<P>& a, const Fe<P>& b) {
    constexpr int N = P::N;
    uint32_t s[N];
    s[0] = add_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < N; i++) s[i] = addc_cc(a.v[i], b.v[i]);
    uint32_t top = addc(0, 0);
    Fe<P> r;
    fe_final_sub<P>(r.v, s, top);
    return r;
}

template <class P>
FB_HD Fe<P> fe_sub(const Fe<P>& a, const Fe<P>& b) {
    constexpr int N = P::N;
    uint32_t s[N];
    s[0] = sub_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < N; i++) s[i] = subc_cc(a.v[i], b.v[i]);
    uint32_t mask = subc(0, 0);  // all-ones on borrow
    Fe<P> r;
    r.v[0] = add_cc(s[0], P::mod(0) & mask);
#pragma unroll
    for (int i = 1; i < N - 1; i++) r.v[i] = addc_cc(s[i], P::mod(i) & mask);
    r.v[N - 1] = addc(s[N - 1], P::mod(N - 1) & mask);
    return r;
}

template <class P>
FB_HD Fe<P> fe_neg(const Fe<P>& a) {
    constexpr int N = P::N;
    Fe<P> r;
    uint32_t mask = a.is_zero() ? 0u : 0xffffffffu;
    r.v[0] = sub_cc(P::mod(0) & mask, a.v[0]);
#pragma unroll
    for (int i = 1; i < N - 1; i++) r.v[i] = subc_cc(P::mod(i) & mask, a.v[i]);
    r.v[N - 1] = subc(P::mod(N - 1) & mask, a.v[N - 1]);
    return r;
}

template <class P>
FB_HD Fe<P> fe_dbl(const Fe<P>& a) { return fe_add<P>(a, a); }

// a >= b as integers (both reduced)
template <class P>
FB_HD bool fe_geq(const uint32_t* a, const uint32_t* b) {
    constexpr int N = P::N;
    (void)sub_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < N; i++) (void)subc_cc(a[i], b[i]);
    return subc(0, 0) == 0;
}

}  // namespace fb
#include "field_mul.cuh"
