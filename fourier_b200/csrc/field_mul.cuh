// field_mul.cuh - Montgomery multiplication for Fe<P> (included by field.cuh).
#pragma once

namespace fb {

// Reference-quality portable CIOS on uint64 accumulators.  Kept as the
// cross-check for the carry-chain version below (tests + GPU self-test).
template <class P>
FB_HD Fe<P> fe_mul_portable(const Fe<P>& a, const Fe<P>& b) {
    constexpr int N = P::N;
    uint32_t t[N + 2];
#pragma unroll
    for (int i = 0; i < N + 2; i++) t[i] = 0;
#pragma unroll
    for (int i = 0; i < N; i++) {
        uint64_t c = 0;
#pragma unroll
        for (int j = 0; j < N; j++) {
            uint64_t x = (uint64_t)a.v[j] * b.v[i] + t[j] + c;
            t[j] = (uint32_t)x;
            c = x >> 32;
        }
        uint64_t x = (uint64_t)t[N] + c;
        t[N] = (uint32_t)x;
        t[N + 1] = (uint32_t)(x >> 32);
        uint32_t m = t[0] * P::INV;
        x = (uint64_t)m * P::mod(0) + t[0];
        c = x >> 32;
#pragma unroll
        for (int j = 1; j < N; j++) {
            x = (uint64_t)m * P::mod(j) + t[j] + c;
            t[j - 1] = (uint32_t)x;
            c = x >> 32;
        }
        x = (uint64_t)t[N] + c;
        t[N - 1] = (uint32_t)x;
        t[N] = t[N + 1] + (uint32_t)(x >> 32);
    }
    // t < 2m; conditional subtract (plain C so it is independent of limbs.cuh)
    uint32_t s[N];
    uint64_t br = 0;
#pragma unroll
    for (int i = 0; i < N; i++) {
        uint64_t d = (uint64_t)t[i] - P::mod(i) - br;
        s[i] = (uint32_t)d;
        br = (d >> 32) & 1;
    }
    bool keep = (t[N] == 0) && br;
    Fe<P> r;
#pragma unroll
    for (int i = 0; i < N; i++) r.v[i] = keep ? t[i] : s[i];
    return r;
}

// Carry-chain CIOS.  State: two limb arrays E ("even", aligned at limb 0) and O
// ("odd", aligned at limb 1) with  V = E + O * 2^32.  Products a_j*b_i with even j
// land on E's aligned limb pairs, odd j on O's, so each row is two carry chains of
// N/2 lo/hi MAD pairs.  After the Montgomery step E[0] == 0; dividing V by 2^32
// swaps the roles of the arrays (O is now aligned at limb 0) and the shift of the
// old E by two limbs is folded into the next row's first chain.
template <class P>
FB_HD Fe<P> fe_mul(const Fe<P>& a, const Fe<P>& b) {
    constexpr int N = P::N;
    static_assert(N % 2 == 0, "even limb count expected");
    uint32_t buf0[N], buf1[N];
    uint32_t* E = buf0;
    uint32_t* O = buf1;
    // ---- row 0: plain products
    {
        const uint32_t bi = b.v[0];
#pragma unroll
        for (int j = 0; j < N; j += 2) {
            E[j] = mul_lo(a.v[j], bi);
            E[j + 1] = mul_hi(a.v[j], bi);
            O[j] = mul_lo(a.v[j + 1], bi);
            O[j + 1] = mul_hi(a.v[j + 1], bi);
        }
        const uint32_t m = E[0] * P::INV;
        O[0] = mad_lo_cc(m, P::mod(1), O[0]);
        O[1] = madc_hi_cc(m, P::mod(1), O[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            O[j - 1] = madc_lo_cc(m, P::mod(j), O[j - 1]);
            O[j] = (j == N - 1) ? madc_hi(m, P::mod(j), O[j]) : madc_hi_cc(m, P::mod(j), O[j]);
        }
        E[0] = mad_lo_cc(m, P::mod(0), E[0]);
        E[1] = madc_hi_cc(m, P::mod(0), E[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            E[j] = madc_lo_cc(m, P::mod(j), E[j]);
            E[j + 1] = madc_hi_cc(m, P::mod(j), E[j + 1]);
        }
        O[N - 1] = addc(O[N - 1], 0);
    }
    // ---- rows 1..N-1
#pragma unroll
    for (int i = 1; i < N; i++) {
        const uint32_t bi = b.v[i];
        // role swap: nE = old O (kept in place), nO = old E shifted down two limbs
        uint32_t* nE = O;
        uint32_t* nO = E;
        nE[0] = add_cc(nE[0], E[1]);
#pragma unroll
        for (int j = 1; j < N; j += 2) {
            const uint32_t lo_in = (j + 1 < N) ? E[j + 1] : 0u;
            const uint32_t hi_in = (j + 2 < N) ? E[j + 2] : 0u;
            nO[j - 1] = madc_lo_cc(a.v[j], bi, lo_in);
            nO[j] = (j == N - 1) ? madc_hi(a.v[j], bi, hi_in) : madc_hi_cc(a.v[j], bi, hi_in);
        }
        nE[0] = mad_lo_cc(a.v[0], bi, nE[0]);
        nE[1] = madc_hi_cc(a.v[0], bi, nE[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            nE[j] = madc_lo_cc(a.v[j], bi, nE[j]);
            nE[j + 1] = madc_hi_cc(a.v[j], bi, nE[j + 1]);
        }
        nO[N - 1] = addc(nO[N - 1], 0);
        const uint32_t m = nE[0] * P::INV;
        nO[0] = mad_lo_cc(m, P::mod(1), nO[0]);
        nO[1] = madc_hi_cc(m, P::mod(1), nO[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            nO[j - 1] = madc_lo_cc(m, P::mod(j), nO[j - 1]);
            nO[j] = (j == N - 1) ? madc_hi(m, P::mod(j), nO[j]) : madc_hi_cc(m, P::mod(j), nO[j]);
        }
        nE[0] = mad_lo_cc(m, P::mod(0), nE[0]);
        nE[1] = madc_hi_cc(m, P::mod(0), nE[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            nE[j] = madc_lo_cc(m, P::mod(j), nE[j]);
            nE[j + 1] = madc_hi_cc(m, P::mod(j), nE[j + 1]);
        }
        nO[N - 1] = addc(nO[N - 1], 0);
        E = nE;
        O = nO;
    }
    // ---- V / 2^32 = (E >> 32) + O
    uint32_t t[N];
    t[0] = add_cc(O[0], E[1]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) t[k] = addc_cc(O[k], E[k + 1]);
    t[N - 1] = addc_cc(O[N - 1], 0);
    const uint32_t top = addc(0, 0);
    Fe<P> r;
    fe_final_sub<P>(r.v, t, top);
    return r;
}

template <class P>
FB_HD Fe<P> fe_sqr(const Fe<P>& a) { return fe_mul<P>(a, a); }

// value -> Montgomery form and back
template <class P>
FB_HD Fe<P> fe_to_mont(const Fe<P>& a) { return fe_mul<P>(a, Fe<P>::rsquared()); }
template <class P>
FB_HD Fe<P> fe_from_mont(const Fe<P>& a) {
    Fe<P> o = Fe<P>::zero();
    o.v[0] = 1;
    return fe_mul<P>(a, o);
}

// a^e for a little-endian limb exponent of EN limbs (square-and-multiply,
// MSB first).  Used for Fermat inversion and the Fp square root.
template <class P, int EN>
FB_HD Fe<P> fe_pow(const Fe<P>& a, const uint32_t* e) {
    Fe<P> r = Fe<P>::one();
    bool started = false;
    for (int i = EN - 1; i >= 0; i--) {
        for (int bit = 31; bit >= 0; bit--) {
            if (started) r = fe_sqr<P>(r);
            if ((e[i] >> bit) & 1) {
                r = started ? fe_mul<P>(r, a) : a;
                started = true;
            }
        }
    }
    return r;
}

template <class P>
FB_HD Fe<P> fe_inv(const Fe<P>& a) {  // a^(m-2); inv(0) = 0
    uint32_t e[P::N];
    uint32_t borrow = 2;
    for (int i = 0; i < P::N; i++) {
        uint32_t m = P::mod(i);
        e[i] = m - borrow;
        borrow = (m < borrow) ? 1u : 0u;
    }
    return fe_pow<P, P::N>(a, e);
}

using Fr = Fe<FrParams>;
using Fp = Fe<FpParams>;

}  // namespace fb
