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

// Montgomery step of one CIOS row: adds m*modulus (m = -X[0]/modulus mod 2^32) to the pair of carry chains
// (Ev aligned at limb 0, Od at limb 1) so that Ev[0] becomes 0.  The caller finishes with Od[N-1] += carry.
//
// Fr has modulus = 1 (mod 2^32) and second limb 2^32 - 1, so INV = -1, m = -Ev[0], and the two lowest
// limb products need no multiplier:  m*1 = m  and  m*(2^32 - 1) = (m - [m != 0]) * 2^32 + Ev[0].
// That takes 2 of the 16 wide MADs of the row (and the IMAD forming m) off the multiplier pipe.
template <class P>
FB_HD void fe_mont_step(uint32_t* Ev, uint32_t* Od) {
    constexpr int N = P::N;
    if constexpr (P::INV == 0xffffffffu && P::mod(0) == 1u && P::mod(1) == 0xffffffffu) {
        const uint32_t e0 = Ev[0];
        const uint32_t m = 0u - e0;
        const uint32_t h1 = m - (m != 0u ? 1u : 0u);
        Od[0] = add_cc(e0, Od[0]);
        Od[1] = addc_cc(h1, Od[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            Od[j - 1] = madc_lo_cc(m, P::mod(j), Od[j - 1]);
            Od[j] = (j == N - 1) ? madc_hi(m, P::mod(j), Od[j]) : madc_hi_cc(m, P::mod(j), Od[j]);
        }
        Ev[0] = add_cc(m, e0);
        Ev[1] = addc_cc(Ev[1], 0u);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            Ev[j] = madc_lo_cc(m, P::mod(j), Ev[j]);
            Ev[j + 1] = madc_hi_cc(m, P::mod(j), Ev[j + 1]);
        }
    } else {
        const uint32_t m = Ev[0] * P::INV;
        Od[0] = mad_lo_cc(m, P::mod(1), Od[0]);
        Od[1] = madc_hi_cc(m, P::mod(1), Od[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            Od[j - 1] = madc_lo_cc(m, P::mod(j), Od[j - 1]);
            Od[j] = (j == N - 1) ? madc_hi(m, P::mod(j), Od[j]) : madc_hi_cc(m, P::mod(j), Od[j]);
        }
        Ev[0] = mad_lo_cc(m, P::mod(0), Ev[0]);
        Ev[1] = madc_hi_cc(m, P::mod(0), Ev[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            Ev[j] = madc_lo_cc(m, P::mod(j), Ev[j]);
            Ev[j + 1] = madc_hi_cc(m, P::mod(j), Ev[j + 1]);
        }
    }
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
        fe_mont_step<P>(E, O);
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
        fe_mont_step<P>(nE, nO);
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

// Montgomery step of a reduction-only row: like fe_mont_step, but limb 0 of the running value is Ev[0] + lone (the
// limb that the division by 2^32 of the previous row left unpaired), and the carry of that addition enters the Od
// chain at limb 1.  On return Ev[0] == 0 (mod 2^32) and the caller adds the final carry into Od[N-1].
template <class P>
FB_HD void fe_mont_step_lone(uint32_t* Ev, uint32_t* Od, uint32_t lone) {
    constexpr int N = P::N;
    if constexpr (P::INV == 0xffffffffu && P::mod(0) == 1u && P::mod(1) == 0xffffffffu) {
        const uint32_t e0 = Ev[0] + lone;   // limb 0 (mod 2^32); the carry is regenerated right below
        const uint32_t m = 0u - e0;
        const uint32_t h1 = m - (m != 0u ? 1u : 0u);
        Ev[0] = add_cc(Ev[0], lone);        // CF = carry into limb 1
        Od[0] = addc_cc(e0, Od[0]);         // m*(2^32-1) = h1*2^32 + e0 at limbs (1, 2)
        Od[1] = addc_cc(h1, Od[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            Od[j - 1] = madc_lo_cc(m, P::mod(j), Od[j - 1]);
            Od[j] = (j == N - 1) ? madc_hi(m, P::mod(j), Od[j]) : madc_hi_cc(m, P::mod(j), Od[j]);
        }
        Ev[0] = add_cc(m, e0);              // = 0, carry 1 unless e0 == 0
        Ev[1] = addc_cc(Ev[1], 0u);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            Ev[j] = madc_lo_cc(m, P::mod(j), Ev[j]);
            Ev[j + 1] = madc_hi_cc(m, P::mod(j), Ev[j + 1]);
        }
    } else {
        const uint32_t m = (Ev[0] + lone) * P::INV;
        Ev[0] = add_cc(Ev[0], lone);
        Od[0] = madc_lo_cc(m, P::mod(1), Od[0]);
        Od[1] = madc_hi_cc(m, P::mod(1), Od[1]);
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            Od[j - 1] = madc_lo_cc(m, P::mod(j), Od[j - 1]);
            Od[j] = (j == N - 1) ? madc_hi(m, P::mod(j), Od[j]) : madc_hi_cc(m, P::mod(j), Od[j]);
        }
        Ev[0] = mad_lo_cc(m, P::mod(0), Ev[0]);
        Ev[1] = madc_hi_cc(m, P::mod(0), Ev[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            Ev[j] = madc_lo_cc(m, P::mod(j), Ev[j]);
            Ev[j + 1] = madc_hi_cc(m, P::mod(j), Ev[j + 1]);
        }
    }
}

// Dedicated Montgomery squaring: N(N-1)/2 off-diagonal products (each once), doubled by a one-bit shift, the N
// squares added by ONE carry chain, then a reduction of the low half only: N(N+1)/2 + N^2 wide MADs (Fp: 222
// against the 288 of fe_mul; Fr: 92 against 112).
//   1. off-diagonal sum D = sum_{i<j} a_i a_j 2^(32(i+j)): the product a_i*a_j occupies limbs (i+j, i+j+1), so the
//      products of one row i with j-i odd tile the odd-aligned array Do from limb 2i+1 on, those with j-i even the
//      even-aligned array De from limb 2i+2 on: two independent lo/hi carry chains per row, each closed by an
//      addc into the limb above its top (which so far holds only such carries).
//   2. S = 2(De + Do) + sum a_i^2 2^(64 i): one add chain, one funnel shift per limb, and the squares - they tile
//      limbs (2i, 2i+1) - as one mad.lo.cc / madc.hi.cc chain whose addend is the shifted sum.
//   3. REDC on the low half S_lo alone (the running value stays below 2^(32(N+1))), then + S_hi: with
//      M = sum m_i 2^(32 i), (S_lo + M p) / R <= p and S_hi < p, so one final conditional subtraction.
//      A reduction-only row is a Montgomery step plus the re-alignment of the two arrays after the division by
//      2^32: the old odd-aligned array becomes the even-aligned one unchanged, the old even-aligned array minus
//      its limbs 0 (now zero) and 1 becomes the odd-aligned one unchanged, and limb 1 - alone at the new limb
//      0 - is carried into the next step (fe_mont_step_lone): no carry ripples between rows.
template <class P>
FB_HD Fe<P> fe_sqr(const Fe<P>& a) {
#ifdef FB_SQR_VIA_MUL  // A/B builds only
    return fe_mul<P>(a, a);
#endif
    constexpr int N = P::N;
    static_assert(N % 2 == 0, "even limb count expected");
    uint32_t De[2 * N + 2], Do[2 * N + 2];
#pragma unroll
    for (int i = 0; i < 2 * N + 2; i++) De[i] = Do[i] = 0;
    // ---- 1. off-diagonal products
#pragma unroll
    for (int i = 0; i < N - 1; i++) {
        // odd offsets j = i+1, i+3, ...: limbs 2i+1, ... of Do
        {
            bool first = true;
            int top = 0;
#pragma unroll
            for (int j = i + 1; j < N; j += 2) {
                const int k = i + j;
                Do[k] = first ? mad_lo_cc(a.v[i], a.v[j], Do[k]) : madc_lo_cc(a.v[i], a.v[j], Do[k]);
                Do[k + 1] = madc_hi_cc(a.v[i], a.v[j], Do[k + 1]);
                first = false;
                top = k + 1;
            }
            Do[top + 1] = addc(Do[top + 1], 0u);
        }
        // even offsets j = i+2, i+4, ...: limbs 2i+2, ... of De
        if (i + 2 < N) {
            bool first = true;
            int top = 0;
#pragma unroll
            for (int j = i + 2; j < N; j += 2) {
                const int k = i + j;
                De[k] = first ? mad_lo_cc(a.v[i], a.v[j], De[k]) : madc_lo_cc(a.v[i], a.v[j], De[k]);
                De[k + 1] = madc_hi_cc(a.v[i], a.v[j], De[k + 1]);
                first = false;
                top = k + 1;
            }
            De[top + 1] = addc(De[top + 1], 0u);
        }
    }
    // ---- 2. S = 2 (De + Do) + squares
    uint32_t T[2 * N];
    T[0] = 0;  // no product reaches limb 0
    T[1] = add_cc(De[1], Do[1]);
#pragma unroll
    for (int k = 2; k < 2 * N - 1; k++) T[k] = addc_cc(De[k], Do[k]);
    T[2 * N - 1] = addc(De[2 * N - 1], Do[2 * N - 1]);
    uint32_t S[2 * N];
#pragma unroll
    for (int k = 2 * N - 1; k >= 1; k--) T[k] = (T[k] << 1) | (T[k - 1] >> 31);  // D < 2^(64N-1): the top bit is free
#pragma unroll
    for (int i = 0; i < N; i++) {
        S[2 * i] = (i == 0) ? mad_lo_cc(a.v[i], a.v[i], T[2 * i]) : madc_lo_cc(a.v[i], a.v[i], T[2 * i]);
        S[2 * i + 1] = (i == N - 1) ? madc_hi(a.v[i], a.v[i], T[2 * i + 1]) : madc_hi_cc(a.v[i], a.v[i], T[2 * i + 1]);
    }
    // ---- 3. REDC of the low half
    uint32_t buf0[N], buf1[N];
    uint32_t* E = buf0;
    uint32_t* O = buf1;
#pragma unroll
    for (int k = 0; k < N; k++) {
        E[k] = S[k];
        O[k] = 0;
    }
    fe_mont_step<P>(E, O);
    O[N - 1] = addc(O[N - 1], 0);
#pragma unroll
    for (int i = 1; i < N; i++) {
        // V / 2^32: the old O is the new even-aligned array; the old E from limb 2 on is the new odd-aligned one;
        // the old E[1] stands alone at the new limb 0
        uint32_t* nE = O;
        uint32_t* nO = E;
        const uint32_t lone = E[1];
#pragma unroll
        for (int k = 0; k < N - 2; k++) nO[k] = E[k + 2];
        nO[N - 2] = 0;
        nO[N - 1] = 0;
        fe_mont_step_lone<P>(nE, nO, lone);
        nO[N - 1] = addc(nO[N - 1], 0);
        E = nE;
        O = nO;
    }
    // ---- (E >> 32) + O + S_hi
    uint32_t t[N];
    t[0] = add_cc(O[0], E[1]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) t[k] = addc_cc(O[k], E[k + 1]);
    t[N - 1] = addc(O[N - 1], 0);
    t[0] = add_cc(t[0], S[N]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) t[k] = addc_cc(t[k], S[N + k]);
    t[N - 1] = addc_cc(t[N - 1], S[2 * N - 1]);
    const uint32_t top = addc(0, 0);
    Fe<P> r;
    fe_final_sub<P>(r.v, t, top);
    return r;
}

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

// Low-latency inversion for ONE element on ONE thread (the per-CTA inversion of the affine rounds and of the opening's
// denominators): Kaliski's almost-Montgomery inverse, a binary extended GCD of
// <= 2*bits shift/subtract rounds on the limbs, instead of the ~1.5*bits dependent
// Montgomery multiplies of Fermat's a^(m-2).  Input and output in Montgomery form;
// inv(0) = 0.  Data-dependent branches: use fe_inv (branch-free) in warp-wide code.
template <class P>
FB_HD Fe<P> fe_inv_gcd(const Fe<P>& a) {
    constexpr int N = P::N;
    if (a.is_zero()) return a;
    uint32_t u[N], v[N], r[N], s[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
        u[i] = P::mod(i);
        v[i] = a.v[i];
        r[i] = 0;
        s[i] = 0;
    }
    s[0] = 1;
    int k = 0;
    auto shr1 = [](uint32_t* x) {
#pragma unroll
        for (int i = 0; i < N - 1; i++) x[i] = (x[i] >> 1) | (x[i + 1] << 31);
        x[N - 1] >>= 1;
    };
    auto shl1 = [](uint32_t* x) {
#pragma unroll
        for (int i = N - 1; i > 0; i--) x[i] = (x[i] << 1) | (x[i - 1] >> 31);
        x[0] <<= 1;
    };
    auto sub = [](uint32_t* x, const uint32_t* y) {  // x -= y
        x[0] = sub_cc(x[0], y[0]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) x[i] = subc_cc(x[i], y[i]);
        x[N - 1] = subc(x[N - 1], y[N - 1]);
    };
    auto add = [](uint32_t* x, const uint32_t* y) {  // x += y
        x[0] = add_cc(x[0], y[0]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) x[i] = addc_cc(x[i], y[i]);
        x[N - 1] = addc(x[N - 1], y[N - 1]);
    };
    for (;;) {
        uint32_t vz = 0;
#pragma unroll
        for (int i = 0; i < N; i++) vz |= v[i];
        if (!vz) break;
        if (!(u[0] & 1)) {
            shr1(u);
            shl1(s);
        } else if (!(v[0] & 1)) {
            shr1(v);
            shl1(r);
        } else if (!fe_geq<P>(v, u)) {  // u > v
            sub(u, v);
            shr1(u);
            add(r, s);
            shl1(s);
        } else {
            sub(v, u);
            shr1(v);
            add(s, r);
            shl1(r);
        }
        k++;
    }
    // r < 2m;  m - r = a^-1 * 2^k (mod m),  bits <= k <= 2*bits
    Fe<P> x;
    {
        uint32_t t[N];
        fe_final_sub<P>(t, r, 0);
        x.v[0] = sub_cc(P::mod(0), t[0]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) x.v[i] = subc_cc(P::mod(i), t[i]);
        x.v[N - 1] = subc(P::mod(N - 1), t[N - 1]);
        uint32_t tmp[N];
        fe_final_sub<P>(tmp, x.v, 0);  // r == 0 cannot happen for a != 0, but keep x < m
#pragma unroll
        for (int i = 0; i < N; i++) x.v[i] = tmp[i];
    }
    // input was A = a*R: x = A^-1 * 2^k.  Want A^-1 * R^2 = mont(mont(x, R^2), 2^(2*32N - k))
    x = fe_mul<P>(x, Fe<P>::rsquared());
    int e = 64 * N - k;  // >= 64N - 2*bits > 0
    int safe = 32 * N - 4;  // 2^safe < m for both fields (m > 2^(32N-3))
    int rem = e > safe ? e - safe : 0;
    e -= rem;
    Fe<P> b = Fe<P>::zero();
    b.v[e >> 5] = 1u << (e & 31);
    x = fe_mul<P>(x, b);
    for (int i = 0; i < rem; i++) x = fe_dbl<P>(x);
    return x;
}

// Inversion by the Bernstein-Yang division steps ("safegcd", ia.cr/2019/266), in the batched form libsecp256k1's
// modinv32 popularised: 30 division steps at a time on the low words of f and g build a 2x2 transition matrix, which is then
// applied once to the full-width f, g (exact division by 2^30) and to d, e (division by 2^30 modulo m).  Signed 30-bit
// limbs; the outer loop runs until g == 0 (data-dependent; the inputs here are not secret).  Input and output in
// Montgomery form; inv(0) = 0.
// Measured against fe_inv_gcd (profiles/r2_v7_ab_inversion.txt): about the same instruction count, but on the wide-MAD pipe
// instead of the ALU.  Alone on the GPU it is 1.4x faster (XYZZ -> affine of a result: 71 -> 27 us); under a concurrent
// multiplier-bound kernel it is slower, which is where the per-CTA inversions of the affine rounds run.  So: this one for
// the end-of-call steps (k_to_affine, k_combine_parts), fe_inv_gcd inside the rounds and the opening.
template <class P>
FB_HD Fe<P> fe_inv_safegcd(const Fe<P>& a) {
    constexpr int N = P::N;
    constexpr int L = (32 * N + 29) / 30 + (((32 * N) % 30 == 0) ? 1 : 0);  // limbs of 30 bits incl. room for the sign: 13 (Fp), 9 (Fr)
    constexpr int32_t M30 = (int32_t)(0xffffffffu >> 2);
    if (a.is_zero()) return a;
    auto to30 = [](const uint32_t* x, int32_t* out) {
#pragma unroll
        for (int j = 0; j < L; j++) {
            const int bit = 30 * j, w = bit >> 5, off = bit & 31;
            uint32_t lo = w < N ? x[w] : 0u;
            uint32_t hi = (w + 1 < N) ? x[w + 1] : 0u;
            uint32_t v = off ? ((lo >> off) | (hi << (32 - off))) : lo;
            out[j] = (int32_t)(v & (uint32_t)M30);
        }
    };
    int32_t m[L], f[L], g[L], d[L], e[L];
    {
        uint32_t mw[N];
#pragma unroll
        for (int i = 0; i < N; i++) mw[i] = P::mod(i);
        to30(mw, m);
        to30(a.v, g);
    }
#pragma unroll
    for (int i = 0; i < L; i++) {
        f[i] = m[i];
        d[i] = 0;
        e[i] = 0;
    }
    e[0] = 1;
    const uint32_t minv30 = (0u - P::INV) & (uint32_t)M30;  // m^-1 mod 2^30
    int32_t zeta = -1;  // -(delta + 1/2), delta = 1/2
    for (int iter = 0; iter < 64; iter++) {  // <= 30 batches for 384-bit inputs; the bound is a backstop
        // 30 division steps on the low words; [[u, v], [q, r]] / 2^30 maps (f, g) to the new (f, g)
        uint32_t u = 1, v = 0, q = 0, r = 1;
        {
            uint32_t fl = (uint32_t)f[0] | ((uint32_t)f[1] << 30), gl = (uint32_t)g[0] | ((uint32_t)g[1] << 30);
#pragma unroll 6
            for (int i = 0; i < 30; i++) {
                const uint32_t c1 = (uint32_t)(zeta >> 31);
                const uint32_t c2 = 0u - (gl & 1u);
                const uint32_t x = (fl ^ c1) - c1, y = (u ^ c1) - c1, z = (v ^ c1) - c1;
                gl += x & c2;
                q += y & c2;
                r += z & c2;
                const uint32_t c3 = c1 & c2;
                zeta = (int32_t)((uint32_t)zeta ^ c3) - 1;
                fl += gl & c3;
                u += q & c3;
                v += r & c3;
                gl >>= 1;
                u <<= 1;
                v <<= 1;
            }
        }
        const int32_t su = (int32_t)u, sv = (int32_t)v, sq = (int32_t)q, sr = (int32_t)r;
        // (d, e) <- [[u, v], [q, r]] (d, e) / 2^30 mod m, kept in (-2m, m)
        {
            const int32_t sd = d[L - 1] >> 31, se = e[L - 1] >> 31;
            int32_t md = (su & sd) + (sv & se);
            int32_t me = (sq & sd) + (sr & se);
            long long cd = (long long)su * d[0] + (long long)sv * e[0];
            long long ce = (long long)sq * d[0] + (long long)sr * e[0];
            md -= (int32_t)((minv30 * (uint32_t)cd + (uint32_t)md) & (uint32_t)M30);
            me -= (int32_t)((minv30 * (uint32_t)ce + (uint32_t)me) & (uint32_t)M30);
            cd += (long long)m[0] * md;
            ce += (long long)m[0] * me;
            cd >>= 30;
            ce >>= 30;
#pragma unroll
            for (int i = 1; i < L; i++) {
                cd += (long long)su * d[i] + (long long)sv * e[i] + (long long)m[i] * md;
                ce += (long long)sq * d[i] + (long long)sr * e[i] + (long long)m[i] * me;
                d[i - 1] = (int32_t)cd & M30;
                e[i - 1] = (int32_t)ce & M30;
                cd >>= 30;
                ce >>= 30;
            }
            d[L - 1] = (int32_t)cd;
            e[L - 1] = (int32_t)ce;
        }
        // (f, g) <- [[u, v], [q, r]] (f, g) / 2^30, exactly
        {
            long long cf = (long long)su * f[0] + (long long)sv * g[0];
            long long cg = (long long)sq * f[0] + (long long)sr * g[0];
            cf >>= 30;
            cg >>= 30;
#pragma unroll
            for (int i = 1; i < L; i++) {
                cf += (long long)su * f[i] + (long long)sv * g[i];
                cg += (long long)sq * f[i] + (long long)sr * g[i];
                f[i - 1] = (int32_t)cf & M30;
                g[i - 1] = (int32_t)cg & M30;
                cf >>= 30;
                cg >>= 30;
            }
            f[L - 1] = (int32_t)cf;
            g[L - 1] = (int32_t)cg;
        }
        int32_t gz = 0;
#pragma unroll
        for (int i = 0; i < L; i++) gz |= g[i];
        if (gz == 0) break;
    }
    // f = +-1 and d = +- 1/a: negate by the sign of f and bring d from (-2m, m) to [0, m)
    {
        const int32_t neg = f[L - 1] >> 31;
        int32_t add = d[L - 1] >> 31;
#pragma unroll
        for (int i = 0; i < L; i++) d[i] = ((d[i] + (m[i] & add)) ^ neg) - neg;
#pragma unroll
        for (int i = 0; i < L - 1; i++) {
            d[i + 1] += d[i] >> 30;
            d[i] &= M30;
        }
        add = d[L - 1] >> 31;
#pragma unroll
        for (int i = 0; i < L; i++) d[i] += m[i] & add;
#pragma unroll
        for (int i = 0; i < L - 1; i++) {
            d[i + 1] += d[i] >> 30;
            d[i] &= M30;
        }
    }
    Fe<P> x;
#pragma unroll
    for (int i = 0; i < N; i++) {  // word i = bits [32 i, 32 i + 32) of sum d[j] 2^(30 j)
        const int bit = 32 * i, j = bit / 30, off = bit % 30;
        uint32_t w = (uint32_t)d[j] >> off;
        if (j + 1 < L) w |= (uint32_t)d[j + 1] << (30 - off);
        if (off > 28 && j + 2 < L) w |= (uint32_t)d[j + 2] << (60 - off);
        x.v[i] = w;
    }
    // the input was A = a R (as an integer), x = 1 / A = R^-1 / a; the Montgomery form of 1 / a is x R^2: two products with R^2
    x = fe_mul<P>(x, Fe<P>::rsquared());
    return fe_mul<P>(x, Fe<P>::rsquared());
}

using Fr = Fe<FrParams>;
using Fp = Fe<FpParams>;

}  // namespace fb
