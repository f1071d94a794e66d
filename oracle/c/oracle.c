/* oracle.c - CPU ORACLE (C restatement) of fourier's KZG / Pianist hot path.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs, never by the product (fourier_b200/).
 *
 * What it restates.  The reference (apollozkp/fourier) holds no arithmetic: the
 * path runs in rust-kzg-blst @abb71933 + blst 0.3.11 (reference Cargo.toml:11-12,
 * 29; Cargo.lock:139-142, 767-769, 1247-1249), neither vendored nor buildable
 * here (no Rust toolchain).  This file restates the published algorithms those
 * crates implement, on 64-bit limbs with unsigned __int128 (an independent code
 * base from the 32-bit-limb XYZZ CUDA kernels it checks):
 *   - Montgomery Fr / Fp                        (blst_fr_*, blst_fp_*)
 *   - G1 Jacobian add / double, mixed add      (blst POINTonE1)
 *   - compress / uncompress, ZCash encoding    (blst_p1_compress / uncompress)
 *   - Pippenger MSM, multi-threaded             (g1_linear_combination; call sites
 *                                                src/engine/piano.rs:150, 228, 287)
 *   - radix-2 FFT over Fr, natural order        (fft_fr; piano.rs:1096-1103)
 *   - Horner and division by (X - a)            (FsPoly::eval / div; piano.rs:202, 215)
 *   - worker_open's scalar pipeline             (piano.rs:193-233)
 * Parity status: pinned against the Python big-int oracle (oracle/*.py), which is
 * itself pinned on the reference's only numeric KAT (src/engine/blst.rs:304-324),
 * the well-known generator encoding and the algebraic identities the reference's
 * tests assert.  No golden G1 bytes exist in the reference: "parity unpinned"
 * against blst's own output bytes.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
typedef uint64_t u64;

/* ------------------------------------------------------------------ fields */
#define FRN 4
#define FPN 6
static const u64 FR_MOD[FRN] = {0xffffffff00000001ULL, 0x53bda402fffe5bfeULL, 0x3339d80809a1d805ULL, 0x73eda753299d7d48ULL};
static const u64 FR_INV = 0xfffffffeffffffffULL;
static const u64 FR_ONE[FRN] = {0x00000001fffffffeULL, 0x5884b7fa00034802ULL, 0x998c4fefecbc4ff5ULL, 0x1824b159acc5056fULL};
static const u64 FR_R2[FRN] = {0xc999e990f3f29c6dULL, 0x2b6cedcb87925c23ULL, 0x05d314967254398fULL, 0x0748d9d99f59ff11ULL};
static const u64 FP_MOD[FPN] = {0xb9feffffffffaaabULL, 0x1eabfffeb153ffffULL, 0x6730d2a0f6b0f624ULL,
                                0x64774b84f38512bfULL, 0x4b1ba7b6434bacd7ULL, 0x1a0111ea397fe69aULL};
static const u64 FP_INV = 0x89f3fffcfffcfffdULL;
static const u64 FP_ONE[FPN] = {0x760900000002fffdULL, 0xebf4000bc40c0002ULL, 0x5f48985753c758baULL,
                                0x77ce585370525745ULL, 0x5c071a97a256ec6dULL, 0x15f65ec3fa80e493ULL};
static const u64 FP_R2[FPN] = {0xf4df1f341c341746ULL, 0x0a76e6a609d104f1ULL, 0x8de5476c4c95b6d5ULL,
                               0x67eb88a9939d83c0ULL, 0x9a793e85b519952dULL, 0x11988fe592cae3aaULL};
static const u64 FP_HALF[FPN] = {0xdcff7fffffffd555ULL, 0x0f55ffff58a9ffffULL, 0xb39869507b587b12ULL,
                                 0xb23ba5c279c2895fULL, 0x258dd3db21a5d66bULL, 0x0d0088f51cbff34dULL};
static const u64 FP_SQRT_E[FPN] = {0xee7fbfffffffeaabULL, 0x07aaffffac54ffffULL, 0xd9cc34a83dac3d89ULL,
                                   0xd91dd2e13ce144afULL, 0x92c6e9ed90d2eb35ULL, 0x0680447a8e5ff9a6ULL};
static const u64 G1X[FPN] = {0xfb3af00adb22c6bbULL, 0x6c55e83ff97a1aefULL, 0xa14e3a3f171bac58ULL,
                             0xc3688c4f9774b905ULL, 0x2695638c4fa9ac0fULL, 0x17f1d3a73197d794ULL};
static const u64 G1Y[FPN] = {0x0caa232946c5e7e1ULL, 0xd03cc744a2888ae4ULL, 0x00db18cb2c04b3edULL,
                             0xfcf5e095d5d00af6ULL, 0xa09e30ed741d8ae4ULL, 0x08b3f481e3aaa0f1ULL};
static const u64 FR_W32[FRN] = {0x3829971f439f0d2bULL, 0xb63683508c2280b9ULL, 0xd09b681922c813b4ULL, 0x16a2a19edfe81f20ULL};

/* generic n-limb helpers (n is a compile-time constant at every call site) */
static inline int ge_n(const u64* a, const u64* b, int n) {
    for (int i = n - 1; i >= 0; i--) {
        if (a[i] > b[i]) return 1;
        if (a[i] < b[i]) return 0;
    }
    return 1;
}
static inline u64 add_n(u64* r, const u64* a, const u64* b, int n) {
    u128 c = 0;
    for (int i = 0; i < n; i++) {
        c += (u128)a[i] + b[i];
        r[i] = (u64)c;
        c >>= 64;
    }
    return (u64)c;
}
static inline u64 sub_n(u64* r, const u64* a, const u64* b, int n) {
    u64 br = 0;
    for (int i = 0; i < n; i++) {
        u128 d = (u128)a[i] - b[i] - br;
        r[i] = (u64)d;
        br = (u64)(d >> 64) & 1;
    }
    return br;
}
static inline int is_zero_n(const u64* a, int n) {
    u64 o = 0;
    for (int i = 0; i < n; i++) o |= a[i];
    return o == 0;
}
static inline void mod_add(u64* r, const u64* a, const u64* b, const u64* m, int n) {
    u64 t[FPN], s[FPN];
    u64 c = add_n(t, a, b, n);
    u64 br = sub_n(s, t, m, n);
    const u64* src = (c || !br) ? s : t;
    for (int i = 0; i < n; i++) r[i] = src[i];
}
static inline void mod_sub(u64* r, const u64* a, const u64* b, const u64* m, int n) {
    u64 t[FPN];
    u64 br = sub_n(t, a, b, n);
    if (br) add_n(t, t, m, n);
    for (int i = 0; i < n; i++) r[i] = t[i];
}
/* Montgomery product, coarsely integrated operand scanning; one fully unrolled
 * instance per limb count (the generic-n form was 5x slower) */
#define DEFINE_MONT_MUL(NAME, N)                                                              \
    static inline __attribute__((always_inline)) void NAME(u64* r, const u64* a, const u64* b, \
                                                           const u64* m, u64 inv) {            \
        u64 t[N + 2];                                                                         \
        _Pragma("GCC unroll 8") for (int i = 0; i < N + 2; i++) t[i] = 0;                     \
        _Pragma("GCC unroll 8") for (int i = 0; i < N; i++) {                                 \
            u128 c = 0;                                                                       \
            _Pragma("GCC unroll 8") for (int j = 0; j < N; j++) {                             \
                c += (u128)a[j] * b[i] + t[j];                                                \
                t[j] = (u64)c;                                                                \
                c >>= 64;                                                                     \
            }                                                                                 \
            c += t[N];                                                                        \
            t[N] = (u64)c;                                                                    \
            t[N + 1] = (u64)(c >> 64);                                                        \
            u64 q = t[0] * inv;                                                               \
            c = (u128)q * m[0] + t[0];                                                        \
            c >>= 64;                                                                         \
            _Pragma("GCC unroll 8") for (int j = 1; j < N; j++) {                             \
                c += (u128)q * m[j] + t[j];                                                   \
                t[j - 1] = (u64)c;                                                            \
                c >>= 64;                                                                     \
            }                                                                                 \
            c += t[N];                                                                        \
            t[N - 1] = (u64)c;                                                                \
            t[N] = t[N + 1] + (u64)(c >> 64);                                                 \
        }                                                                                     \
        u64 s[N];                                                                             \
        u64 br = 0;                                                                           \
        _Pragma("GCC unroll 8") for (int i = 0; i < N; i++) {                                 \
            u128 d = (u128)t[i] - m[i] - br;                                                  \
            s[i] = (u64)d;                                                                    \
            br = (u64)(d >> 64) & 1;                                                          \
        }                                                                                     \
        u64 keep = (t[N] == 0 && br) ? ~(u64)0 : 0;                                           \
        _Pragma("GCC unroll 8") for (int i = 0; i < N; i++) r[i] = (t[i] & keep) | (s[i] & ~keep); \
    }
DEFINE_MONT_MUL(mont_mul4, 4)
DEFINE_MONT_MUL(mont_mul6, 6)
#if defined(__x86_64__) && defined(__BMI2__) && defined(__ADX__)
/* t[0..5] += x[0..3] * y with the two carry chains of mulx/adcx/adox */
static inline __attribute__((always_inline)) void mac_row4(u64* t, const u64* x, u64 y) {
    u64 lo, hi, z;
    __asm__(
        "xorq %[z], %[z]\n\t"
        "mulxq 0(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t0]\n\t adoxq %[hi], %[t1]\n\t"
        "mulxq 8(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t1]\n\t adoxq %[hi], %[t2]\n\t"
        "mulxq 16(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t2]\n\t adoxq %[hi], %[t3]\n\t"
        "mulxq 24(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t3]\n\t adoxq %[hi], %[t4]\n\t"
        "adcxq %[z], %[t4]\n\t adoxq %[z], %[t5]\n\t adcxq %[z], %[t5]\n\t"
        : [t0] "+r"(t[0]), [t1] "+r"(t[1]), [t2] "+r"(t[2]), [t3] "+r"(t[3]), [t4] "+r"(t[4]), [t5] "+r"(t[5]), [lo] "=&r"(lo), [hi] "=&r"(hi), [z] "=&r"(z)
        : [x] "r"(x), "d"(y), "m"(*(const u64(*)[4])x)
        : "cc");
}
static inline __attribute__((always_inline)) void mont_mul4x(u64* r, const u64* a, const u64* b, const u64* m, u64 inv) {
    u64 t[6] = {0};
    for (int i = 0; i < 4; i++) {
        mac_row4(t, a, b[i]);
        mac_row4(t, m, t[0] * inv);
        for (int k = 0; k < 5; k++) t[k] = t[k + 1];
        t[5] = 0;
    }
    u64 s[4];
    u64 br = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)t[i] - m[i] - br;
        s[i] = (u64)d;
        br = (u64)(d >> 64) & 1;
    }
    u64 keep = (t[4] == 0 && br) ? ~(u64)0 : 0;
    for (int i = 0; i < 4; i++) r[i] = (t[i] & keep) | (s[i] & ~keep);
}
#else
#define mont_mul4x mont_mul4
#endif
#if defined(__x86_64__) && defined(__BMI2__) && defined(__ADX__)
/* t[0..7] += x[0..5] * y with the two carry chains of mulx/adcx/adox */
static inline __attribute__((always_inline)) void mac_row6(u64* t, const u64* x, u64 y) {
    u64 lo, hi, z;
    __asm__(
        "xorq %[z], %[z]\n\t"
        "mulxq 0(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t0]\n\t adoxq %[hi], %[t1]\n\t"
        "mulxq 8(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t1]\n\t adoxq %[hi], %[t2]\n\t"
        "mulxq 16(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t2]\n\t adoxq %[hi], %[t3]\n\t"
        "mulxq 24(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t3]\n\t adoxq %[hi], %[t4]\n\t"
        "mulxq 32(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t4]\n\t adoxq %[hi], %[t5]\n\t"
        "mulxq 40(%[x]), %[lo], %[hi]\n\t adcxq %[lo], %[t5]\n\t adoxq %[hi], %[t6]\n\t"
        "adcxq %[z], %[t6]\n\t adoxq %[z], %[t7]\n\t adcxq %[z], %[t7]\n\t"
        : [t0] "+r"(t[0]), [t1] "+r"(t[1]), [t2] "+r"(t[2]), [t3] "+r"(t[3]), [t4] "+r"(t[4]), [t5] "+r"(t[5]), [t6] "+r"(t[6]), [t7] "+r"(t[7]), [lo] "=&r"(lo), [hi] "=&r"(hi), [z] "=&r"(z)
        : [x] "r"(x), "d"(y), "m"(*(const u64(*)[6])x)
        : "cc");
}
static inline __attribute__((always_inline)) void mont_mul6x(u64* r, const u64* a, const u64* b, const u64* m, u64 inv) {
    u64 t[8] = {0};
    for (int i = 0; i < 6; i++) {
        mac_row6(t, a, b[i]);
        mac_row6(t, m, t[0] * inv);
        for (int k = 0; k < 7; k++) t[k] = t[k + 1];
        t[7] = 0;
    }
    u64 s[6];
    u64 br = 0;
    for (int i = 0; i < 6; i++) {
        u128 d = (u128)t[i] - m[i] - br;
        s[i] = (u64)d;
        br = (u64)(d >> 64) & 1;
    }
    u64 keep = (t[6] == 0 && br) ? ~(u64)0 : 0;
    for (int i = 0; i < 6; i++) r[i] = (t[i] & keep) | (s[i] & ~keep);
}
#else
#define mont_mul6x mont_mul6
#endif

typedef struct { u64 v[FRN]; } fr_t;
typedef struct { u64 v[FPN]; } fp_t;

static inline void fr_mul(fr_t* r, const fr_t* a, const fr_t* b) { mont_mul4x(r->v, a->v, b->v, FR_MOD, FR_INV); }
static inline void fr_add(fr_t* r, const fr_t* a, const fr_t* b) { mod_add(r->v, a->v, b->v, FR_MOD, FRN); }
static inline void fr_sub(fr_t* r, const fr_t* a, const fr_t* b) { mod_sub(r->v, a->v, b->v, FR_MOD, FRN); }
static inline void fp_mul(fp_t* r, const fp_t* a, const fp_t* b) { mont_mul6x(r->v, a->v, b->v, FP_MOD, FP_INV); }
static inline void fp_sqr(fp_t* r, const fp_t* a) { mont_mul6x(r->v, a->v, a->v, FP_MOD, FP_INV); }
static inline void fp_add(fp_t* r, const fp_t* a, const fp_t* b) { mod_add(r->v, a->v, b->v, FP_MOD, FPN); }
static inline void fp_sub(fp_t* r, const fp_t* a, const fp_t* b) { mod_sub(r->v, a->v, b->v, FP_MOD, FPN); }
static inline int fp_is_zero(const fp_t* a) { return is_zero_n(a->v, FPN); }
static inline int fp_eq(const fp_t* a, const fp_t* b) { return memcmp(a->v, b->v, sizeof a->v) == 0; }
static inline void fp_neg(fp_t* r, const fp_t* a) {
    if (fp_is_zero(a)) { *r = *a; return; }
    sub_n(r->v, FP_MOD, a->v, FPN);
}
static void fr_set_one(fr_t* r) { memcpy(r->v, FR_ONE, sizeof r->v); }
static void fp_set_one(fp_t* r) { memcpy(r->v, FP_ONE, sizeof r->v); }

static void fr_pow(fr_t* r, const fr_t* a, const u64* e, int en) {
    fr_t acc;
    fr_set_one(&acc);
    for (int i = en - 1; i >= 0; i--)
        for (int b = 63; b >= 0; b--) {
            fr_mul(&acc, &acc, &acc);
            if ((e[i] >> b) & 1) fr_mul(&acc, &acc, a);
        }
    *r = acc;
}
static void fp_pow(fp_t* r, const fp_t* a, const u64* e, int en) {
    fp_t acc;
    fp_set_one(&acc);
    for (int i = en - 1; i >= 0; i--)
        for (int b = 63; b >= 0; b--) {
            fp_sqr(&acc, &acc);
            if ((e[i] >> b) & 1) fp_mul(&acc, &acc, a);
        }
    *r = acc;
}
static void fr_inv(fr_t* r, const fr_t* a) {
    u64 e[FRN];
    u64 two[FRN] = {2, 0, 0, 0};
    sub_n(e, FR_MOD, two, FRN);
    fr_pow(r, a, e, FRN);
}
static void fp_inv(fp_t* r, const fp_t* a) {
    u64 e[FPN];
    u64 two[FPN] = {2, 0, 0, 0, 0, 0};
    sub_n(e, FP_MOD, two, FPN);
    fp_pow(r, a, e, FPN);
}

/* FsFr::from_bytes: 32 bytes big-endian, canonical (piano.rs:62, 70); -> Montgomery.
 * returns 0 ok, 1 "Invalid scalar" */
static int fr_from_be(fr_t* r, const uint8_t* b) {
    fr_t x;
    for (int i = 0; i < FRN; i++) {
        u64 w = 0;
        for (int k = 0; k < 8; k++) w = (w << 8) | b[8 * (FRN - 1 - i) + k];
        x.v[i] = w;
    }
    if (ge_n(x.v, FR_MOD, FRN)) return 1;
    fr_t r2;
    memcpy(r2.v, FR_R2, sizeof r2.v);
    fr_mul(r, &x, &r2);
    return 0;
}
static void fr_canon(u64* out, const fr_t* a) {
    fr_t one = {{1, 0, 0, 0}}, t;
    fr_mul(&t, a, &one);
    memcpy(out, t.v, sizeof t.v);
}
static void fr_to_be(uint8_t* b, const fr_t* a) {
    u64 c[FRN];
    fr_canon(c, a);
    for (int i = 0; i < FRN; i++)
        for (int k = 0; k < 8; k++) b[8 * (FRN - 1 - i) + k] = (uint8_t)(c[i] >> (56 - 8 * k));
}
static void fp_from_be_raw(fp_t* x, const uint8_t* b) {
    for (int i = 0; i < FPN; i++) {
        u64 w = 0;
        for (int k = 0; k < 8; k++) w = (w << 8) | b[8 * (FPN - 1 - i) + k];
        x->v[i] = w;
    }
}
static void fp_to_mont(fp_t* r, const fp_t* x) {
    fp_t r2;
    memcpy(r2.v, FP_R2, sizeof r2.v);
    fp_mul(r, x, &r2);
}
static void fp_canon(u64* out, const fp_t* a) {
    fp_t one = {{1, 0, 0, 0, 0, 0}}, t;
    fp_mul(&t, a, &one);
    memcpy(out, t.v, sizeof t.v);
}
static void fp_to_be(uint8_t* b, const fp_t* a) {
    u64 c[FPN];
    fp_canon(c, a);
    for (int i = 0; i < FPN; i++)
        for (int k = 0; k < 8; k++) b[8 * (FPN - 1 - i) + k] = (uint8_t)(c[i] >> (56 - 8 * k));
}

/* ---------------------------------------------------------------------- G1 */
typedef struct { fp_t x, y; int inf; } g1a_t;      /* affine */
typedef struct { fp_t x, y, z; } g1j_t;            /* Jacobian, z == 0: identity */

static void g1j_set_inf(g1j_t* p) { memset(p, 0, sizeof *p); }
static int g1j_is_inf(const g1j_t* p) { return fp_is_zero(&p->z); }

/* dbl-2009-l (a = 0) */
static void g1j_dbl(g1j_t* r, const g1j_t* p) {
    if (g1j_is_inf(p)) { *r = *p; return; }
    fp_t A, B, C, D, E, F, t;
    fp_sqr(&A, &p->x);
    fp_sqr(&B, &p->y);
    fp_sqr(&C, &B);
    fp_add(&t, &p->x, &B);
    fp_sqr(&t, &t);
    fp_sub(&t, &t, &A);
    fp_sub(&t, &t, &C);
    fp_add(&D, &t, &t);
    fp_add(&E, &A, &A);
    fp_add(&E, &E, &A);
    fp_sqr(&F, &E);
    g1j_t o;
    fp_sub(&o.x, &F, &D);
    fp_sub(&o.x, &o.x, &D);
    fp_sub(&t, &D, &o.x);
    fp_mul(&o.y, &E, &t);
    fp_add(&C, &C, &C);
    fp_add(&C, &C, &C);
    fp_add(&C, &C, &C);
    fp_sub(&o.y, &o.y, &C);
    fp_mul(&o.z, &p->y, &p->z);
    fp_add(&o.z, &o.z, &o.z);
    *r = o;
}

/* add-2007-bl, complete by case analysis */
static void g1j_add(g1j_t* r, const g1j_t* p, const g1j_t* q) {
    if (g1j_is_inf(p)) { *r = *q; return; }
    if (g1j_is_inf(q)) { *r = *p; return; }
    fp_t z1z1, z2z2, u1, u2, s1, s2, h, rr, hh, hhh, v, t;
    fp_sqr(&z1z1, &p->z);
    fp_sqr(&z2z2, &q->z);
    fp_mul(&u1, &p->x, &z2z2);
    fp_mul(&u2, &q->x, &z1z1);
    fp_mul(&s1, &p->y, &q->z);
    fp_mul(&s1, &s1, &z2z2);
    fp_mul(&s2, &q->y, &p->z);
    fp_mul(&s2, &s2, &z1z1);
    fp_sub(&h, &u2, &u1);
    fp_sub(&rr, &s2, &s1);
    if (fp_is_zero(&h)) {
        if (fp_is_zero(&rr)) g1j_dbl(r, p);
        else g1j_set_inf(r);
        return;
    }
    fp_sqr(&hh, &h);
    fp_mul(&hhh, &h, &hh);
    fp_mul(&v, &u1, &hh);
    g1j_t o;
    fp_sqr(&o.x, &rr);
    fp_sub(&o.x, &o.x, &hhh);
    fp_sub(&o.x, &o.x, &v);
    fp_sub(&o.x, &o.x, &v);
    fp_sub(&t, &v, &o.x);
    fp_mul(&o.y, &rr, &t);
    fp_mul(&t, &s1, &hhh);
    fp_sub(&o.y, &o.y, &t);
    fp_mul(&o.z, &p->z, &q->z);
    fp_mul(&o.z, &o.z, &h);
    *r = o;
}

/* mixed addition (q affine) */
static void g1j_add_affine(g1j_t* r, const g1j_t* p, const g1a_t* q) {
    if (q->inf) { *r = *p; return; }
    if (g1j_is_inf(p)) {
        r->x = q->x;
        r->y = q->y;
        fp_set_one(&r->z);
        return;
    }
    fp_t z1z1, u2, s2, h, rr, hh, hhh, v, t;
    fp_sqr(&z1z1, &p->z);
    fp_mul(&u2, &q->x, &z1z1);
    fp_mul(&s2, &q->y, &p->z);
    fp_mul(&s2, &s2, &z1z1);
    fp_sub(&h, &u2, &p->x);
    fp_sub(&rr, &s2, &p->y);
    if (fp_is_zero(&h)) {
        if (fp_is_zero(&rr)) g1j_dbl(r, p);
        else g1j_set_inf(r);
        return;
    }
    fp_sqr(&hh, &h);
    fp_mul(&hhh, &h, &hh);
    fp_mul(&v, &p->x, &hh);
    g1j_t o;
    fp_sqr(&o.x, &rr);
    fp_sub(&o.x, &o.x, &hhh);
    fp_sub(&o.x, &o.x, &v);
    fp_sub(&o.x, &o.x, &v);
    fp_sub(&t, &v, &o.x);
    fp_mul(&o.y, &rr, &t);
    fp_mul(&t, &p->y, &hhh);
    fp_sub(&o.y, &o.y, &t);
    fp_mul(&o.z, &p->z, &h);
    *r = o;
}

static void g1j_to_affine(g1a_t* r, const g1j_t* p) {
    if (g1j_is_inf(p)) { memset(r, 0, sizeof *r); r->inf = 1; return; }
    fp_t zi, zi2, zi3;
    fp_inv(&zi, &p->z);
    fp_sqr(&zi2, &zi);
    fp_mul(&zi3, &zi2, &zi);
    fp_mul(&r->x, &p->x, &zi2);
    fp_mul(&r->y, &p->y, &zi3);
    r->inf = 0;
}

static int fp_lex_larger(const fp_t* y) {
    u64 c[FPN];
    fp_canon(c, y);
    for (int i = FPN - 1; i >= 0; i--) {
        if (c[i] > FP_HALF[i]) return 1;
        if (c[i] < FP_HALF[i]) return 0;
    }
    return 0;
}

/* blst_p1_compress (reference src/rpc.rs:241, 268-269, 311, 330) */
static void g1a_compress(uint8_t out[48], const g1a_t* p) {
    if (p->inf) { memset(out, 0, 48); out[0] = 0xC0; return; }
    fp_to_be(out, &p->x);
    out[0] |= 0x80;
    if (fp_lex_larger(&p->y)) out[0] |= 0x20;
}
/* blst_p1_serialize (piano.rs:655) */
static void g1a_serialize(uint8_t out[96], const g1a_t* p) {
    if (p->inf) { memset(out, 0, 96); out[0] = 0x40; return; }
    fp_to_be(out, &p->x);
    fp_to_be(out + 48, &p->y);
}
static int g1a_on_curve(const g1a_t* p) {
    if (p->inf) return 1;
    fp_t l, r, four;
    fp_sqr(&l, &p->y);
    fp_sqr(&r, &p->x);
    fp_mul(&r, &r, &p->x);
    fp_set_one(&four);
    fp_add(&four, &four, &four);
    fp_add(&four, &four, &four);
    fp_add(&r, &r, &four);
    return fp_eq(&l, &r);
}
/* blst_p1_deserialize (piano.rs:728, 770); 0 ok */
static int g1a_deserialize(g1a_t* r, const uint8_t in[96]) {
    if (in[0] & 0x80) return 1;
    if (in[0] & 0x40) { memset(r, 0, sizeof *r); r->inf = 1; return 0; }
    fp_t x, y;
    fp_from_be_raw(&x, in);
    fp_from_be_raw(&y, in + 48);
    if (ge_n(x.v, FP_MOD, FPN) || ge_n(y.v, FP_MOD, FPN)) return 3;
    fp_to_mont(&r->x, &x);
    fp_to_mont(&r->y, &y);
    r->inf = 0;
    return g1a_on_curve(r) ? 0 : 4;
}
/* blst_p1_uncompress (piano.rs:57, 724, 759); 0 ok */
static int g1a_uncompress(g1a_t* r, const uint8_t in[48]) {
    uint8_t f = in[0], tmp[48];
    if (!(f & 0x80)) return 1;
    memcpy(tmp, in, 48);
    tmp[0] &= 0x1f;
    fp_t x;
    fp_from_be_raw(&x, tmp);
    if (f & 0x40) {
        if (!fp_is_zero(&x) || (f & 0x20)) return 2;
        memset(r, 0, sizeof *r);
        r->inf = 1;
        return 0;
    }
    if (ge_n(x.v, FP_MOD, FPN)) return 3;
    fp_t xm, rhs, four, y, y2;
    fp_to_mont(&xm, &x);
    fp_sqr(&rhs, &xm);
    fp_mul(&rhs, &rhs, &xm);
    fp_set_one(&four);
    fp_add(&four, &four, &four);
    fp_add(&four, &four, &four);
    fp_add(&rhs, &rhs, &four);
    fp_pow(&y, &rhs, FP_SQRT_E, FPN);
    fp_sqr(&y2, &y);
    if (!fp_eq(&y2, &rhs)) return 4;
    if (fp_lex_larger(&y) != ((f & 0x20) != 0)) fp_neg(&y, &y);
    r->x = xm;
    r->y = y;
    r->inf = 0;
    return 0;
}
static void g1_generator(g1a_t* g) {
    fp_t x, y;
    memcpy(x.v, G1X, sizeof x.v);
    memcpy(y.v, G1Y, sizeof y.v);
    fp_to_mont(&g->x, &x);
    fp_to_mont(&g->y, &y);
    g->inf = 0;
}
/* double-and-add with a canonical 4-limb scalar */
static void g1_mul(g1j_t* r, const g1a_t* p, const u64* k) {
    g1j_t acc;
    g1j_set_inf(&acc);
    for (int i = FRN - 1; i >= 0; i--)
        for (int b = 63; b >= 0; b--) {
            g1j_dbl(&acc, &acc);
            if ((k[i] >> b) & 1) g1j_add_affine(&acc, &acc, p);
        }
    *r = acc;
}

/* --------------------------------------------------------------------- MSM */
static int pick_window(size_t n) {
    int lg = 0;
    while (((size_t)1 << (lg + 1)) <= n) lg++;
    int c = lg > 4 ? lg - 3 : 2;
    if (c > 16) c = 16;
    return c;
}

/* signed c-bit digit w of canonical scalar k (4 limbs); carries propagate from
 * below, so digits are produced in order into out[nwin] */
static void signed_digits(const u64* k, int c, int nwin, int32_t* out) {
    u64 carry = 0;
    for (int w = 0; w < nwin; w++) {
        int pos = w * c;
        u64 raw = 0;
        if (pos < 256) {
            int limb = pos >> 6, off = pos & 63;
            raw = k[limb] >> off;
            if (off + c > 64 && limb + 1 < FRN) raw |= k[limb + 1] << (64 - off);
            raw &= ((u64)1 << c) - 1;
        }
        raw += carry;
        if (raw > ((u64)1 << (c - 1))) {
            out[w] = (int32_t)((int64_t)raw - ((int64_t)1 << c));
            carry = 1;
        } else {
            out[w] = (int32_t)raw;
            carry = 0;
        }
    }
}

/* sum_k scalars[k] * pts[k]: tiles = windows x point chunks (the "parallel tiling
 * Pippenger" shape of rust-kzg's msm), then per-window sums and a Horner ladder */
static void msm_core(g1j_t* out, const g1a_t* pts, const u64 (*ks)[FRN], size_t n) {
    g1j_set_inf(out);
    if (n == 0) return;
    if (n < 8) { /* naive for tiny inputs, as the reference's msm does */
        for (size_t i = 0; i < n; i++) {
            g1j_t t;
            g1_mul(&t, &pts[i], ks[i]);
            g1j_add(out, out, &t);
        }
        return;
    }
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    size_t nchunks = 1;
    if (n >= 4096) {
        nchunks = (size_t)nthreads;
        while (nchunks > 1 && n / nchunks < 2048) nchunks--;
    }
    size_t chunk = (n + nchunks - 1) / nchunks;
    int c = pick_window(chunk);
    int nwin = (256 + c - 1) / c;
    int32_t* digits = (int32_t*)malloc(sizeof(int32_t) * n * (size_t)nwin);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) signed_digits(ks[i], c, nwin, digits + (size_t)i * nwin);
    size_t ntiles = nchunks * (size_t)nwin;
    g1j_t* tile_sum = (g1j_t*)malloc(sizeof(g1j_t) * ntiles);
    size_t nb = ((size_t)1 << (c - 1)) + 1;
#pragma omp parallel
    {
        g1j_t* buckets = (g1j_t*)malloc(sizeof(g1j_t) * nb);
#pragma omp for schedule(dynamic, 1)
        for (long t = 0; t < (long)ntiles; t++) {
            int w = (int)(t % nwin);
            size_t ch = (size_t)t / nwin;
            size_t lo = ch * chunk, hi = lo + chunk < n ? lo + chunk : n;
            for (size_t b = 0; b < nb; b++) g1j_set_inf(&buckets[b]);
            for (size_t i = lo; i < hi; i++) {
                int32_t d = digits[i * nwin + w];
                if (d > 0) {
                    g1j_add_affine(&buckets[d], &buckets[d], &pts[i]);
                } else if (d < 0) {
                    g1a_t neg = pts[i];
                    fp_neg(&neg.y, &neg.y);
                    g1j_add_affine(&buckets[-d], &buckets[-d], &neg);
                }
            }
            g1j_t run, acc;
            g1j_set_inf(&run);
            g1j_set_inf(&acc);
            for (size_t b = nb - 1; b >= 1; b--) {
                g1j_add(&run, &run, &buckets[b]);
                g1j_add(&acc, &acc, &run);
            }
            tile_sum[t] = acc;
        }
        free(buckets);
    }
    g1j_t total;
    g1j_set_inf(&total);
    for (int w = nwin - 1; w >= 0; w--) {
        for (int d = 0; d < c; d++) g1j_dbl(&total, &total);
        for (size_t ch = 0; ch < nchunks; ch++) g1j_add(&total, &total, &tile_sum[ch * nwin + w]);
    }
    *out = total;
    free(tile_sum);
    free(digits);
}

/* ---------------------------------------------------------------------- FFT */
static void fr_root(fr_t* w, int log_n, int inverse) {
    fr_t r, r2;
    memcpy(r.v, FR_W32, sizeof r.v);
    memcpy(r2.v, FR_R2, sizeof r2.v);
    fr_mul(&r, &r, &r2);
    for (int i = log_n; i < 32; i++) fr_mul(&r, &r, &r);
    if (inverse) fr_inv(&r, &r);
    *w = r;
}

/* natural-order in/out radix-2 transform (bit reversal + decimation in time) */
static void fft_core(fr_t* a, size_t n, int inverse) {
    if (n <= 1) return;
    int log_n = 0;
    while (((size_t)1 << log_n) < n) log_n++;
    for (size_t i = 1, j = 0; i < n; i++) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j |= bit;
        if (i < j) { fr_t t = a[i]; a[i] = a[j]; a[j] = t; }
    }
    fr_t w;
    fr_root(&w, log_n, inverse);
    fr_t* tw = (fr_t*)malloc(sizeof(fr_t) * (n / 2));
    fr_set_one(&tw[0]);
    for (size_t k = 1; k < n / 2; k++) fr_mul(&tw[k], &tw[k - 1], &w);
    for (size_t len = 2; len <= n; len <<= 1) {
        size_t half = len >> 1, step = n / len;
#pragma omp parallel for schedule(static) if (n >= 4096)
        for (long p = 0; p < (long)(n / 2); p++) {
            size_t blk = (size_t)p / half, k = (size_t)p % half;
            size_t i = blk * len + k;
            fr_t v, u = a[i];
            fr_mul(&v, &a[i + half], &tw[k * step]);
            fr_add(&a[i], &u, &v);
            fr_sub(&a[i + half], &u, &v);
        }
    }
    free(tw);
    if (inverse) {
        fr_t ninv, nn = {{(u64)n, 0, 0, 0}}, r2;
        memcpy(r2.v, FR_R2, sizeof r2.v);
        fr_mul(&nn, &nn, &r2);
        fr_inv(&ninv, &nn);
#pragma omp parallel for schedule(static) if (n >= 4096)
        for (long i = 0; i < (long)n; i++) fr_mul(&a[i], &a[i], &ninv);
    }
}

/* FsPoly::eval (Horner) */
static void poly_eval(fr_t* y, const fr_t* p, size_t n, const fr_t* x) {
    fr_t acc;
    memset(&acc, 0, sizeof acc);
    for (size_t k = n; k > 0; k--) {
        fr_mul(&acc, &acc, x);
        fr_add(&acc, &acc, &p[k - 1]);
    }
    *y = acc;
}
/* FsPoly::div by (X - a): q has n-1 coefficients, remainder dropped */
static void poly_div_linear(fr_t* q, const fr_t* p, size_t n, const fr_t* a) {
    if (n < 2) return;
    q[n - 2] = p[n - 1];
    for (size_t k = n - 2; k > 0; k--) {
        fr_t t;
        fr_mul(&t, a, &q[k]);
        fr_add(&q[k - 1], &p[k], &t);
    }
}
/* field half of worker_open (piano.rs:193-233): coeffs -> y, NTT(q || 0) in place */
static void open_scalars_core(fr_t* coeffs, size_t n, const fr_t* alpha, fr_t* y) {
    fft_core(coeffs, n, 1);
    poly_eval(y, coeffs, n, alpha);
    fr_t* q = (fr_t*)calloc(n ? n : 1, sizeof(fr_t));
    poly_div_linear(q, coeffs, n, alpha);
    memcpy(coeffs, q, sizeof(fr_t) * n);
    free(q);
    fft_core(coeffs, n, 0);
}

/* ---------------------------------------------------------------- exports */
static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int or_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void or_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static int decode_frs(fr_t* out, const uint8_t* be, size_t n) {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad) if (n >= 4096)
    for (long i = 0; i < (long)n; i++) bad |= fr_from_be(&out[i], be + 32 * (size_t)i);
    return bad;
}
static void encode_frs(uint8_t* be, const fr_t* in, size_t n) {
#pragma omp parallel for schedule(static) if (n >= 4096)
    for (long i = 0; i < (long)n; i++) fr_to_be(be + 32 * (size_t)i, &in[i]);
}
static int decode_points96(g1a_t* out, const uint8_t* in, size_t n) {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad) if (n >= 1024)
    for (long i = 0; i < (long)n; i++) bad |= g1a_deserialize(&out[i], in + 96 * (size_t)i);
    return bad;
}
static void canon_scalars(u64 (*ks)[FRN], const fr_t* in, size_t n) {
#pragma omp parallel for schedule(static) if (n >= 4096)
    for (long i = 0; i < (long)n; i++) fr_canon(ks[i], &in[i]);
}

/* fft_fr on wire bytes; 0 ok, 1 invalid scalar, 2 bad length */
int or_ntt(const uint8_t* in_be, uint64_t n, int inverse, uint8_t* out_be) {
    if (n == 0 || (n & (n - 1))) return 2;
    fr_t* a = (fr_t*)malloc(sizeof(fr_t) * n);
    if (decode_frs(a, in_be, n)) { free(a); return 1; }
    fft_core(a, n, inverse);
    encode_frs(out_be, a, n);
    free(a);
    return 0;
}
int or_eval(const uint8_t* poly_be, uint64_t n, const uint8_t* x_be, uint8_t* y_be) {
    fr_t* a = (fr_t*)malloc(sizeof(fr_t) * (n ? n : 1));
    fr_t x, y;
    if (decode_frs(a, poly_be, n) || fr_from_be(&x, x_be)) { free(a); return 1; }
    poly_eval(&y, a, n, &x);
    fr_to_be(y_be, &y);
    free(a);
    return 0;
}
int or_open_scalars(const uint8_t* coeffs_be, uint64_t n, const uint8_t* alpha_be, uint8_t* y_be, uint8_t* qhat_be) {
    if (n == 0 || (n & (n - 1))) return 2;
    fr_t* a = (fr_t*)malloc(sizeof(fr_t) * n);
    fr_t alpha, y;
    if (decode_frs(a, coeffs_be, n) || fr_from_be(&alpha, alpha_be)) { free(a); return 1; }
    open_scalars_core(a, n, &alpha, &y);
    fr_to_be(y_be, &y);
    encode_frs(qhat_be, a, n);
    free(a);
    return 0;
}

/* out[j] = L_j(tau) for the size-2^log_n domain, closed form with one batch inversion
 * (same field elements as the reference's iFFT(e_j)+Horner, piano.rs:956-988) */
int or_lagrange_at(const uint8_t* tau_be, int log_n, uint8_t* out_be) {
    size_t n = (size_t)1 << log_n;
    fr_t tau, w, zn, ninv, one;
    if (fr_from_be(&tau, tau_be)) return 1;
    fr_set_one(&one);
    fr_root(&w, log_n, 0);
    zn = tau;
    for (int i = 0; i < log_n; i++) fr_mul(&zn, &zn, &zn);
    fr_sub(&zn, &zn, &one);
    fr_t* roots = (fr_t*)malloc(sizeof(fr_t) * n);
    fr_t* d = (fr_t*)malloc(sizeof(fr_t) * n);
    fr_t* pre = (fr_t*)malloc(sizeof(fr_t) * (n + 1));
    roots[0] = one;
    for (size_t j = 1; j < n; j++) fr_mul(&roots[j], &roots[j - 1], &w);
    fr_t* out = (fr_t*)calloc(n, sizeof(fr_t));
    if (is_zero_n(zn.v, FRN)) {
        for (size_t j = 0; j < n; j++)
            if (memcmp(roots[j].v, tau.v, sizeof tau.v) == 0) out[j] = one;
    } else {
        fr_t nn = {{(u64)n, 0, 0, 0}}, r2;
        memcpy(r2.v, FR_R2, sizeof r2.v);
        fr_mul(&nn, &nn, &r2);
        fr_inv(&ninv, &nn);
        pre[0] = one;
        for (size_t j = 0; j < n; j++) {
            fr_sub(&d[j], &tau, &roots[j]);
            fr_mul(&pre[j + 1], &pre[j], &d[j]);
        }
        fr_t inv;
        fr_inv(&inv, &pre[n]);
        fr_t c;
        fr_mul(&c, &zn, &ninv);
        for (size_t j = n; j > 0; j--) {
            fr_t dinv;
            fr_mul(&dinv, &inv, &pre[j - 1]);
            fr_mul(&inv, &inv, &d[j - 1]);
            fr_mul(&out[j - 1], &roots[j - 1], &c);
            fr_mul(&out[j - 1], &out[j - 1], &dinv);
        }
    }
    encode_frs(out_be, out, n);
    free(roots); free(d); free(pre); free(out);
    return 0;
}

/* out96[k] = scalars[k] * G, uncompressed (generate_trusted_setup's G1::mul,
 * piano.rs:916-1007), 8-bit fixed-base comb + batch normalisation */
int or_g1_mul_gen(const uint8_t* scalars_be, uint64_t n, uint8_t* out96) {
    static g1a_t* table = NULL; /* [32][255] */
    if (!table) {
        g1a_t* t = (g1a_t*)malloc(sizeof(g1a_t) * 32 * 255);
        g1a_t g;
        g1_generator(&g);
        g1j_t base;
        base.x = g.x; base.y = g.y; fp_set_one(&base.z);
        for (int w = 0; w < 32; w++) {
            g1j_t acc = base;
            for (int d = 1; d <= 255; d++) {
                g1j_to_affine(&t[w * 255 + d - 1], &acc);
                g1j_add(&acc, &acc, &base);
            }
            for (int i = 0; i < 8; i++) g1j_dbl(&base, &base);
        }
        table = t;
    }
    fr_t* s = (fr_t*)malloc(sizeof(fr_t) * (n ? n : 1));
    if (decode_frs(s, scalars_be, n)) { free(s); return 1; }
    int bad = 0;
#pragma omp parallel for schedule(dynamic, 256)
    for (long i = 0; i < (long)n; i++) {
        u64 k[FRN];
        fr_canon(k, &s[i]);
        g1j_t acc;
        g1j_set_inf(&acc);
        for (int w = 0; w < 32; w++) {
            unsigned d = (unsigned)(k[w >> 3] >> (8 * (w & 7))) & 0xff;
            if (d) g1j_add_affine(&acc, &acc, &table[w * 255 + d - 1]);
        }
        g1a_t a;
        g1j_to_affine(&a, &acc);
        g1a_serialize(out96 + 96 * (size_t)i, &a);
    }
    free(s);
    return bad;
}

/* g1_linear_combination on wire bytes: points uncompressed (96) */
int or_msm(const uint8_t* points96, const uint8_t* scalars_be, uint64_t n, uint8_t* out48) {
    g1a_t* pts = (g1a_t*)malloc(sizeof(g1a_t) * (n ? n : 1));
    fr_t* s = (fr_t*)malloc(sizeof(fr_t) * (n ? n : 1));
    u64 (*ks)[FRN] = (u64(*)[FRN])malloc(sizeof(u64[FRN]) * (n ? n : 1));
    int rc = 0;
    if (decode_points96(pts, points96, n)) rc = 4;
    else if (decode_frs(s, scalars_be, n)) rc = 1;
    else {
        canon_scalars(ks, s, n);
        g1j_t r;
        msm_core(&r, pts, ks, n);
        g1a_t a;
        g1j_to_affine(&a, &r);
        g1a_compress(out48, &a);
    }
    free(pts); free(s); free(ks);
    return rc;
}

int or_g1_sum(const uint8_t* points48, uint64_t n, uint8_t* out48) {
    g1j_t acc;
    g1j_set_inf(&acc);
    for (uint64_t i = 0; i < n; i++) {
        g1a_t p;
        int rc = g1a_uncompress(&p, points48 + 48 * i);
        if (rc) return rc;
        g1j_add_affine(&acc, &acc, &p);
    }
    g1a_t a;
    g1j_to_affine(&a, &acc);
    g1a_compress(out48, &a);
    return 0;
}
int or_g1_recode(const uint8_t* in, int in_fmt, uint8_t* out, int out_fmt) {
    g1a_t p;
    int rc = in_fmt == 48 ? g1a_uncompress(&p, in) : g1a_deserialize(&p, in);
    if (rc) return rc;
    if (out_fmt == 48) g1a_compress(out, &p);
    else g1a_serialize(out, &p);
    return 0;
}

/* ---- timed entry points (cpu_baseline / --impl reference).  Decoding the wire
 * bytes happens outside the timed region; each returns seconds per repetition. */
double or_bench_ntt(const uint8_t* in_be, uint64_t n, int inverse, int reps) {
    fr_t* a = (fr_t*)malloc(sizeof(fr_t) * n);
    fr_t* b = (fr_t*)malloc(sizeof(fr_t) * n);
    decode_frs(a, in_be, n);
    double t = 0;
    for (int r = 0; r < reps; r++) {
        memcpy(b, a, sizeof(fr_t) * n);
        double t0 = now_s();
        fft_core(b, n, inverse);
        t += now_s() - t0;
    }
    free(a); free(b);
    return t / reps;
}
double or_bench_msm(const uint8_t* points96, const uint8_t* scalars_be, uint64_t n, int reps, uint8_t* out48) {
    g1a_t* pts = (g1a_t*)malloc(sizeof(g1a_t) * n);
    fr_t* s = (fr_t*)malloc(sizeof(fr_t) * n);
    u64 (*ks)[FRN] = (u64(*)[FRN])malloc(sizeof(u64[FRN]) * n);
    decode_points96(pts, points96, n);
    decode_frs(s, scalars_be, n);
    double t = 0;
    g1j_t r;
    for (int k = 0; k < reps; k++) {
        double t0 = now_s();
        canon_scalars(ks, s, n);
        msm_core(&r, pts, ks, n);
        t += now_s() - t0;
    }
    g1a_t a;
    g1j_to_affine(&a, &r);
    g1a_compress(out48, &a);
    free(pts); free(s); free(ks);
    return t / reps;
}
/* one worker_commit + one worker_open (piano.rs:143-161, 187-238) on the same row;
 * times[0] = commit seconds, times[1] = open seconds (per repetition) */
int or_bench_commit_open(const uint8_t* points96, const uint8_t* coeffs_be, uint64_t n, const uint8_t* alpha_be, int reps,
                         uint8_t* commit48, uint8_t* y32, uint8_t* proof48, double* times) {
    g1a_t* pts = (g1a_t*)malloc(sizeof(g1a_t) * n);
    fr_t* s = (fr_t*)malloc(sizeof(fr_t) * n);
    fr_t* w = (fr_t*)malloc(sizeof(fr_t) * n);
    u64 (*ks)[FRN] = (u64(*)[FRN])malloc(sizeof(u64[FRN]) * n);
    fr_t alpha, y;
    int rc = 0;
    if (decode_points96(pts, points96, n)) rc = 4;
    else if (decode_frs(s, coeffs_be, n) || fr_from_be(&alpha, alpha_be)) rc = 1;
    else {
        g1j_t c, p;
        times[0] = times[1] = 0;
        for (int k = 0; k < reps; k++) {
            double t0 = now_s();
            canon_scalars(ks, s, n);
            msm_core(&c, pts, ks, n);
            double t1 = now_s();
            memcpy(w, s, sizeof(fr_t) * n);
            open_scalars_core(w, n, &alpha, &y);
            canon_scalars(ks, w, n);
            msm_core(&p, pts, ks, n);
            double t2 = now_s();
            times[0] += t1 - t0;
            times[1] += t2 - t1;
        }
        times[0] /= reps;
        times[1] /= reps;
        g1a_t a;
        g1j_to_affine(&a, &c);
        g1a_compress(commit48, &a);
        g1j_to_affine(&a, &p);
        g1a_compress(proof48, &a);
        fr_to_be(y32, &y);
    }
    free(pts); free(s); free(w); free(ks);
    return rc;
}
