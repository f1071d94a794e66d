// host_arith.cpp - TEST HARNESS ONLY.  Compiles the product's host+device
// arithmetic headers (fourier_b200/csrc/*.cuh) for the HOST with g++, where the
// PTX carry-chain primitives are emulated instruction by instruction
// (limbs.cuh), and exports them over a C ABI so pytest can compare every routine
// with the Python big-int oracle on a machine without a GPU.  Nothing here is
// linked into libfourier_b200.so.
#include <string.h>

#include "../fourier_b200/csrc/g1.cuh"

using namespace fb;

extern "C" {

void h_fr_mul(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fr x, y;
    memcpy(x.v, a, 32);
    memcpy(y.v, b, 32);
    Fr z = fe_mul<FrParams>(x, y);
    memcpy(r, z.v, 32);
}
void h_fr_mul_portable(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fr x, y;
    memcpy(x.v, a, 32);
    memcpy(y.v, b, 32);
    Fr z = fe_mul_portable<FrParams>(x, y);
    memcpy(r, z.v, 32);
}
void h_fp_mul(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fp x, y;
    memcpy(x.v, a, 48);
    memcpy(y.v, b, 48);
    Fp z = fe_mul<FpParams>(x, y);
    memcpy(r, z.v, 48);
}
void h_fp_mul_portable(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fp x, y;
    memcpy(x.v, a, 48);
    memcpy(y.v, b, 48);
    Fp z = fe_mul_portable<FpParams>(x, y);
    memcpy(r, z.v, 48);
}
void h_fr_sqr(const uint32_t* a, uint32_t* r) {
    Fr x;
    memcpy(x.v, a, 32);
    Fr z = fe_sqr<FrParams>(x);
    memcpy(r, z.v, 32);
}
void h_fp_sqr(const uint32_t* a, uint32_t* r) {
    Fp x;
    memcpy(x.v, a, 48);
    Fp z = fe_sqr<FpParams>(x);
    memcpy(r, z.v, 48);
}
// op: 0 add, 1 sub, 2 neg(a), 3 inv(a), 4 to_mont(a), 5 from_mont(a), 6 inv_gcd(a)
void h_fr_op(int op, const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fr x, y, z;
    memcpy(x.v, a, 32);
    memcpy(y.v, b, 32);
    switch (op) {
        case 0: z = fe_add<FrParams>(x, y); break;
        case 1: z = fe_sub<FrParams>(x, y); break;
        case 2: z = fe_neg<FrParams>(x); break;
        case 3: z = fe_inv<FrParams>(x); break;
        case 4: z = fe_to_mont<FrParams>(x); break;
        case 6: z = fe_inv_gcd<FrParams>(x); break;
        case 7: z = fe_inv_safegcd<FrParams>(x); break;
        default: z = fe_from_mont<FrParams>(x); break;
    }
    memcpy(r, z.v, 32);
}
void h_fp_op(int op, const uint32_t* a, const uint32_t* b, uint32_t* r) {
    Fp x, y, z;
    memcpy(x.v, a, 48);
    memcpy(y.v, b, 48);
    switch (op) {
        case 0: z = fe_add<FpParams>(x, y); break;
        case 1: z = fe_sub<FpParams>(x, y); break;
        case 2: z = fe_neg<FpParams>(x); break;
        case 3: z = fe_inv<FpParams>(x); break;
        case 4: z = fe_to_mont<FpParams>(x); break;
        case 6: z = fe_inv_gcd<FpParams>(x); break;
        case 7: z = fe_inv_safegcd<FpParams>(x); break;
        default: z = fe_from_mont<FpParams>(x); break;
    }
    memcpy(r, z.v, 48);
}

// points cross the harness as 48-byte compressed strings
int h_g1_decompress_compress(const uint8_t* in48, uint8_t* out48) {
    G1Affine p;
    int rc = g1_decompress(in48, &p);
    if (rc) return rc;
    if (!g1_on_curve(p)) return 100;
    g1_compress(p, out48);
    return 0;
}
int h_g1_serialize_roundtrip(const uint8_t* in48, uint8_t* out96, uint8_t* back48) {
    G1Affine p, q;
    int rc = g1_decompress(in48, &p);
    if (rc) return rc;
    g1_serialize(p, out96);
    rc = g1_deserialize(out96, &q);
    if (rc) return 50 + rc;
    g1_compress(q, back48);
    return 0;
}
// out = a + b with a via XYZZ accumulator and b affine (mixed) when mode == 0,
// both XYZZ when mode == 1 (a is first doubled-and-halved through a general add
// with the identity to exercise that path), out = 2a when mode == 2.
int h_g1_add(int mode, const uint8_t* a48, const uint8_t* b48, uint8_t* out48) {
    G1Affine a, b;
    int rc = g1_decompress(a48, &a);
    if (rc) return rc;
    rc = g1_decompress(b48, &b);
    if (rc) return rc;
    G1Xyzz acc = G1Xyzz::from_affine(a);
    if (mode == 0) {
        g1_add_mixed(acc, b);
    } else if (mode == 1) {
        // make both operands non-trivially projective: acc = 2a - a, q = 2b - b
        G1Xyzz a2 = g1_dbl(acc);
        g1_add_mixed(a2, g1_neg(a));
        G1Xyzz q = G1Xyzz::from_affine(b);
        G1Xyzz b2 = g1_dbl(q);
        g1_add_mixed(b2, g1_neg(b));
        acc = a2;
        g1_add(acc, b2);
    } else {
        acc = g1_dbl(acc);
    }
    g1_compress(g1_to_affine(acc), out48);
    return 0;
}
int h_g1_mul(const uint8_t* a48, const uint32_t* k8, uint8_t* out48) {
    G1Affine a;
    int rc = g1_decompress(a48, &a);
    if (rc) return rc;
    g1_compress(g1_to_affine(g1_mul_scalar(a, k8)), out48);
    return 0;
}
void h_g1_generator(uint8_t* out48) { g1_compress(g1_generator(), out48); }

}  // extern "C"
