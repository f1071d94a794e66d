// verify.cu - worker_verify / master_verify of the C ABI (reference
// src/engine/piano.rs:309-465).  The G1 side (the MSM over g^{tau_Y^j}, the scalar
// multiples and the subtraction) runs on the GPU through the same msm_run as the
// commit path; the G2 side and the pairing product are a handful of field
// operations and run on the host (g2host.cpp).
#include <string.h>

#include "ctx.cuh"

namespace fb {
Fr* ctx_upload_fr(fb_ctx* c, fb_lane* L, DevBuf& bytes, DevBuf& dst, const uint8_t* h_be, size_t count);
void ctx_reset_err(fb_ctx* c, fb_lane* L);
int ctx_check_err(fb_ctx* c, fb_lane* L);
// g2host.cpp
bool g1_deserialize_host(const uint8_t in[96], G1Affine* out);
bool g2_sub_mul_host(const std::vector<uint8_t>& base, const std::vector<uint8_t>& g2, const uint32_t k[8], uint8_t out192[192]);
bool g2_to_serialized_host(const std::vector<uint8_t>& in, uint8_t out192[192]);
int pairing_product_is_one_host(const G1Affine* g1, const uint8_t* g2_192, size_t n);
bool pairing_value_host(const G1Affine* g1, const uint8_t* g2_192, size_t n, uint8_t out576[576]);
}  // namespace fb

using namespace fb;

namespace {

// r - 1, big-endian: the scalar of the term that is subtracted
const uint8_t kMinusOneBe[32] = {0x73, 0xed, 0xa7, 0x53, 0x29, 0x9d, 0x7d, 0x48, 0x33, 0x39, 0xd8, 0x08, 0x09, 0xa1, 0xd8, 0x05,
                                 0x53, 0xbd, 0xa4, 0x02, 0xff, 0xfe, 0x5b, 0xfe, 0xff, 0xff, 0xff, 0xff, 0x00, 0x00, 0x00, 0x00};

// 32 big-endian bytes -> canonical limbs; false when >= r (FsFr::from_bytes, "Invalid scalar")
bool scalar_limbs(const uint8_t be[32], uint32_t k[8]) {
    for (int i = 0; i < 8; i++) {
        const uint8_t* b = be + 4 * (7 - i);
        k[i] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    Fr m = Fr::modulus();
    for (int i = 7; i >= 0; i--)
        if (k[i] != m.v[i]) return k[i] < m.v[i];
    return false;
}

// d_out (XYZZ) = s * pts[0] - pts[1] for the device affine points pts[0..1]
void scaled_difference(fb_ctx* c, fb_lane* L, const G1Affine* d_pts, const uint8_t s_be[32], G1Xyzz* d_out) {
    uint8_t h[64];
    memcpy(h, s_be, 32);
    memcpy(h + 32, kMinusOneBe, 32);
    Fr* d_s = ctx_upload_fr(c, L, L->small, L->eng_a, h, 2);
    MsmBases two;
    two.pts = (void*)d_pts;
    two.n = 2;
    msm_run(two, d_s, 2, 1, true, d_out, L->msm, L->stream);
}

// the G1 points of the product go back as Montgomery affine structs; the check itself is host work
int finish_pairing(fb_ctx* c, fb_lane* L, const G1Affine* d_pts, size_t n, const uint8_t* g2_192, int* out_valid) {
    FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_pts, n * sizeof(G1Affine), cudaMemcpyDeviceToHost, L->stream));
    int rc = ctx_check_err(c, L);  // synchronises
    if (rc) return rc;
    G1Affine pts[3];
    memcpy(pts, L->h_pinned, n * sizeof(G1Affine));
    int ok = pairing_product_is_one_host(pts, g2_192, n);
    if (ok < 0) return ctx_fail(c, FB_ERR_STATE, "bad G2 element in the setup");
    *out_valid = ok;
    return FB_OK;
}

}  // namespace

extern "C" {

// e(C - y * g^{R_i(tau_Y)}, g2) == e(pi, g2^{tau_X} - alpha * g2)   (piano.rs:309-356)
int fb_worker_verify(fb_ctx* c, uint64_t i, const uint8_t commitment48[48], const uint8_t alpha_be[32], const uint8_t y_be[32],
                     const uint8_t proof48[48], int* out_valid) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
        if (!out_valid) return ctx_fail(c, FB_ERR_ARG, "out_valid is null");
        if (i >= c->M) return ctx_fail(c, FB_ERR_ARG, "machine index out of range");
        uint32_t alpha[8];
        if (!scalar_limbs(alpha_be, alpha)) return ctx_fail(c, FB_ERR_SCALAR, "Invalid scalar");
        ctx_reset_err(c, L);
        cudaStream_t st = L->stream;
        // R_i in the standard basis = iNTT_right(e_i) (right_lagrange_poly, piano.rs:1130-1135)
        std::vector<uint8_t> unit(c->M * 32, 0);
        unit[i * 32 + 31] = 1;
        Fr* d_r = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, unit.data(), c->M);
        if (c->M > 1) {
            Fr* d_coef = L->fr_b.as<Fr>(c->M);
            ntt_run(ctx_ntt_tables(c, L, (int)c->m), c->tile_tw, d_r, d_coef, L->fr_c.as<Fr>(c->M), true, st);
            d_r = d_coef;
        }
        // g^{R_i(tau_Y)} = sum_j r_j g^{tau_Y^j}: no base change (piano.rs:326-335)
        G1Xyzz* d_x = (G1Xyzz*)L->eng_b.get(4 * sizeof(G1Xyzz));
        msm_run(c->y_bases, d_r, c->M, 1, true, d_x, L->msm, st);
        G1Affine* pts = L->pt_a.as<G1Affine>(8);  // [R, C, pi | -lhs, pi]
        g1_xyzz_to_affine_batch(d_x, pts, 1, st);
        uint8_t h[96];
        memcpy(h, commitment48, 48);
        memcpy(h + 48, proof48, 48);
        uint8_t* d_pb = L->out_bytes.as<uint8_t>(96 + 16);
        FB_CUDA(cudaMemcpyAsync(d_pb, h, 96, cudaMemcpyHostToDevice, st));
        g1_decompress_batch(d_pb, pts + 1, 2, L->d_err + 1, st);
        // y * R - C = -(C - g^{y'}): the negated left side of pairings_verify_single (:338-341, :369-372)
        scaled_difference(c, L, pts, y_be, d_x + 1);
        g1_xyzz_to_affine_batch(d_x + 1, pts + 3, 1, st);
        FB_CUDA(cudaMemcpyAsync(pts + 4, pts + 2, sizeof(G1Affine), cudaMemcpyDeviceToDevice, st));
        uint8_t g2s[2 * 192];
        if (!g2_to_serialized_host(c->g2_bytes[0], g2s) || !g2_sub_mul_host(c->g2_bytes[1], c->g2_bytes[0], alpha, g2s + 192))
            return ctx_fail(c, FB_ERR_STATE, "bad G2 element in the setup");
        return finish_pairing(c, L, pts + 3, 2, g2s, out_valid);
    });
}

// e(C - z * g, g2) == e(pi0, g2^{tau_X} - alpha g2) * e(pi1, g2^{tau_Y} - beta g2)   (piano.rs:393-420)
int fb_master_verify(fb_ctx* c, const uint8_t commitment48[48], const uint8_t beta_be[32], const uint8_t alpha_be[32],
                     const uint8_t z_be[32], const uint8_t pi0_48[48], const uint8_t pi1_48[48], int* out_valid) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
        if (!out_valid) return ctx_fail(c, FB_ERR_ARG, "out_valid is null");
        uint32_t alpha[8], beta[8];
        if (!scalar_limbs(alpha_be, alpha) || !scalar_limbs(beta_be, beta)) return ctx_fail(c, FB_ERR_SCALAR, "Invalid scalar");
        ctx_reset_err(c, L);
        cudaStream_t st = L->stream;
        G1Affine* pts = L->pt_a.as<G1Affine>(8);  // [g, C | -lhs, pi0, pi1]
        FB_CUDA(cudaMemcpyAsync(pts, &c->g, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
        uint8_t h[144];
        memcpy(h, commitment48, 48);
        memcpy(h + 48, pi0_48, 48);
        memcpy(h + 96, pi1_48, 48);
        uint8_t* d_pb = L->out_bytes.as<uint8_t>(144 + 16);
        FB_CUDA(cudaMemcpyAsync(d_pb, h, 144, cudaMemcpyHostToDevice, st));
        g1_decompress_batch(d_pb, pts + 1, 1, L->d_err + 1, st);
        g1_decompress_batch(d_pb + 48, pts + 3, 2, L->d_err + 1, st);
        G1Xyzz* d_x = (G1Xyzz*)L->eng_b.get(4 * sizeof(G1Xyzz));
        scaled_difference(c, L, pts, z_be, d_x);  // z * g - C  (:402)
        g1_xyzz_to_affine_batch(d_x, pts + 2, 1, st);
        uint8_t g2s[3 * 192];
        if (!g2_to_serialized_host(c->g2_bytes[0], g2s) || !g2_sub_mul_host(c->g2_bytes[1], c->g2_bytes[0], alpha, g2s + 192) ||
            !g2_sub_mul_host(c->g2_bytes[2], c->g2_bytes[0], beta, g2s + 384))
            return ctx_fail(c, FB_ERR_STATE, "bad G2 element in the setup");
        return finish_pairing(c, L, pts + 2, 3, g2s, out_valid);
    });
}

// Context-free pairing product on serialized points (host arithmetic only):
// g1 = n x 96 bytes (blst_p1_serialize), g2 = n x 192 bytes (blst_p2_serialize).
int fb_pairing_check(const uint8_t* g1_96, const uint8_t* g2_192, uint64_t n, int* out_is_one) {
    if (!out_is_one || (n && (!g1_96 || !g2_192))) return FB_ERR_ARG;
    std::vector<G1Affine> pts(n);
    for (uint64_t k = 0; k < n; k++)
        if (!g1_deserialize_host(g1_96 + 96 * k, &pts[k])) return FB_ERR_POINT;
    int ok = pairing_product_is_one_host(pts.data(), g2_192, n);
    if (ok < 0) return FB_ERR_POINT;
    *out_is_one = ok;
    return FB_OK;
}

int fb_pairing_value(const uint8_t* g1_96, const uint8_t* g2_192, uint64_t n, uint8_t out576[576]) {
    if (!out576 || (n && (!g1_96 || !g2_192))) return FB_ERR_ARG;
    std::vector<G1Affine> pts(n);
    for (uint64_t k = 0; k < n; k++)
        if (!g1_deserialize_host(g1_96 + 96 * k, &pts[k])) return FB_ERR_POINT;
    return pairing_value_host(pts.data(), g2_192, n, out576) ? FB_OK : FB_ERR_POINT;
}

}  // extern "C"
