// engine.cu - the PianoBackend calls of the C ABI: SRS residency and the
// worker / master commit + open paths (reference src/engine/piano.rs:143-298,
// 587-846, 883-1037).
#include <stdlib.h>
#include <string.h>

#include <fstream>

#include "ctx.cuh"

namespace fb {
Fr* ctx_upload_fr(fb_ctx* c, fb_lane* L, DevBuf& bytes, DevBuf& dst, const uint8_t* h_be, size_t count);
void ctx_reset_err(fb_ctx* c, fb_lane* L);
int ctx_check_err(fb_ctx* c, fb_lane* L);
void ctx_points_out(fb_ctx* c, fb_lane* L, const G1Xyzz* d_xyzz, size_t count, uint8_t* d_out48);
bool g2_generate_host(const uint8_t sx[32], const uint8_t sy[32], std::vector<uint8_t> out[3]);  // g2host.cpp
bool g2_recode_host(const std::vector<uint8_t>& in, bool to_compressed, std::vector<uint8_t>& out);
}  // namespace fb

using namespace fb;

namespace {

void srs_free(fb_ctx* c) {
    if (c->u_bases.pts && c->u_bases.pts != c->d_u) cudaFree(c->u_bases.pts);
    if (c->u_bases.xs) cudaFree((void*)c->u_bases.xs);
    if (c->y_bases.pts && c->y_bases.pts != c->d_g_tau_y) cudaFree(c->y_bases.pts);
    if (c->y_bases.xs) cudaFree((void*)c->y_bases.xs);
    if (c->d_u) cudaFree(c->d_u);
    if (c->d_g_tau_x) cudaFree(c->d_g_tau_x);
    if (c->d_g_tau_y) cudaFree(c->d_g_tau_y);
    c->d_u = c->d_g_tau_x = c->d_g_tau_y = nullptr;
    c->u_bases = MsmBases();
    c->y_bases = MsmBases();
    c->srs_ready = false;
}

// build the resident window tables (PianoPrecomputation::generate, piano.rs:477-504)
void srs_build_tables(fb_ctx* c, fb_lane* L) {
    c->u_bases = MsmBases();
    c->u_bases.n = c->ncols;
    c->u_bases.pts = c->d_u;
    c->y_bases = MsmBases();
    c->y_bases.n = c->M;
    c->y_bases.pts = c->d_g_tau_y;
    if (c->window_pref < 0) return;
    if (c->ncols >= 16) {
        int cc = c->window_pref > 0 ? c->window_pref : msm_pick_window(c->ncols, true);
        if (cc < 2 || cc > 24) throw CudaError{"window bits out of range (2..24)"};
        int nwin = msm_num_windows(cc);
        void* tab;
        FB_CUDA(cudaMalloc(&tab, (size_t)nwin * c->ncols * c->nrows * kTableStride));
        msm_build_table(c->d_u, c->ncols, c->nrows, cc, nwin, tab, L->pt_a, L->stream);
        c->u_bases.pts = tab;
        c->u_bases.c = cc;
        c->u_bases.nwin = nwin;
        c->u_bases.xs = msm_alloc_xtable(tab, (size_t)nwin * c->ncols * c->nrows, L->stream);
    }
    if (c->M >= 16) {
        int cc = msm_pick_window(c->M, true);
        int nwin = msm_num_windows(cc);
        void* tab;
        FB_CUDA(cudaMalloc(&tab, (size_t)nwin * c->M * kTableStride));
        msm_build_table(c->d_g_tau_y, c->M, 1, cc, nwin, tab, L->pt_a, L->stream);
        c->y_bases.pts = tab;
        c->y_bases.c = cc;
        c->y_bases.nwin = nwin;
        c->y_bases.xs = msm_alloc_xtable(tab, (size_t)nwin * c->M, L->stream);
    }
    FB_CUDA(cudaStreamSynchronize(L->stream));
}

int check_rows(fb_ctx* c, uint64_t& row0, uint64_t& nrows) {
    if (nrows == 0) {
        row0 = 0;
        nrows = c->M;
    }
    if (row0 + nrows > c->M) return ctx_fail(c, FB_ERR_ARG, "row range exceeds 2^machines_scale");
    uint64_t c0 = c->want_col0, nc = c->want_ncols ? c->want_ncols : c->T;
    if (c0 + nc > c->T) return ctx_fail(c, FB_ERR_ARG, "column range exceeds 2^(scale - machines_scale)");
    return FB_OK;
}

MsmBases row_bases(fb_ctx* c, uint64_t i) {
    MsmBases b = c->u_bases;
    size_t per_row = (size_t)(b.c > 0 ? b.nwin : 1) * c->ncols;
    b.pts = (char*)c->u_bases.pts + (i - c->row0) * per_row * b.stride();
    if (b.xs) b.xs += (i - c->row0) * per_row * 64;
    return b;
}

int need_row(fb_ctx* c, uint64_t i, uint64_t cnt) {
    if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
    if (i + cnt > c->M) return ctx_fail(c, FB_ERR_ARG, "machine index out of range");
    if (i < c->row0 || i + cnt > c->row0 + c->nrows) return ctx_fail(c, FB_ERR_STATE, "row not resident on this context");
    return FB_OK;
}

bool col_sharded(fb_ctx* c) { return c->ncols != c->T; }
int need_full_cols(fb_ctx* c) {
    if (col_sharded(c)) return ctx_fail(c, FB_ERR_STATE, "column-sharded context: use the *_partial calls and add the parts");
    return FB_OK;
}

// MSM of one row over the resident column range: scalars are the FULL row (len)
void row_msm(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_scalars, size_t len, G1Xyzz* d_out) {
    size_t lo = c->col0 < len ? c->col0 : len;
    size_t hi = c->col0 + c->ncols < len ? c->col0 + c->ncols : len;
    lane_msm(c, L, row_bases(c, i), d_scalars + lo, hi - lo, 1, d_out);
}

// the field half of worker_open on device for `batch` rows at once: d_coeffs (Montgomery,
// batch x len) -> y (d_y[batch]) and q_hat = NTT(q || 0) per row in L->fr_b.  Returns q_hat.
Fr* open_scalars(fb_ctx* c, fb_lane* L, const NttTables& tab, const Fr* d_coeffs, size_t len, const Fr* d_alpha, Fr* d_y,
                 size_t batch = 1) {
    // Default: evaluation form, no transform (evalform.cu; same field elements).  fb_ctx_set_open_mode(ctx, 1)
    // keeps the reference's literal sequence iNTT -> eval -> divide -> NTT (piano.rs:193-236).
    if (c->open_mode == 0 && len >= 2) {
        Fr* d_q = L->fr_b.as<Fr>(len * batch);
        evalform_open(tab, d_coeffs, len, batch, d_alpha, d_y, d_q, L->scan, L->stream);
        return d_q;
    }
    Fr* d_poly = L->fr_b.as<Fr>(len * batch);
    Fr* d_tmp = L->fr_c.as<Fr>(len * batch);
    ntt_run(tab, c->tile_tw, d_coeffs, d_poly, d_tmp, true, L->stream, batch);          // piano.rs:194
    Fr* d_q = L->fr_a.as<Fr>(len * batch);  // NB: callers keep d_coeffs elsewhere or accept the overwrite
    horner_quotient(d_poly, len, d_alpha, d_y, d_q, L->scan, L->stream, batch);           // :202-225
    ntt_run(tab, c->tile_tw, d_q, d_poly, d_tmp, false, L->stream, batch);               // :233
    return d_poly;
}

}  // namespace

extern "C" {

int fb_srs_generate(fb_ctx* c, const uint8_t sx[32], const uint8_t sy[32], uint64_t row0, uint64_t nrows) {
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane* L) -> int {
        int rc = check_rows(c, row0, nrows);
        if (rc) return rc;
        srs_free(c);
        c->row0 = row0;
        c->nrows = nrows;
        c->col0 = c->want_col0;
        c->ncols = c->want_ncols ? c->want_ncols : c->T;
        const size_t NC = c->ncols;
        cudaStream_t st = L->stream;
        uint8_t* d_sec = L->in_bytes.as<uint8_t>(64);
        FB_CUDA(cudaMemcpyAsync(d_sec, sx, 32, cudaMemcpyHostToDevice, st));
        FB_CUDA(cudaMemcpyAsync(d_sec + 32, sy, 32, cudaMemcpyHostToDevice, st));
        Fr* d_small = L->small.as<Fr>(16);  // [0] tau_x [1] tau_y [2..4] hdr
        srs_secret_to_fr(d_sec, d_small, st);
        srs_secret_to_fr(d_sec + 32, d_small + 1, st);
        Fr* d_l = L->fr_a.as<Fr>(c->T);
        Fr* d_r = L->fr_b.as<Fr>(c->M);
        srs_lagrange_at(d_small, (int)c->t, d_l, d_small + 2, st);      // piano.rs:973-988
        srs_lagrange_at(d_small + 1, (int)c->m, d_r, d_small + 5, st);  // :956-971
        FB_CUDA(cudaMalloc(&c->d_u, nrows * NC * sizeof(G1Affine)));
        FB_CUDA(cudaMalloc(&c->d_g_tau_x, c->T * sizeof(G1Affine)));
        FB_CUDA(cudaMalloc(&c->d_g_tau_y, c->M * sizeof(G1Affine)));
        // U[i][j] = g^(R_i * L_j)  (:995-1007), in slabs of rows to bound the scalar scratch
        size_t slab = ((size_t)1 << 22) / NC;
        if (slab < 1) slab = 1;
        for (uint64_t r = 0; r < nrows; r += slab) {
            size_t rc2 = nrows - r < slab ? nrows - r : slab;
            Fr* d_prod = L->fr_c.as<Fr>(rc2 * NC);
            srs_outer(d_r + row0 + r, rc2, d_l + c->col0, NC, d_prod, st);
            g1_fixed_base_mul(d_prod, rc2 * NC, c->d_u + r * NC, L->pt_a, st);
        }
        if (col_sharded(c)) {  // a column shard does not need g^(tau_X^k) (only the setup file does)
            FB_CUDA(cudaMemsetAsync(c->d_g_tau_x, 0, c->T * sizeof(G1Affine), st));
            Fr* d_pw = L->fr_c.as<Fr>(c->M);
            srs_powers(d_small + 1, c->M, d_pw, st);
            g1_fixed_base_mul(d_pw, c->M, c->d_g_tau_y, L->pt_a, st);
        } else {
            Fr* d_pw = L->fr_c.as<Fr>(c->T > c->M ? c->T : c->M);
            srs_powers(d_small, c->T, d_pw, st);  // :916-931
            g1_fixed_base_mul(d_pw, c->T, c->d_g_tau_x, L->pt_a, st);
            srs_powers(d_small + 1, c->M, d_pw, st);  // :934-949
            g1_fixed_base_mul(d_pw, c->M, c->d_g_tau_y, L->pt_a, st);
        }
        c->g = g1_generator();
        FB_CUDA(cudaStreamSynchronize(st));
        g2_generate_host(sx, sy, c->g2_bytes);
        c->g2_compressed = 1;
        srs_build_tables(c, L);
        c->srs_ready = true;
        return FB_OK;
    });
}

int fb_srs_load_bytes(fb_ctx* c, const uint8_t* buf, size_t len, int compressed, uint64_t row0, uint64_t nrows) {
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane* L) -> int {
        int rc = check_rows(c, row0, nrows);
        if (rc) return rc;
        const size_t e1 = compressed ? 48 : 96, e2 = compressed ? 96 : 192;
        size_t pos = 0;
        auto need = [&](size_t k) { return pos + k <= len; };
        auto rd64 = [&](uint64_t& v) {
            if (!need(8)) return false;
            memcpy(&v, buf + pos, 8);  // u64 little-endian (piano.rs:671-673)
            pos += 8;
            return true;
        };
        const char* trunc = "failed to fill whole buffer";
        if (!need(e1)) return ctx_fail(c, FB_ERR_IO, trunc);
        size_t off_g = pos;
        pos += e1;
        uint64_t nx, ny, rows, cols;
        if (!rd64(nx) || !need(nx * e1)) return ctx_fail(c, FB_ERR_IO, trunc);
        size_t off_x = pos;
        pos += nx * e1;
        if (!rd64(ny) || !need(ny * e1)) return ctx_fail(c, FB_ERR_IO, trunc);
        size_t off_y = pos;
        pos += ny * e1;
        if (!rd64(rows) || !rd64(cols) || !need(rows * cols * e1)) return ctx_fail(c, FB_ERR_IO, trunc);
        size_t off_u = pos;
        pos += rows * cols * e1;
        if (!need(3 * e2)) return ctx_fail(c, FB_ERR_IO, trunc);
        if (nx != c->T || ny != c->M || rows != c->M || cols != c->T)
            return ctx_fail(c, FB_ERR_ARG, "setup dimensions do not match scale / machines_scale");
        srs_free(c);
        c->row0 = row0;
        c->nrows = nrows;
        c->col0 = c->want_col0;
        c->ncols = c->want_ncols ? c->want_ncols : c->T;
        cudaStream_t st = L->stream;
        ctx_reset_err(c, L);
        FB_CUDA(cudaMalloc(&c->d_u, nrows * c->ncols * sizeof(G1Affine)));
        FB_CUDA(cudaMalloc(&c->d_g_tau_x, c->T * sizeof(G1Affine)));
        FB_CUDA(cudaMalloc(&c->d_g_tau_y, c->M * sizeof(G1Affine)));
        auto decode = [&](size_t off, size_t count, G1Affine* dst) {
            // upload in slabs of <= 64 MiB and decode on the GPU
            const size_t slab = ((size_t)64 << 20) / e1;
            for (size_t k = 0; k < count; k += slab) {
                size_t cnt = count - k < slab ? count - k : slab;
                uint8_t* d_b = L->in_bytes.as<uint8_t>(cnt * e1 + 16);
                lane_h2d(c, L, d_b, buf + off + k * e1, cnt * e1);
                if (compressed) g1_decompress_batch(d_b, dst + k, cnt, L->d_err + 1, st);
                else g1_deserialize_batch(d_b, dst + k, cnt, L->d_err + 1, st);
            }
        };
        G1Affine* d_g = L->pt_b.as<G1Affine>(2);
        decode(off_g, 1, d_g);
        decode(off_x, c->T, c->d_g_tau_x);
        decode(off_y, c->M, c->d_g_tau_y);
        if (!col_sharded(c)) {
            decode(off_u + row0 * c->T * e1, nrows * c->T, c->d_u);
        } else {
            for (uint64_t r = 0; r < nrows; r++)
                decode(off_u + ((row0 + r) * c->T + c->col0) * e1, c->ncols, c->d_u + r * c->ncols);
        }
        FB_CUDA(cudaMemcpyAsync(&c->g, d_g, sizeof(G1Affine), cudaMemcpyDeviceToHost, st));
        rc = ctx_check_err(c, L);
        if (rc) {
            srs_free(c);
            return rc;
        }
        for (int k = 0; k < 3; k++) c->g2_bytes[k].assign(buf + pos + k * e2, buf + pos + (k + 1) * e2);
        c->g2_compressed = compressed ? 1 : 0;
        srs_build_tables(c, L);
        c->srs_ready = true;
        return FB_OK;
    });
}

int fb_srs_load_file(fb_ctx* c, const char* path, int compressed, uint64_t row0, uint64_t nrows) {
    if (!c || !path) return FB_ERR_ARG;
    std::vector<uint8_t> data;
    {
        std::ifstream f(path, std::ios::binary | std::ios::ate);
        if (!f) {
            return ctx_fail(c, FB_ERR_IO, std::string("cannot open ") + path);
        }
        std::streamsize sz = f.tellg();
        f.seekg(0);
        data.resize((size_t)sz);
        if (sz && !f.read((char*)data.data(), sz)) {
            return ctx_fail(c, FB_ERR_IO, std::string("cannot read ") + path);
        }
    }
    return fb_srs_load_bytes(c, data.data(), data.size(), compressed, row0, nrows);
}

int fb_srs_to_bytes(fb_ctx* c, int compressed, uint8_t* out, size_t cap, size_t* written) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
        if (c->row0 != 0 || c->nrows != c->M || col_sharded(c))
            return ctx_fail(c, FB_ERR_STATE, "saving needs every row and column resident");
        const size_t e1 = compressed ? 48 : 96, e2 = compressed ? 96 : 192;
        size_t total = e1 + 8 + c->T * e1 + 8 + c->M * e1 + 16 + c->M * c->T * e1 + 3 * e2;
        if (written) *written = total;
        if (!out) return FB_OK;
        if (cap < total) return ctx_fail(c, FB_ERR_ARG, "output buffer too small");
        cudaStream_t st = L->stream;
        size_t pos = 0;
        auto encode = [&](const G1Affine* src, size_t count) {
            const size_t slab = ((size_t)64 << 20) / e1;
            for (size_t k = 0; k < count; k += slab) {
                size_t cnt = count - k < slab ? count - k : slab;
                uint8_t* d_b = L->out_bytes.as<uint8_t>(cnt * e1);
                if (compressed) g1_compress_batch(src + k, d_b, cnt, st);
                else g1_serialize_batch(src + k, d_b, cnt, st);
                FB_CUDA(cudaMemcpyAsync(out + pos, d_b, cnt * e1, cudaMemcpyDeviceToHost, st));
                FB_CUDA(cudaStreamSynchronize(st));
                pos += cnt * e1;
            }
        };
        auto wr64 = [&](uint64_t v) {
            memcpy(out + pos, &v, 8);
            pos += 8;
        };
        G1Affine* d_g = L->pt_b.as<G1Affine>(2);
        FB_CUDA(cudaMemcpyAsync(d_g, &c->g, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
        encode(d_g, 1);
        wr64(c->T);
        encode(c->d_g_tau_x, c->T);
        wr64(c->M);
        encode(c->d_g_tau_y, c->M);
        wr64(c->M);
        wr64(c->T);
        encode(c->d_u, c->M * c->T);
        for (int k = 0; k < 3; k++) {
            std::vector<uint8_t> tmp;
            const std::vector<uint8_t>* src = &c->g2_bytes[k];
            if ((c->g2_compressed != 0) != (compressed != 0)) {
                if (!g2_recode_host(c->g2_bytes[k], compressed != 0, tmp)) return ctx_fail(c, FB_ERR_STATE, "bad G2 element");
                src = &tmp;
            }
            if (src->size() != e2) return ctx_fail(c, FB_ERR_STATE, "G2 elements unavailable");
            memcpy(out + pos, src->data(), e2);
            pos += e2;
        }
        return FB_OK;
    });
}

int fb_srs_save_file(fb_ctx* c, const char* path, int compressed) {
    if (!c || !path) return FB_ERR_ARG;
    size_t total = 0;
    int rc = fb_srs_to_bytes(c, compressed, nullptr, 0, &total);
    if (rc) return rc;
    std::vector<uint8_t> data(total);
    rc = fb_srs_to_bytes(c, compressed, data.data(), total, &total);
    if (rc) return rc;
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    if (!f || !f.write((const char*)data.data(), (std::streamsize)total)) {
        return ctx_fail(c, FB_ERR_IO, std::string("cannot write ") + path);
    }
    return FB_OK;
}

// ---------------------------------------------------------------- worker side
int fb_commit_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len, uint8_t* out48s) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        int rc = need_row(c, row0, nrows);
        if (rc) return rc;
        if ((rc = need_full_cols(c))) return rc;
        if (len > c->T) return ctx_fail(c, FB_ERR_ARG, "more coefficients than columns");
        if (nrows == 0) return FB_OK;
        ctx_reset_err(c, L);
        Fr* d_s = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, coeffs_be, nrows * len);
        G1Xyzz* d_r = L->small.as<G1Xyzz>(nrows + 1);
        lane_msm(c, L, row_bases(c, row0), d_s, len, nrows, d_r);  // piano.rs:150-158
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(48 * nrows);
        ctx_points_out(c, L, d_r, nrows, d_ob);
        FB_CUDA(cudaMemcpyAsync(out48s, d_ob, 48 * nrows, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

int fb_worker_commit(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, uint8_t out48[48]) {
    return fb_commit_rows(c, i, 1, coeffs_be, len, out48);
}

int fb_open_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                 uint8_t* out_ys_be, uint8_t* out_proof48s) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        int rc = need_row(c, row0, nrows);
        if (rc) return rc;
        if ((rc = need_full_cols(c))) return rc;
        if (len > c->T) return ctx_fail(c, FB_ERR_ARG, "Supplied list is longer than the available max width");
        int lg = ilog2_exact(len);
        if (lg < 0) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        if (nrows == 0) return FB_OK;
        ctx_reset_err(c, L);
        const NttTables& tab = ctx_ntt_tables(c, L, lg);
        // all rows' coefficients go up in one copy; rows are then processed in turn
        uint8_t* d_b = L->in_bytes.as<uint8_t>(nrows * len * 32 + 64);
        lane_h2d(c, L, d_b, coeffs_be, nrows * len * 32);
        Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
        Fr* d_all = (Fr*)L->eng_a.get((nrows * len + nrows + 1) * sizeof(Fr));
        Fr* d_ys = d_all + nrows * len;
        fr_from_be(d_b, d_all, nrows * len, L->d_err, L->stream);
        G1Xyzz* d_r = (G1Xyzz*)L->eng_b.get((nrows + 1) * sizeof(G1Xyzz));
        // rows are independent: one batched iNTT / scan / NTT / MSM over slabs of rows
        const uint64_t slab = 4096;
        for (uint64_t r = 0; r < nrows; r += slab) {
            const uint64_t rc2 = nrows - r < slab ? nrows - r : slab;
            Fr* d_qhat = open_scalars(c, L, tab, d_all + r * len, len, d_alpha, d_ys + r, rc2);
            lane_msm(c, L, row_bases(c, row0 + r), d_qhat, len, rc2, d_r + r);  // piano.rs:228-236
        }
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(80 * nrows);
        ctx_points_out(c, L, d_r, nrows, d_ob + 32 * nrows);
        fr_to_be(d_ys, d_ob, nrows, L->stream);
        FB_CUDA(cudaMemcpyAsync(out_ys_be, d_ob, 32 * nrows, cudaMemcpyDeviceToHost, L->stream));
        FB_CUDA(cudaMemcpyAsync(out_proof48s, d_ob + 32 * nrows, 48 * nrows, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

int fb_worker_open(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                   uint8_t out_y_be[32], uint8_t out_proof48[48]) {
    return fb_open_rows(c, i, 1, coeffs_be, len, alpha_be, out_y_be, out_proof48);
}

// ---------------------------------------------------------------- master side
int fb_master_commit(fb_ctx* c, const uint8_t* commitments48, uint64_t count, uint8_t out48[48]) {
    return fb_g1_sum(c, commitments48, count, 48, out48);  // all-ones MSM == point sum (piano.rs:164-175)
}

int fb_master_open(fb_ctx* c, const uint8_t* evals_be, const uint8_t* proofs48, uint64_t count, const uint8_t beta_be[32],
                   uint8_t out_z_be[32], uint8_t out_pi0_48[48], uint8_t out_pi1_48[48]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
        if (count > c->M) return ctx_fail(c, FB_ERR_ARG, "Supplied list is longer than the available max width");
        int lg = ilog2_exact(count);
        if (lg < 0) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        ctx_reset_err(c, L);
        cudaStream_t st = L->stream;
        // pi_0 = sum of the worker proofs (piano.rs:258-265)
        uint8_t* d_pb = L->in_bytes.as<uint8_t>(count * 48 + 16);
        lane_h2d(c, L, d_pb, proofs48, count * 48);
        G1Affine* pts = L->pt_a.as<G1Affine>(count + 1);
        g1_decompress_batch(d_pb, pts, count, L->d_err + 1, st);
        G1Xyzz* d_r = (G1Xyzz*)L->eng_b.get(4 * sizeof(G1Xyzz));
        g1_sum_affine(pts, count, d_r, L->scan, st);
        // f(Y, alpha) = iNTT_right(evals); z = f(beta); q1 = (f - z)/(Y - beta)  (:268-282)
        const NttTables& tab = ctx_ntt_tables(c, L, lg);
        Fr* d_ev = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, evals_be, count);
        Fr* d_beta = (Fr*)L->eng_a.get(4 * sizeof(Fr));
        {
            uint8_t* d_bb = L->small.as<uint8_t>(64);
            FB_CUDA(cudaMemcpyAsync(d_bb, beta_be, 32, cudaMemcpyHostToDevice, st));
            fr_from_be(d_bb, d_beta, 1, L->d_err, st);
        }
        Fr* d_z = d_beta + 1;
        Fr* d_poly = L->fr_b.as<Fr>(count);
        Fr* d_tmp = L->fr_c.as<Fr>(count);
        ntt_run(tab, c->tile_tw, d_ev, d_poly, d_tmp, true, st);
        horner_quotient(d_poly, count, d_beta, d_z, d_ev, L->scan, st);  // q (zero-padded) overwrites evals
        // pi_1 = sum_j q_j * g^(tau_Y^j)  (:286-295)
        msm_run(c->y_bases, d_ev, count, 1, true, d_r + 1, L->msm, st);
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(32 + 96);
        fr_to_be(d_z, d_ob, 1, st);
        ctx_points_out(c, L, d_r, 2, d_ob + 32);
        FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_ob, 128, cudaMemcpyDeviceToHost, st));
        int rc = ctx_check_err(c, L);
        if (rc) return rc;
        memcpy(out_z_be, L->h_pinned, 32);
        memcpy(out_pi0_48, L->h_pinned + 32, 48);
        memcpy(out_pi1_48, L->h_pinned + 80, 48);
        return FB_OK;
    });
}

// ------------------------------------------- partial (column-shard) + device calls
// out_fmt 48: compressed final result (whole-row contexts only); 96: uncompressed
// partial over the resident column range, to be added across shards (fb_g1_sum).
static void points_out_fmt(fb_ctx* c, fb_lane* L, const G1Xyzz* d_xyzz, uint8_t* d_out, int fmt) {
    if (fmt == 48) {
        ctx_points_out(c, L, d_xyzz, 1, d_out);
        return;
    }
    G1Affine* aff = L->pt_b.as<G1Affine>(2);
    prof_mark("out.affine_compress", L->stream);
    g1_xyzz_to_affine_batch(d_xyzz, aff, 1, L->stream);
    g1_serialize_batch(aff, d_out, 1, L->stream);
}

static int dev_commit(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_coeffs, uint64_t len, uint8_t* d_out, int fmt) {
    int rc = need_row(c, i, 1);
    if (rc) return rc;
    if (fmt == 48 && (rc = need_full_cols(c))) return rc;
    if (len > c->T) return ctx_fail(c, FB_ERR_ARG, "more coefficients than columns");
    G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
    row_msm(c, L, i, d_coeffs, len, d_r);
    points_out_fmt(c, L, d_r, d_out, fmt);
    return FB_OK;
}

// d_out: 32 bytes y (big-endian) then the point in `fmt`
static int dev_open(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_coeffs, uint64_t len, const Fr* d_alpha, uint8_t* d_out, int fmt) {
    int rc = need_row(c, i, 1);
    if (rc) return rc;
    if (fmt == 48 && (rc = need_full_cols(c))) return rc;
    int lg = ilog2_exact(len);
    if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
    const NttTables& tab = ctx_ntt_tables(c, L, lg);
    Fr* d_y = (Fr*)L->eng_a.get(4 * sizeof(Fr));
    Fr* d_qhat = open_scalars(c, L, tab, d_coeffs, len, d_alpha, d_y);
    G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
    row_msm(c, L, i, d_qhat, len, d_r);
    fr_to_be(d_y, d_out, 1, L->stream);
    points_out_fmt(c, L, d_r, d_out + 32, fmt);
    return FB_OK;
}

// batched rows, operands in HBM: d_coeffs = nrows x len Montgomery Fr
int fb_dev_commit_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const void* d_coeffs, uint64_t len, void* d_out48s) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        int rc = need_row(c, row0, nrows);
        if (rc) return rc;
        if ((rc = need_full_cols(c))) return rc;
        if (len > c->T) return ctx_fail(c, FB_ERR_ARG, "more coefficients than columns");
        if (nrows == 0) return FB_OK;
        G1Xyzz* d_r = (G1Xyzz*)L->eng_b.get((nrows + 1) * sizeof(G1Xyzz));
        lane_msm(c, L, row_bases(c, row0), (const Fr*)d_coeffs, len, nrows, d_r);
        ctx_points_out(c, L, d_r, nrows, (uint8_t*)d_out48s);
        return FB_OK;
    });
}

// d_out_ys: nrows x 32 bytes (big-endian), d_out_proofs: nrows x 48 bytes
int fb_dev_open_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const void* d_coeffs, uint64_t len, const void* d_alpha,
                     void* d_out_ys, void* d_out_proofs) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        int rc = need_row(c, row0, nrows);
        if (rc) return rc;
        if ((rc = need_full_cols(c))) return rc;
        int lg = ilog2_exact(len);
        if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        if (nrows == 0) return FB_OK;
        const NttTables& tab = ctx_ntt_tables(c, L, lg);
        Fr* d_ys = (Fr*)L->eng_a.get((nrows + 1) * sizeof(Fr));
        G1Xyzz* d_r = (G1Xyzz*)L->eng_b.get((nrows + 1) * sizeof(G1Xyzz));
        const uint64_t slab = 4096;
        for (uint64_t r = 0; r < nrows; r += slab) {
            const uint64_t rc2 = nrows - r < slab ? nrows - r : slab;
            Fr* d_qhat = open_scalars(c, L, tab, (const Fr*)d_coeffs + r * len, len, (const Fr*)d_alpha, d_ys + r, rc2);
            lane_msm(c, L, row_bases(c, row0 + r), d_qhat, len, rc2, d_r + r);
        }
        fr_to_be(d_ys, (uint8_t*)d_out_ys, nrows, L->stream);
        ctx_points_out(c, L, d_r, nrows, (uint8_t*)d_out_proofs);
        return FB_OK;
    });
}

int fb_ctx_set_col_shard(fb_ctx* c, uint64_t col0, uint64_t ncols) {
    if (!c) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (col0 + ncols > c->T) return ctx_fail(c, FB_ERR_ARG, "column range exceeds 2^(scale - machines_scale)");
    c->want_col0 = ncols ? col0 : 0;
    c->want_ncols = ncols;
    return FB_OK;
}

int fb_dev_worker_commit(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, void* d_out48) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int { return dev_commit(c, L, i, (const Fr*)d_coeffs, len, (uint8_t*)d_out48, 48); });
}
int fb_dev_worker_commit_partial(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, void* d_out96) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int { return dev_commit(c, L, i, (const Fr*)d_coeffs, len, (uint8_t*)d_out96, 96); });
}
int fb_dev_worker_open(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, const void* d_alpha, void* d_out80) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        return dev_open(c, L, i, (const Fr*)d_coeffs, len, (const Fr*)d_alpha, (uint8_t*)d_out80, 48);
    });
}
int fb_dev_worker_open_partial(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, const void* d_alpha,
                               void* d_out128) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        return dev_open(c, L, i, (const Fr*)d_coeffs, len, (const Fr*)d_alpha, (uint8_t*)d_out128, 96);
    });
}

int fb_worker_commit_partial(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, uint8_t out96[96]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        ctx_reset_err(c, L);
        // only the resident column range of the coefficients takes part in this shard's MSM: upload
        // that slice (at its own offset, so the indices line up) instead of the whole row
        // NB: scalars outside the slice are therefore not range-checked by this shard; the shard
        // that owns them rejects them ("Invalid scalar").
        const uint64_t lo = c->col0 < len ? c->col0 : len;
        const uint64_t hi = c->col0 + c->ncols < len ? c->col0 + c->ncols : len;
        Fr* d_s = L->fr_a.as<Fr>(len + 1);
        if (hi > lo) {
            uint8_t* d_b = L->in_bytes.as<uint8_t>((hi - lo) * 32 + 32);
            prof_mark("h2d", L->stream);
            lane_h2d(c, L, d_b, coeffs_be + lo * 32, (hi - lo) * 32);
            fr_from_be(d_b, d_s + lo, hi - lo, L->d_err, L->stream);
        }
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(128);
        int rc = dev_commit(c, L, i, d_s, len, d_ob, 96);
        if (rc) return rc;
        FB_CUDA(cudaMemcpyAsync(out96, d_ob, 96, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

int fb_worker_open_partial(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                           uint8_t out_y_be[32], uint8_t out96[96]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        ctx_reset_err(c, L);
        Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
        Fr* d_s = (Fr*)L->eng_b.get((len + 1) * sizeof(Fr));
        uint8_t* d_b = L->in_bytes.as<uint8_t>(len * 32 + 64);
        lane_h2d(c, L, d_b, coeffs_be, len * 32);
        fr_from_be(d_b, d_s, len, L->d_err, L->stream);
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(128);
        int rc = dev_open(c, L, i, d_s, len, d_alpha, d_ob, 96);
        if (rc) return rc;
        FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_ob, 128, cudaMemcpyDeviceToHost, L->stream));
        rc = ctx_check_err(c, L);
        if (rc) return rc;
        memcpy(out_y_be, L->h_pinned, 32);
        memcpy(out96, L->h_pinned + 32, 96);
        return FB_OK;
    });
}

// ---- commit + open of one row in ONE pass (an extension over the reference's two calls, which stay):
// both MSMs run over the same resident row, so the two scalar vectors (the evaluations and the
// quotient's evaluations) go through msm_run as a batch of two over shared bases and the coefficients
// are uploaded once.  Same bytes as fb_worker_commit + fb_worker_open.
// d_out: point | 32 bytes y | point, points in `fmt` (48 compressed / 96 uncompressed partial)
static int dev_commit_open(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_coeffs, uint64_t len, const Fr* d_alpha, uint8_t* d_out,
                           int fmt) {
    int rc = need_row(c, i, 1);
    if (rc) return rc;
    if (fmt == 48 && (rc = need_full_cols(c))) return rc;
    int lg = ilog2_exact(len);
    if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
    const NttTables& tab = ctx_ntt_tables(c, L, lg);
    Fr* d_y = (Fr*)L->eng_a.get(4 * sizeof(Fr));
    G1Xyzz* d_r = L->small.as<G1Xyzz>(3);
    const size_t lo = c->col0 < len ? c->col0 : len;
    const size_t hi = c->col0 + c->ncols < len ? c->col0 + c->ncols : len;
    const MsmBases bases = row_bases(c, i);
    const bool forked = lane_pair_fork(c, L, bases, d_coeffs + lo, hi - lo, d_r);  // piano.rs:150-158 (may start now)
    Fr* d_qhat = open_scalars(c, L, tab, d_coeffs, len, d_alpha, d_y);
    lane_pair_join(c, L, forked, bases, d_coeffs + lo, d_qhat + lo, hi - lo, d_r);  // :228-236
    fr_to_be(d_y, d_out + fmt, 1, L->stream);
    G1Affine* aff = L->pt_b.as<G1Affine>(3);
    prof_mark("out.affine_compress", L->stream);
    g1_xyzz_to_affine_batch(d_r, aff, 2, L->stream);
    if (fmt == 48) {
        g1_compress_batch(aff, d_out, 1, L->stream);
        g1_compress_batch(aff + 1, d_out + 48 + 32, 1, L->stream);
    } else {
        g1_serialize_batch(aff, d_out, 1, L->stream);
        g1_serialize_batch(aff + 1, d_out + 96 + 32, 1, L->stream);
    }
    return FB_OK;
}

int fb_dev_worker_commit_open(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, const void* d_alpha, void* d_out128) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        return dev_commit_open(c, L, i, (const Fr*)d_coeffs, len, (const Fr*)d_alpha, (uint8_t*)d_out128, 48);
    });
}
int fb_dev_worker_commit_open_partial(fb_ctx* c, uint64_t i, const void* d_coeffs, uint64_t len, const void* d_alpha,
                                      void* d_out224) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        return dev_commit_open(c, L, i, (const Fr*)d_coeffs, len, (const Fr*)d_alpha, (uint8_t*)d_out224, 96);
    });
}

static int host_commit_open(fb_ctx* c, fb_lane* L, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                            uint8_t* out_commit, uint8_t out_y_be[32], uint8_t* out_proof, int fmt) {
    ctx_reset_err(c, L);
    Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
    Fr* d_s = (Fr*)L->eng_b.get((len + 1) * sizeof(Fr));
    uint8_t* d_b = L->in_bytes.as<uint8_t>(len * 32 + 64);
    prof_mark("h2d", L->stream);
    lane_h2d(c, L, d_b, coeffs_be, len * 32);
    fr_from_be(d_b, d_s, len, L->d_err, L->stream);
    uint8_t* d_ob = L->out_bytes.as<uint8_t>(256);
    int rc = dev_commit_open(c, L, i, d_s, len, d_alpha, d_ob, fmt);
    if (rc) return rc;
    FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_ob, 2 * fmt + 32, cudaMemcpyDeviceToHost, L->stream));
    rc = ctx_check_err(c, L);
    if (rc) return rc;
    memcpy(out_commit, L->h_pinned, fmt);
    memcpy(out_y_be, L->h_pinned + fmt, 32);
    memcpy(out_proof, L->h_pinned + fmt + 32, fmt);
    return FB_OK;
}

int fb_worker_commit_open(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                          uint8_t out_commit48[48], uint8_t out_y_be[32], uint8_t out_proof48[48]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        return host_commit_open(c, L, i, coeffs_be, len, alpha_be, out_commit48, out_y_be, out_proof48, 48);
    });
}
int fb_worker_commit_open_partial(fb_ctx* c, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                                  uint8_t out_commit96[96], uint8_t out_y_be[32], uint8_t out_proof96[96]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        return host_commit_open(c, L, i, coeffs_be, len, alpha_be, out_commit96, out_y_be, out_proof96, 96);
    });
}

// ---- commit + open of one row sharded by columns, field half included.
// With fb_worker_commit_open_partial every shard still runs the field half over the WHOLE row (and uploads it).  Here a
// shard works on its slice only, in two halves around one exchange of 128 bytes per shard:
//   begin   inv_j = 1/(w^j - alpha), h_j = w^j inv_j, partial S1 = sum c_j h_j and S0 = sum h_j over the slice -> state
//   (the caller gathers the states of all shards, e.g. one all-gather of 128 bytes)
//   finish  S1, S0, y from all states; q_hat_j over the slice; commit and proof partials = one batch-of-two MSM
// The slice is columns [col0, col0 + ncols) cut to the row length `len`: d_slice holds exactly those coefficients.
static void shard_span(fb_ctx* c, uint64_t len, uint64_t& lo, uint64_t& cnt) {
    lo = c->col0 < len ? c->col0 : len;
    const uint64_t hi = c->col0 + c->ncols < len ? c->col0 + c->ncols : len;
    cnt = hi - lo;
}

static int dev_shard_begin(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_slice, uint64_t len, const Fr* d_alpha, Fr* d_state) {
    int rc = need_row(c, i, 1);
    if (rc) return rc;
    int lg = ilog2_exact(len);
    if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
    uint64_t lo, cnt;
    shard_span(c, len, lo, cnt);
    const NttTables& tab = ctx_ntt_tables(c, L, lg);
    evalform_shard_begin(tab, d_slice, len, lo, cnt, d_alpha, d_state, c->shard_buf, L->stream);
    return FB_OK;
}

// d_out: 96 bytes partial commitment | 32 bytes y | 96 bytes partial proof
static int dev_shard_finish(fb_ctx* c, fb_lane* L, uint64_t i, const Fr* d_slice, uint64_t len, const Fr* d_alpha,
                            const Fr* d_states, uint64_t nparts, uint8_t* d_out) {
    int rc = need_row(c, i, 1);
    if (rc) return rc;
    int lg = ilog2_exact(len);
    if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
    if (nparts == 0 || nparts > 4096) return ctx_fail(c, FB_ERR_ARG, "number of shard states out of range");
    uint64_t lo, cnt;
    shard_span(c, len, lo, cnt);
    const NttTables& tab = ctx_ntt_tables(c, L, lg);
    Fr* d_y = (Fr*)L->eng_a.get(4 * sizeof(Fr));
    Fr* d_qhat = L->fr_b.as<Fr>(cnt + 1);
    G1Xyzz* d_r = L->small.as<G1Xyzz>(3);
    const MsmBases bases = row_bases(c, i);
    const bool forked = lane_pair_fork(c, L, bases, d_slice, cnt, d_r);
    evalform_shard_finish(tab, d_slice, len, lo, cnt, d_alpha, d_states, (int)nparts, d_y, d_qhat, c->shard_buf, L->stream);
    lane_pair_join(c, L, forked, bases, d_slice, d_qhat, cnt, d_r);
    fr_to_be(d_y, d_out + 96, 1, L->stream);
    G1Affine* aff = L->pt_b.as<G1Affine>(3);
    prof_mark("out.affine_compress", L->stream);
    g1_xyzz_to_affine_batch(d_r, aff, 2, L->stream);
    g1_serialize_batch(aff, d_out, 1, L->stream);
    g1_serialize_batch(aff + 1, d_out + 128, 1, L->stream);
    return FB_OK;
}

int fb_dev_commit_open_shard_begin(fb_ctx* c, uint64_t i, const void* d_coeffs_slice, uint64_t len, const void* d_alpha,
                                   void* d_out_state128) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        return dev_shard_begin(c, L, i, (const Fr*)d_coeffs_slice, len, (const Fr*)d_alpha, (Fr*)d_out_state128);
    });
}
int fb_dev_commit_open_shard_finish(fb_ctx* c, uint64_t i, const void* d_coeffs_slice, uint64_t len, const void* d_alpha,
                                    const void* d_states, uint64_t nparts, void* d_out224) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        return dev_shard_finish(c, L, i, (const Fr*)d_coeffs_slice, len, (const Fr*)d_alpha, (const Fr*)d_states, nparts,
                                (uint8_t*)d_out224);
    });
}

// host forms: the slice's coefficients (wire bytes) go up once in `begin` and stay in the context until `finish`
int fb_worker_commit_open_shard_begin(fb_ctx* c, uint64_t i, const uint8_t* coeffs_slice_be, uint64_t slice_len, uint64_t len,
                                      const uint8_t alpha_be[32], uint8_t out_state128[128]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        uint64_t lo, cnt;
        shard_span(c, len, lo, cnt);
        if (slice_len != cnt) return ctx_fail(c, FB_ERR_ARG, "the slice must hold exactly the resident columns of the row");
        ctx_reset_err(c, L);
        Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
        Fr* d_s = c->shard_coeffs.as<Fr>(cnt + 1);
        if (cnt) {
            uint8_t* d_b = L->in_bytes.as<uint8_t>(cnt * 32 + 64);
            prof_mark("h2d", L->stream);
            lane_h2d(c, L, d_b, coeffs_slice_be, cnt * 32);
            fr_from_be(d_b, d_s, cnt, L->d_err, L->stream);
        }
        Fr* d_state = (Fr*)L->eng_a.get(8 * sizeof(Fr));
        int rc = dev_shard_begin(c, L, i, d_s, len, d_alpha, d_state + 4);
        if (rc) return rc;
        c->shard_len = len;
        c->shard_cnt = cnt;
        FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_state + 4, 128, cudaMemcpyDeviceToHost, L->stream));
        rc = ctx_check_err(c, L);
        if (rc) return rc;
        memcpy(out_state128, L->h_pinned, 128);
        return FB_OK;
    });
}

int fb_worker_commit_open_shard_finish(fb_ctx* c, uint64_t i, const uint8_t* states128, uint64_t nparts, const uint8_t alpha_be[32],
                                       uint8_t out_commit96[96], uint8_t out_y_be[32], uint8_t out_proof96[96]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!c->shard_len) return ctx_fail(c, FB_ERR_STATE, "no sharded open in flight: call fb_worker_commit_open_shard_begin first");
        if (nparts == 0 || nparts > 4096) return ctx_fail(c, FB_ERR_ARG, "number of shard states out of range");
        ctx_reset_err(c, L);
        Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
        uint8_t* d_st = L->in_bytes.as<uint8_t>(nparts * 128 + 64);
        FB_CUDA(cudaMemcpyAsync(d_st, states128, nparts * 128, cudaMemcpyHostToDevice, L->stream));
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(256);
        int rc = dev_shard_finish(c, L, i, (const Fr*)c->shard_coeffs.p, c->shard_len, d_alpha, (const Fr*)d_st, nparts, d_ob);
        c->shard_len = 0;
        if (rc) return rc;
        FB_CUDA(cudaMemcpyAsync(L->h_pinned, d_ob, 224, cudaMemcpyDeviceToHost, L->stream));
        rc = ctx_check_err(c, L);
        if (rc) return rc;
        memcpy(out_commit96, L->h_pinned, 96);
        memcpy(out_y_be, L->h_pinned + 96, 32);
        memcpy(out_proof96, L->h_pinned + 128, 96);
        return FB_OK;
    });
}

// the last step of every column-sharded commit + open: add the shards' partials.  d_parts = nparts x (96-byte partial
// commitment | 32-byte y | 96-byte partial proof) as gathered from the shards; d_out = 48 | 32 | 48 (final, compressed)
int fb_dev_combine_commit_open(fb_ctx* c, const void* d_parts224, uint64_t nparts, void* d_out128) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        if (nparts == 0 || nparts > 4096) return ctx_fail(c, FB_ERR_ARG, "number of parts out of range");
        prof_mark("out.combine_parts", L->stream);
        g1_combine_parts((const uint8_t*)d_parts224, (size_t)nparts, (uint8_t*)d_out128, L->d_err + 1, L->stream);
        return FB_OK;
    });
}

// ---- commit + open of a block of rows in one call (the bivariate row loop of src/bipoly.rs / benches/piano.rs:41-63 does
// worker_commit and worker_open of every row): the coefficient matrix goes up once, and the commit MSMs (a batch over the
// rows) run on a sub-stream of the lane while the field half and the proof MSMs run on the lane's stream.
static int dev_commit_open_rows(fb_ctx* c, fb_lane* L, uint64_t row0, uint64_t nrows, const Fr* d_all, uint64_t len,
                                const Fr* d_alpha, uint8_t* d_c48, uint8_t* d_ys_be, uint8_t* d_p48) {
    int rc = need_row(c, row0, nrows);
    if (rc) return rc;
    if ((rc = need_full_cols(c))) return rc;
    int lg = ilog2_exact(len);
    if (lg < 0 || len > c->T) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
    if (nrows == 0) return FB_OK;
    const NttTables& tab = ctx_ntt_tables(c, L, lg);
    Fr* d_ys = (Fr*)L->eng_a.get((nrows + 1) * sizeof(Fr));
    G1Xyzz* d_rc = (G1Xyzz*)L->eng_b.get((2 * nrows + 2) * sizeof(G1Xyzz));
    G1Xyzz* d_rp = d_rc + nrows;
    const bool profiling = g_prof && g_prof->on;
    const bool forked = !profiling && c->msm_split <= 1 && c->pair_mode != 1;
    if (forked) {
        FB_CUDA(cudaEventRecord(L->sub_fork, L->stream));
        FB_CUDA(cudaStreamWaitEvent(L->sub_stream[0], L->sub_fork, 0));
        msm_run(row_bases(c, row0), d_all, len, nrows, true, d_rc, L->msm_sub[0], L->sub_stream[0]);  // piano.rs:150-158
        FB_CUDA(cudaEventRecord(L->sub_join[0], L->sub_stream[0]));
    } else {
        lane_msm(c, L, row_bases(c, row0), d_all, len, nrows, d_rc);
    }
    const uint64_t slab = 4096;
    for (uint64_t r = 0; r < nrows; r += slab) {
        const uint64_t rc2 = nrows - r < slab ? nrows - r : slab;
        Fr* d_qhat = open_scalars(c, L, tab, d_all + r * len, len, d_alpha, d_ys + r, rc2);
        lane_msm(c, L, row_bases(c, row0 + r), d_qhat, len, rc2, d_rp + r);  // :228-236
    }
    if (forked) FB_CUDA(cudaStreamWaitEvent(L->stream, L->sub_join[0], 0));
    fr_to_be(d_ys, d_ys_be, nrows, L->stream);
    ctx_points_out(c, L, d_rc, nrows, d_c48);
    ctx_points_out(c, L, d_rp, nrows, d_p48);
    return FB_OK;
}

int fb_dev_commit_open_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const void* d_coeffs, uint64_t len, const void* d_alpha,
                            void* d_out_commit48s, void* d_out_ys, void* d_out_proof48s) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        return dev_commit_open_rows(c, L, row0, nrows, (const Fr*)d_coeffs, len, (const Fr*)d_alpha, (uint8_t*)d_out_commit48s,
                                    (uint8_t*)d_out_ys, (uint8_t*)d_out_proof48s);
    });
}

int fb_commit_open_rows(fb_ctx* c, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                        uint8_t* out_commit48s, uint8_t* out_ys_be, uint8_t* out_proof48s) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (nrows == 0) return FB_OK;
        if (len > c->T) return ctx_fail(c, FB_ERR_ARG, "Supplied list is longer than the available max width");
        ctx_reset_err(c, L);
        uint8_t* d_b = L->in_bytes.as<uint8_t>(nrows * len * 32 + 64);
        prof_mark("h2d", L->stream);
        lane_h2d(c, L, d_b, coeffs_be, nrows * len * 32);
        Fr* d_alpha = ctx_upload_fr(c, L, L->small, L->pt_a, alpha_be, 1);
        Fr* d_all = L->fr_a.as<Fr>(nrows * len + 1);
        fr_from_be(d_b, d_all, nrows * len, L->d_err, L->stream);
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(128 * nrows);
        int rc = dev_commit_open_rows(c, L, row0, nrows, d_all, len, d_alpha, d_ob, d_ob + 48 * nrows, d_ob + 80 * nrows);
        if (rc) return rc;
        FB_CUDA(cudaMemcpyAsync(out_commit48s, d_ob, 48 * nrows, cudaMemcpyDeviceToHost, L->stream));
        FB_CUDA(cudaMemcpyAsync(out_ys_be, d_ob + 48 * nrows, 32 * nrows, cudaMemcpyDeviceToHost, L->stream));
        FB_CUDA(cudaMemcpyAsync(out_proof48s, d_ob + 80 * nrows, 48 * nrows, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

}  // extern "C"
