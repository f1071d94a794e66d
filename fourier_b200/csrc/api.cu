// api.cu - the C ABI (include/fourier_b200.h): context, field-side calls, raw
// kernels and device-resident entry points.  Engine calls (worker_* / master_* /
// SRS) are in engine.cu.
#include <string.h>

#include "ctx.cuh"

namespace fb {

std::atomic<unsigned long long> g_launches{0};
thread_local Prof* g_prof = nullptr;

int ctx_fail(fb_ctx* c, int code, const std::string& msg) {
    c->err = msg;
    return code;
}

const NttTables& ctx_ntt_tables(fb_ctx* c, int log_n) {
    if (!c->tile_tw_ready) {
        ntt_build_tile_twiddles(c->tile_tw, c->stream);
        c->tile_tw_ready = true;
    }
    auto it = c->ntt.find(log_n);
    if (it == c->ntt.end()) {
        NttTables t;
        ntt_build_tables(t, log_n, c->stream);
        it = c->ntt.emplace(log_n, t).first;
    }
    return it->second;
}

// host bytes -> device Fr (Montgomery); returns FB_ERR_SCALAR lazily via d_err
Fr* ctx_upload_fr(fb_ctx* c, DevBuf& bytes, DevBuf& dst, const uint8_t* h_be, size_t count) {
    uint8_t* d_b = bytes.as<uint8_t>(count * 32 + 32);
    Fr* d = dst.as<Fr>(count + 1);
    if (count) {
        prof_mark("h2d", c->stream);
        FB_CUDA(cudaMemcpyAsync(d_b, h_be, count * 32, cudaMemcpyHostToDevice, c->stream));
        fr_from_be(d_b, d, count, c->d_err, c->stream);
    }
    return d;
}

void ctx_reset_err(fb_ctx* c) { FB_CUDA(cudaMemsetAsync(c->d_err, 0, 2 * sizeof(int), c->stream)); }

// read the two error flags (scalar, point) after a sync
int ctx_check_err(fb_ctx* c) {
    int h[2];
    FB_CUDA(cudaMemcpyAsync(h, c->d_err, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    FB_CUDA(cudaStreamSynchronize(c->stream));
    if (h[0]) return ctx_fail(c, FB_ERR_SCALAR, "Invalid scalar");
    if (h[1]) return ctx_fail(c, FB_ERR_POINT, "Failed to uncompress G1 point (code " + std::to_string(h[1]) + ")");
    return FB_OK;
}

// XYZZ results -> 48-byte compressed strings on the device
void ctx_points_out(fb_ctx* c, const G1Xyzz* d_xyzz, size_t count, uint8_t* d_out48) {
    G1Affine* aff = c->pt_b.as<G1Affine>(count + 1);
    prof_mark("out.affine_compress", c->stream);
    g1_xyzz_to_affine_batch(d_xyzz, aff, count, c->stream);
    g1_compress_batch(aff, d_out48, count, c->stream);
}

}  // namespace fb

using namespace fb;

extern "C" {

const char* fb_version(void) { return "fourier_b200 0.1 (sm_100a)"; }

int fb_ctx_create(uint32_t n, uint32_t m, int device, fb_ctx** out) {
    if (!out) return FB_ERR_ARG;
    *out = nullptr;
    if (m > n || n > 30) return FB_ERR_ARG;  // "m must be less than or equal to n" (piano.rs:1062-1064)
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return FB_ERR_CUDA;
    fb_ctx* c = new fb_ctx();
    c->device = device;
    c->n = n;
    c->m = m;
    c->t = n - m;
    c->T = (size_t)1 << c->t;
    c->M = (size_t)1 << m;
    try {
        FB_CUDA(cudaSetDevice(device));
        FB_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
        c->stream = c->own_stream;
        FB_CUDA(cudaMalloc(&c->d_err, 2 * sizeof(int)));
        FB_CUDA(cudaMemset(c->d_err, 0, 2 * sizeof(int)));
        c->h_pinned_cap = 1 << 16;
        FB_CUDA(cudaMallocHost(&c->h_pinned, c->h_pinned_cap));
    } catch (const CudaError&) {
        delete c;
        return FB_ERR_CUDA;
    }
    *out = c;
    return FB_OK;
}

void fb_ctx_destroy(fb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (auto& kv : c->ntt) ntt_free_tables(kv.second);
    for (int d = 0; d < 2; d++)
        if (c->tile_tw.base[d]) cudaFree(c->tile_tw.base[d]);
    if (c->d_g_tau_x) cudaFree(c->d_g_tau_x);
    if (c->d_g_tau_y) cudaFree(c->d_g_tau_y);
    if (c->u_bases.pts && c->u_bases.pts != c->d_u) cudaFree(c->u_bases.pts);
    if (c->y_bases.pts && c->y_bases.pts != c->d_g_tau_y) cudaFree(c->y_bases.pts);
    if (c->d_u) cudaFree(c->d_u);
    DevBuf* bufs[] = {&c->msm.counts, &c->msm.offsets, &c->msm.cursor, &c->msm.sorted, &c->msm.buckets, &c->msm.block_sums,
                      &c->msm.seg_a, &c->msm.seg_t, &c->msm.order, &c->msm.misc, &c->msm.pts_a, &c->msm.pts_b, &c->msm.pref,
                      &c->msm.where, &c->msm.stage, &c->msm.inv, &c->in_bytes, &c->fr_a, &c->fr_b, &c->fr_c,
                      &c->pt_a, &c->pt_b, &c->out_bytes, &c->scan, &c->small, &c->eng_a, &c->eng_b};
    for (DevBuf* b : bufs) b->release();
    if (c->d_err) cudaFree(c->d_err);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

const char* fb_last_error(const fb_ctx* c) { return c ? c->err.c_str() : "null context"; }

int fb_ctx_set_stream(fb_ctx* c, void* s) {
    return ctx_call(c, [&]() -> int {
        FB_CUDA(cudaStreamSynchronize(c->stream));
        c->stream = s ? (cudaStream_t)s : c->own_stream;
        return FB_OK;
    });
}

int fb_ctx_sync(fb_ctx* c) {
    return ctx_call(c, [&]() -> int {
        FB_CUDA(cudaStreamSynchronize(c->stream));
        return FB_OK;
    });
}

int fb_ctx_set_window(fb_ctx* c, int w) {
    if (!c || w < -1 || (w > 0 && w < 4) || w > 24) return FB_ERR_ARG;
    c->window_pref = w;
    return FB_OK;
}

uint64_t fb_launch_count(const fb_ctx*) { return g_launches.load(); }

int fb_prof_enable(fb_ctx* c, int on) {
    if (!c) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->prof.on = on != 0;
    return FB_OK;
}

// JSON object {"phase": [total_ms, count], ...} of everything marked since the last read
int fb_prof_read(fb_ctx* c, char* buf, size_t cap) {
    if (!c || !buf || cap < 3) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    std::map<std::string, std::pair<double, int>> acc;
    auto& m = c->prof.marks;
    for (size_t i = 0; i + 1 < m.size(); i++) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, m[i].second, m[i + 1].second) != cudaSuccess) ms = 0;
        auto& a = acc[m[i].first];
        a.first += ms;
        a.second += 1;
    }
    for (auto& kv : m) c->prof.pool.push_back(kv.second);
    m.clear();
    cudaGetLastError();
    std::string out = "{";
    bool first = true;
    for (auto& kv : acc) {
        if (kv.first == "idle") continue;
        char tmp[160];
        snprintf(tmp, sizeof tmp, "%s\"%s\": [%.6f, %d]", first ? "" : ", ", kv.first.c_str(), kv.second.first, kv.second.second);
        out += tmp;
        first = false;
    }
    out += "}";
    if (out.size() + 1 > cap) return FB_ERR_ARG;
    memcpy(buf, out.c_str(), out.size() + 1);
    return FB_OK;
}

// ------------------------------------------------------------ field-side calls
static int do_ntt_host(fb_ctx* c, int log_n, bool inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    ctx_reset_err(c);
    const NttTables& tab = ctx_ntt_tables(c, log_n);
    Fr* d_in = ctx_upload_fr(c, c->in_bytes, c->fr_a, in_be, len);
    Fr* d_out = c->fr_b.as<Fr>(len);
    Fr* d_tmp = c->fr_c.as<Fr>(len);
    ntt_run(tab, c->tile_tw, d_in, d_out, d_tmp, inverse, c->stream);
    uint8_t* d_ob = c->out_bytes.as<uint8_t>(len * 32);
    fr_to_be(d_out, d_ob, len, c->stream);
    FB_CUDA(cudaMemcpyAsync(out_be, d_ob, len * 32, cudaMemcpyDeviceToHost, c->stream));
    return ctx_check_err(c);
}

int fb_ntt(fb_ctx* c, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    return ctx_call(c, [&]() -> int {
        int lg = ilog2_exact(len);
        if (lg < 0 || lg > 30) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        return do_ntt_host(c, lg, inverse != 0, in_be, len, out_be);
    });
}

int fb_fft(fb_ctx* c, int left, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    return ctx_call(c, [&]() -> int {
        size_t width = left ? c->T : c->M;
        if (len > width) return ctx_fail(c, FB_ERR_ARG, "Supplied list is longer than the available max width");
        int lg = ilog2_exact(len);
        if (lg < 0) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        return do_ntt_host(c, lg, inverse != 0, in_be, len, out_be);
    });
}

int fb_evaluate(fb_ctx* c, const uint8_t* poly_be, uint64_t len, const uint8_t x_be[32], uint8_t out_y_be[32]) {
    return ctx_call(c, [&]() -> int {
        ctx_reset_err(c);
        Fr* d_p = ctx_upload_fr(c, c->in_bytes, c->fr_a, poly_be, len);
        Fr* d_x = ctx_upload_fr(c, c->small, c->fr_c, x_be, 1);
        Fr* d_y = d_x + 1;  // fr_c holds count+1 elements
        horner_quotient(d_p, len, d_x, d_y, nullptr, c->scan, c->stream);
        uint8_t* d_ob = c->out_bytes.as<uint8_t>(32);
        fr_to_be(d_y, d_ob, 1, c->stream);
        FB_CUDA(cudaMemcpyAsync(out_y_be, d_ob, 32, cudaMemcpyDeviceToHost, c->stream));
        return ctx_check_err(c);
    });
}

// ----------------------------------------------------------------- raw kernels
static int bases_finish(fb_ctx* c, G1Affine* pts, uint64_t count, fb_bases** out) {
    fb_bases* b = new fb_bases();
    b->raw = pts;
    b->b.pts = pts;
    b->b.n = count;
    b->b.c = 0;
    b->b.nwin = 0;
    *out = b;
    return FB_OK;
}

int fb_bases_from_points(fb_ctx* c, const uint8_t* points, uint64_t count, int format, fb_bases** out) {
    return ctx_call(c, [&]() -> int {
        if (!out || (format != 48 && format != 96) || count == 0) return ctx_fail(c, FB_ERR_ARG, "bad base-set arguments");
        ctx_reset_err(c);
        uint8_t* d_b = c->in_bytes.as<uint8_t>(count * format);
        FB_CUDA(cudaMemcpyAsync(d_b, points, count * format, cudaMemcpyHostToDevice, c->stream));
        G1Affine* pts;
        FB_CUDA(cudaMalloc(&pts, count * sizeof(G1Affine)));
        if (format == 48) g1_decompress_batch(d_b, pts, count, c->d_err + 1, c->stream);
        else g1_deserialize_batch(d_b, pts, count, c->d_err + 1, c->stream);
        int rc = ctx_check_err(c);
        if (rc) {
            cudaFree(pts);
            return rc;
        }
        return bases_finish(c, pts, count, out);
    });
}

int fb_bases_from_scalars(fb_ctx* c, const uint8_t* scalars_be, uint64_t count, fb_bases** out) {
    return ctx_call(c, [&]() -> int {
        if (!out || count == 0) return ctx_fail(c, FB_ERR_ARG, "bad base-set arguments");
        ctx_reset_err(c);
        Fr* d_s = ctx_upload_fr(c, c->in_bytes, c->fr_a, scalars_be, count);
        G1Affine* pts;
        FB_CUDA(cudaMalloc(&pts, count * sizeof(G1Affine)));
        g1_fixed_base_mul(d_s, count, pts, c->pt_a, c->stream);
        int rc = ctx_check_err(c);
        if (rc) {
            cudaFree(pts);
            return rc;
        }
        return bases_finish(c, pts, count, out);
    });
}

int fb_bases_precompute(fb_ctx* c, fb_bases* b, int window_bits) {
    return ctx_call(c, [&]() -> int {
        if (!b) return ctx_fail(c, FB_ERR_ARG, "null base set");
        int cc = window_bits > 0 ? window_bits : msm_pick_window(b->b.n, true);
        if (cc < 4 || cc > 24) return ctx_fail(c, FB_ERR_ARG, "window bits out of range");
        int nwin = msm_num_windows(cc);
        void* tab;
        FB_CUDA(cudaMalloc(&tab, (size_t)nwin * b->b.n * kTableStride));
        msm_build_table(b->raw, b->b.n, 1, cc, nwin, tab, c->pt_a, c->stream);
        FB_CUDA(cudaStreamSynchronize(c->stream));
        if (b->b.pts != b->raw) cudaFree(b->b.pts);
        b->b.pts = tab;
        b->b.c = cc;
        b->b.nwin = nwin;
        return FB_OK;
    });
}

void fb_bases_free(fb_ctx* c, fb_bases* b) {
    if (!b) return;
    if (c) cudaSetDevice(c->device);
    if (b->b.pts && b->b.pts != b->raw) cudaFree(b->b.pts);
    if (b->raw) cudaFree(b->raw);
    delete b;
}

int fb_bases_get(fb_ctx* c, const fb_bases* b, uint64_t index, uint8_t out48[48]) {
    return ctx_call(c, [&]() -> int {
        if (!b || index >= b->b.n) return ctx_fail(c, FB_ERR_ARG, "base index out of range");
        uint8_t* d_ob = c->out_bytes.as<uint8_t>(48);
        g1_compress_batch(b->raw + index, d_ob, 1, c->stream);
        FB_CUDA(cudaMemcpyAsync(out48, d_ob, 48, cudaMemcpyDeviceToHost, c->stream));
        FB_CUDA(cudaStreamSynchronize(c->stream));
        return FB_OK;
    });
}

static int msm_host(fb_ctx* c, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t* out, int fmt) {
    if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
    ctx_reset_err(c);
    Fr* d_s = ctx_upload_fr(c, c->in_bytes, c->fr_a, scalars_be, count);
    G1Xyzz* d_r = c->small.as<G1Xyzz>(2);
    msm_run(b->b, d_s, count, 1, true, d_r, c->msm, c->stream);
    uint8_t* d_ob = c->out_bytes.as<uint8_t>(96);
    G1Affine* aff = c->pt_b.as<G1Affine>(2);
    g1_xyzz_to_affine_batch(d_r, aff, 1, c->stream);
    if (fmt == 48) g1_compress_batch(aff, d_ob, 1, c->stream);
    else g1_serialize_batch(aff, d_ob, 1, c->stream);
    FB_CUDA(cudaMemcpyAsync(out, d_ob, fmt, cudaMemcpyDeviceToHost, c->stream));
    return ctx_check_err(c);
}

int fb_msm(fb_ctx* c, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out48[48]) {
    return ctx_call(c, [&]() -> int { return msm_host(c, b, scalars_be, count, out48, 48); });
}
int fb_msm_partial(fb_ctx* c, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out96[96]) {
    return ctx_call(c, [&]() -> int { return msm_host(c, b, scalars_be, count, out96, 96); });
}

int fb_g1_sum(fb_ctx* c, const uint8_t* points, uint64_t count, int format, uint8_t out48[48]) {
    return ctx_call(c, [&]() -> int {
        if (format != 48 && format != 96) return ctx_fail(c, FB_ERR_ARG, "format must be 48 or 96");
        ctx_reset_err(c);
        uint8_t* d_b = c->in_bytes.as<uint8_t>(count * format + 16);
        G1Affine* pts = c->pt_a.as<G1Affine>(count + 1);
        if (count) {
            FB_CUDA(cudaMemcpyAsync(d_b, points, count * format, cudaMemcpyHostToDevice, c->stream));
            if (format == 48) g1_decompress_batch(d_b, pts, count, c->d_err + 1, c->stream);
            else g1_deserialize_batch(d_b, pts, count, c->d_err + 1, c->stream);
        }
        G1Xyzz* d_r = c->small.as<G1Xyzz>(2);
        g1_sum_affine(pts, count, d_r, c->scan, c->stream);
        uint8_t* d_ob = c->out_bytes.as<uint8_t>(48);
        ctx_points_out(c, d_r, 1, d_ob);
        FB_CUDA(cudaMemcpyAsync(out48, d_ob, 48, cudaMemcpyDeviceToHost, c->stream));
        return ctx_check_err(c);
    });
}

// ------------------------------------------------------- device-resident calls
int fb_dev_fr_from_be(fb_ctx* c, const void* d_in, void* d_out, uint64_t count) {
    return ctx_call(c, [&]() -> int {
        fr_from_be((const uint8_t*)d_in, (Fr*)d_out, count, c->d_err, c->stream);
        return FB_OK;
    });
}
int fb_dev_fr_to_be(fb_ctx* c, const void* d_in, void* d_out, uint64_t count) {
    return ctx_call(c, [&]() -> int {
        fr_to_be((const Fr*)d_in, (uint8_t*)d_out, count, c->stream);
        return FB_OK;
    });
}
int fb_dev_ntt(fb_ctx* c, int inverse, const void* d_in, void* d_out, uint64_t len) {
    return ctx_call(c, [&]() -> int {
        int lg = ilog2_exact(len);
        if (lg < 0 || lg > 30) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        const NttTables& tab = ctx_ntt_tables(c, lg);
        Fr* d_tmp = c->fr_c.as<Fr>(len);
        ntt_run(tab, c->tile_tw, (const Fr*)d_in, (Fr*)d_out, d_tmp, inverse != 0, c->stream);
        return FB_OK;
    });
}
int fb_dev_msm(fb_ctx* c, const fb_bases* b, const void* d_scalars, uint64_t count, void* d_out48) {
    return ctx_call(c, [&]() -> int {
        if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
        G1Xyzz* d_r = c->small.as<G1Xyzz>(2);
        msm_run(b->b, (const Fr*)d_scalars, count, 1, true, d_r, c->msm, c->stream);
        ctx_points_out(c, d_r, 1, (uint8_t*)d_out48);
        return FB_OK;
    });
}
int fb_dev_msm_partial(fb_ctx* c, const fb_bases* b, const void* d_scalars, uint64_t count, void* d_out96) {
    return ctx_call(c, [&]() -> int {
        if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
        G1Xyzz* d_r = c->small.as<G1Xyzz>(2);
        msm_run(b->b, (const Fr*)d_scalars, count, 1, true, d_r, c->msm, c->stream);
        G1Affine* aff = c->pt_b.as<G1Affine>(2);
        prof_mark("out.affine_compress", c->stream);
        g1_xyzz_to_affine_batch(d_r, aff, 1, c->stream);
        g1_serialize_batch(aff, (uint8_t*)d_out96, 1, c->stream);
        return FB_OK;
    });
}

}  // extern "C"
