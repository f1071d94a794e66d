// ctx.cuh - the context object behind the C ABI (include/fourier_b200.h).
#pragma once
#include <memory>

#include "../../include/fourier_b200.h"
#include "common.cuh"

struct fb_bases {
    fb::MsmBases b;        // pts owned
    fb::G1Affine* raw = nullptr;  // untabulated copy (== b.pts when no table)
};

struct fb_ctx {
    int device = 0;
    uint32_t n = 0, m = 0, t = 0;
    size_t T = 1, M = 1;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    std::string err;
    int window_pref = 0;

    fb::TileTwiddles tile_tw;
    bool tile_tw_ready = false;
    std::map<int, fb::NttTables> ntt;

    // SRS (device, affine Montgomery)
    bool srs_ready = false;
    uint64_t row0 = 0, nrows = 0;
    uint64_t col0 = 0, ncols = 0;      // resident column range of every row (ncols == T: whole rows)
    uint64_t want_col0 = 0, want_ncols = 0;  // requested through fb_ctx_set_col_shard (0,0 = all)
    fb::G1Affine g;                    // host copy of the generator element of the setup
    fb::G1Affine* d_g_tau_x = nullptr; // [T]
    fb::G1Affine* d_g_tau_y = nullptr; // [M]
    fb::G1Affine* d_u = nullptr;       // [nrows][T] untabulated
    fb::MsmBases u_bases;              // [nrows][nwin][T] (or aliases d_u when no table)
    fb::MsmBases y_bases;              // g_tau_y with its table
    std::vector<uint8_t> g2_bytes[3];  // g2, g2^tauX, g2^tauY as read / generated (host only)
    int g2_compressed = -1;

    // scratch
    fb::MsmScratch msm;
    fb::DevBuf in_bytes, fr_a, fr_b, fr_c, pt_a, pt_b, out_bytes, scan, small, eng_a, eng_b;
    fb::Prof prof;
    int* d_err = nullptr;
    uint8_t* h_pinned = nullptr;  // small pinned result buffer
    size_t h_pinned_cap = 0;
};

namespace fb {

extern std::atomic<unsigned long long> g_launches;

struct CtxGuard {
    fb_ctx* c;
    std::unique_lock<std::mutex> lk;
    explicit CtxGuard(fb_ctx* ctx) : c(ctx), lk(ctx->mu) {
        cudaSetDevice(ctx->device);
        g_prof = &ctx->prof;
    }
    ~CtxGuard() {
        prof_mark("idle", c->stream);
        g_prof = nullptr;
    }
};

const NttTables& ctx_ntt_tables(fb_ctx* c, int log_n);
int ctx_fail(fb_ctx* c, int code, const std::string& msg);

// run f(), mapping exceptions to status codes
template <class F>
int ctx_call(fb_ctx* c, F&& f) {
    if (!c) return FB_ERR_ARG;
    CtxGuard g(c);
    try {
        return f();
    } catch (const CudaError& e) {
        cudaGetLastError();
        return ctx_fail(c, FB_ERR_CUDA, e.msg);
    } catch (const std::exception& e) {
        return ctx_fail(c, FB_ERR_CUDA, e.what());
    }
}

inline int ilog2_exact(uint64_t v) {  // -1 unless v is a power of two
    if (v == 0 || (v & (v - 1))) return -1;
    int l = 0;
    while ((1ull << l) < v) l++;
    return l;
}

}  // namespace fb
