// api.cu - the C ABI (include/fourier_b200.h): context and lanes, field-side calls, raw
// kernels and device-resident entry points.  Engine calls (worker_* / master_* /
// SRS) are in engine.cu.
#include <string.h>

#include "ctx.cuh"

namespace fb {

std::atomic<unsigned long long> g_launches{0};
thread_local Prof* g_prof = nullptr;
// the message of the calling thread's last failing call (several threads share one context: a per-context
// string could be overwritten, or its buffer reallocated, between the failing call's return and the read)
thread_local std::string g_last_error;

int ctx_fail(fb_ctx*, int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

static fb_lane* lane_create(int index, bool own_stream) {
    fb_lane* L = new fb_lane();
    L->index = index;
    if (own_stream) {
        FB_CUDA(cudaStreamCreateWithFlags(&L->own_stream, cudaStreamNonBlocking));
        L->stream = L->own_stream;
    }
    for (int k = 0; k < kMaxSplit - 1; k++) {
        FB_CUDA(cudaStreamCreateWithFlags(&L->sub_stream[k], cudaStreamNonBlocking));
        FB_CUDA(cudaEventCreateWithFlags(&L->sub_join[k], cudaEventDisableTiming));
    }
    FB_CUDA(cudaEventCreateWithFlags(&L->sub_fork, cudaEventDisableTiming));
    FB_CUDA(cudaEventCreateWithFlags(&L->ev_fork, cudaEventDisableTiming));
    FB_CUDA(cudaEventCreateWithFlags(&L->ev_done, cudaEventDisableTiming));
    FB_CUDA(cudaMalloc(&L->d_err, 2 * sizeof(int)));
    FB_CUDA(cudaMemset(L->d_err, 0, 2 * sizeof(int)));
    L->h_pinned_cap = 1 << 16;
    FB_CUDA(cudaMallocHost(&L->h_pinned, L->h_pinned_cap));
    return L;
}

static void lane_destroy(fb_lane* L) {
    if (!L) return;
    DevBuf* bufs[] = {&L->msm.counts, &L->msm.offsets, &L->msm.cursor, &L->msm.sorted, &L->msm.buckets, &L->msm.block_sums,
                      &L->msm.seg_a, &L->msm.seg_t, &L->msm.order, &L->msm.misc, &L->msm.pts_a, &L->msm.pts_b, &L->msm.pref,
                      &L->msm.where, &L->msm.stage, &L->msm.inv, &L->msm.stats, &L->in_bytes, &L->fr_a, &L->fr_b, &L->fr_c,
                      &L->pt_a, &L->pt_b, &L->out_bytes, &L->scan, &L->small, &L->eng_a, &L->eng_b};
    for (DevBuf* b : bufs) b->release();
    for (int k = 0; k < kMaxSplit - 1; k++) {
        MsmScratch& s = L->msm_sub[k];
        DevBuf* sb[] = {&s.counts, &s.offsets, &s.cursor, &s.sorted, &s.buckets, &s.block_sums, &s.seg_a, &s.seg_t, &s.order,
                        &s.misc, &s.pts_a, &s.pts_b, &s.pref, &s.where, &s.stage, &s.inv, &s.stats};
        for (DevBuf* b : sb) b->release();
        if (L->sub_stream[k]) cudaStreamDestroy(L->sub_stream[k]);
        if (L->sub_join[k]) cudaEventDestroy(L->sub_join[k]);
    }
    L->msm_parts.release();
    if (L->sub_fork) cudaEventDestroy(L->sub_fork);
    for (auto& kv : L->prof.marks) cudaEventDestroy(kv.second);
    for (auto e : L->prof.pool) cudaEventDestroy(e);
    if (L->d_err) cudaFree(L->d_err);
    if (L->h_pinned) cudaFreeHost(L->h_pinned);
    for (int k = 0; k < 2; k++) {
        if (L->h_stage[k]) cudaFreeHost(L->h_stage[k]);
        if (L->ev_stage[k]) cudaEventDestroy(L->ev_stage[k]);
    }
    if (L->ev_fork) cudaEventDestroy(L->ev_fork);
    if (L->ev_done) cudaEventDestroy(L->ev_done);
    if (L->own_stream) cudaStreamDestroy(L->own_stream);
    delete L;
}

// the caller's stream waits for every asynchronous call in flight on the lanes (c->mu held)
static void ctx_join_locked(fb_ctx* c) {
    for (int k = 0; k < kMaxLanes; k++) {
        fb_lane* L = c->lanes[k];
        if (L && L->dirty) {
            cudaStreamWaitEvent(c->caller_stream(), L->ev_done, 0);
            L->dirty = false;
        }
    }
}

fb_lane* ctx_acquire(fb_ctx* c, fb_call_kind kind) {
    std::unique_lock<std::mutex> lk(c->mu);
    fb_lane* L = nullptr;
    if (kind == FB_CALL_EXCL) {
        c->cv.wait(lk, [&] { return !c->excl && c->active == 0; });
        c->excl = true;
        c->active = 1;
        L = c->lanes[0];
        L->busy = true;
        lk.unlock();
        cudaDeviceSynchronize();  // nothing asynchronous may still read the tables a setup call replaces
        return L;
    }
    auto pick = [&]() -> fb_lane* {
        if (c->excl) return nullptr;
        if (kind == FB_CALL_UTIL || (kind == FB_CALL_ASYNC && c->async_lanes <= 1)) return c->ulane->busy ? nullptr : c->ulane;
        if (kind == FB_CALL_ASYNC) {
            for (int k = 0; k < c->async_lanes; k++) {
                fb_lane* cand = c->lanes[(c->rr + k) % c->async_lanes];
                if (!cand->busy) {
                    c->rr = (c->rr + k + 1) % c->async_lanes;
                    return cand;
                }
            }
            return nullptr;
        }
        for (int k = 0; k < kMaxLanes; k++)
            if (!c->lanes[k]->busy) return c->lanes[k];
        return nullptr;
    };
    c->cv.wait(lk, [&] { return (L = pick()) != nullptr; });
    L->busy = true;
    c->active++;
    if (L == c->ulane) {
        L->stream = c->caller_stream();
        ctx_join_locked(c);  // stream order against everything enqueued before this call
    } else if (kind == FB_CALL_ASYNC) {
        // fork: the lane starts after everything the caller has enqueued on S so far (its inputs)
        cudaEventRecord(L->ev_fork, c->caller_stream());
        cudaStreamWaitEvent(L->stream, L->ev_fork, 0);
    }
    return L;
}

void ctx_release(fb_ctx* c, fb_lane* L, fb_call_kind kind) {
    std::lock_guard<std::mutex> lk(c->mu);
    if (kind == FB_CALL_ASYNC && L != c->ulane) {
        cudaEventRecord(L->ev_done, L->stream);
        L->dirty = true;
    }
    L->busy = false;
    c->active--;
    if (kind == FB_CALL_EXCL) c->excl = false;
    c->cv.notify_all();
}

void lane_h2d(fb_ctx* c, fb_lane* L, void* d_dst, const void* h_src, size_t bytes) {
    if (!bytes) return;
    bool direct = bytes < (256u << 10) || c->staging == 1;
    if (!direct) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, h_src) == cudaSuccess) direct = at.type != cudaMemoryTypeUnregistered;
        else cudaGetLastError();
    }
    if (direct) {
        FB_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, L->stream));
        return;
    }
    // pageable source: copy chunk k+1 into pinned memory while chunk k is on the wire
    for (int k = 0; k < 2; k++)
        if (!L->h_stage[k]) {
            FB_CUDA(cudaMallocHost(&L->h_stage[k], kStageChunk));
            FB_CUDA(cudaEventCreateWithFlags(&L->ev_stage[k], cudaEventDisableTiming));
        }
    size_t off = 0;
    for (int k = 0; off < bytes; k ^= 1) {
        const size_t cnt = bytes - off < kStageChunk ? bytes - off : kStageChunk;
        FB_CUDA(cudaEventSynchronize(L->ev_stage[k]));  // the previous copy out of this buffer is done
        memcpy(L->h_stage[k], (const uint8_t*)h_src + off, cnt);
        FB_CUDA(cudaMemcpyAsync((uint8_t*)d_dst + off, L->h_stage[k], cnt, cudaMemcpyHostToDevice, L->stream));
        FB_CUDA(cudaEventRecord(L->ev_stage[k], L->stream));
        off += cnt;
    }
}

void lane_msm(fb_ctx* c, fb_lane* L, const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, G1Xyzz* d_out,
              const Fr* d_scalars2) {
    const int S = (g_prof && g_prof->on) ? 1 : c->msm_split;  // per-phase timing needs one stream
    if (S <= 1) {
        L->msm.overlapped = !(g_prof && g_prof->on) && c->async_lanes > 1 && L != c->ulane;  // heavy calls rotate over lanes
        msm_run(bases, d_scalars, len, batch, true, d_out, L->msm, L->stream, d_scalars2);
        return;
    }
    MsmExec ex;
    ex.S = S;
    ex.scratch[0] = &L->msm;
    ex.stream[0] = L->stream;
    for (int k = 1; k < kMaxSplit; k++) {
        ex.scratch[k] = &L->msm_sub[k - 1];
        ex.stream[k] = L->sub_stream[k - 1];
        ex.join[k] = L->sub_join[k - 1];
    }
    ex.fork = L->sub_fork;
    ex.parts = &L->msm_parts;
    msm_run_split(bases, d_scalars, len, batch, true, d_out, ex, d_scalars2);
}

bool lane_pair_fork(fb_ctx* c, fb_lane* L, const MsmBases& bases, const Fr* d_a, size_t cnt, G1Xyzz* d_r) {
    const bool profiling = g_prof && g_prof->on;  // per-phase timing needs one stream
    bool two = c->pair_mode == 2 || (c->pair_mode == 0 && cnt >= ((size_t)1 << 18));
    if (profiling || c->msm_split > 1 || cnt == 0) two = false;
    if (!two) return false;
    FB_CUDA(cudaEventRecord(L->sub_fork, L->stream));
    FB_CUDA(cudaStreamWaitEvent(L->sub_stream[0], L->sub_fork, 0));
    L->msm_sub[0].overlapped = true;
    msm_run(bases, d_a, cnt, 1, true, d_r, L->msm_sub[0], L->sub_stream[0]);
    FB_CUDA(cudaEventRecord(L->sub_join[0], L->sub_stream[0]));
    return true;
}

void lane_pair_join(fb_ctx* c, fb_lane* L, bool forked, const MsmBases& bases, const Fr* d_a, const Fr* d_b, size_t cnt, G1Xyzz* d_r) {
    if (!forked) {
        lane_msm(c, L, bases, d_a, cnt, 2, d_r, d_b);
        return;
    }
    L->msm.overlapped = true;
    msm_run(bases, d_b, cnt, 1, true, d_r + 1, L->msm, L->stream);
    L->msm.overlapped = false;
    FB_CUDA(cudaStreamWaitEvent(L->stream, L->sub_join[0], 0));
}

const NttTables& ctx_ntt_tables(fb_ctx* c, fb_lane* L, int log_n) {
    std::lock_guard<std::mutex> lk(c->tab_mu);
    bool built = false;
    if (!c->tile_tw_ready) {
        ntt_build_tile_twiddles(c->tile_tw, L->stream);
        c->tile_tw_ready = true;
        built = true;
    }
    auto it = c->ntt.find(log_n);
    if (it == c->ntt.end()) {
        NttTables t;
        ntt_build_tables(t, log_n, L->stream);
        it = c->ntt.emplace(log_n, t).first;
        built = true;
    }
    if (built) FB_CUDA(cudaStreamSynchronize(L->stream));  // other lanes may use the tables as soon as the lock drops
    return it->second;
}

// host bytes -> device Fr (Montgomery); returns FB_ERR_SCALAR lazily via d_err
Fr* ctx_upload_fr(fb_ctx* c, fb_lane* L, DevBuf& bytes, DevBuf& dst, const uint8_t* h_be, size_t count) {
    uint8_t* d_b = bytes.as<uint8_t>(count * 32 + 32);
    Fr* d = dst.as<Fr>(count + 1);
    if (count) {
        prof_mark("h2d", L->stream);
        lane_h2d(c, L, d_b, h_be, count * 32);
        fr_from_be(d_b, d, count, L->d_err, L->stream);
    }
    return d;
}

void ctx_reset_err(fb_ctx*, fb_lane* L) { FB_CUDA(cudaMemsetAsync(L->d_err, 0, 2 * sizeof(int), L->stream)); }

// read the two error flags (scalar, point) after a sync
int ctx_check_err(fb_ctx* c, fb_lane* L) {
    int* h = reinterpret_cast<int*>(L->h_pinned + L->h_pinned_cap - 16);
    FB_CUDA(cudaMemcpyAsync(h, L->d_err, 2 * sizeof(int), cudaMemcpyDeviceToHost, L->stream));
    FB_CUDA(cudaStreamSynchronize(L->stream));
    if (h[0]) return ctx_fail(c, FB_ERR_SCALAR, "Invalid scalar");
    if (h[1]) return ctx_fail(c, FB_ERR_POINT, "Failed to uncompress G1 point (code " + std::to_string(h[1]) + ")");
    return FB_OK;
}

// XYZZ results -> 48-byte compressed strings on the device
void ctx_points_out(fb_ctx*, fb_lane* L, const G1Xyzz* d_xyzz, size_t count, uint8_t* d_out48) {
    G1Affine* aff = L->pt_b.as<G1Affine>(count + 1);
    prof_mark("out.affine_compress", L->stream);
    g1_xyzz_to_affine_batch(d_xyzz, aff, count, L->stream);
    g1_compress_batch(aff, d_out48, count, L->stream);
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
        for (int k = 0; k < kMaxLanes; k++) c->lanes[k] = lane_create(k, true);
        c->ulane = lane_create(-1, false);
        c->ulane->stream = c->own_stream;
    } catch (const CudaError&) {
        fb_ctx_destroy(c);
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
    if (c->u_bases.xs) cudaFree((void*)c->u_bases.xs);
    if (c->y_bases.pts && c->y_bases.pts != c->d_g_tau_y) cudaFree(c->y_bases.pts);
    if (c->y_bases.xs) cudaFree((void*)c->y_bases.xs);
    if (c->d_u) cudaFree(c->d_u);
    c->shard_buf.release();
    c->shard_coeffs.release();
    for (int k = 0; k < kMaxLanes; k++) lane_destroy(c->lanes[k]);
    lane_destroy(c->ulane);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

// the calling thread's last failure (any context): valid until that thread's next failing call
const char* fb_last_error(const fb_ctx*) { return g_last_error.c_str(); }

static void ctx_sync_all(fb_ctx* c) {
    for (int k = 0; k < kMaxLanes; k++) FB_CUDA(cudaStreamSynchronize(c->lanes[k]->stream));
    FB_CUDA(cudaStreamSynchronize(c->caller_stream()));
}

int fb_ctx_set_stream(fb_ctx* c, void* s) {
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane*) -> int {
        c->user_stream = (cudaStream_t)s;
        c->ulane->stream = c->caller_stream();
        return FB_OK;
    });
}

int fb_ctx_sync(fb_ctx* c) {
    if (!c) return FB_ERR_ARG;
    cudaSetDevice(c->device);
    try {
        {
            std::lock_guard<std::mutex> lk(c->mu);
            ctx_join_locked(c);
        }
        ctx_sync_all(c);
        // The device-resident decoders (fb_dev_fr_from_be, fb_dev_combine_commit_open, ...) are asynchronous: a malformed
        // scalar or point sets a sticky flag of the caller-stream lane, reported - and cleared - here.
        fb_lane* U = c->ulane;
        int flags[2] = {0, 0};
        FB_CUDA(cudaMemcpy(flags, U->d_err, sizeof(flags), cudaMemcpyDeviceToHost));
        if (flags[0] || flags[1]) {
            FB_CUDA(cudaMemset(U->d_err, 0, sizeof(flags)));
            if (flags[0]) return ctx_fail(c, FB_ERR_SCALAR, "Invalid scalar (in an earlier device-resident call)");
            return ctx_fail(c, FB_ERR_POINT,
                            "Failed to uncompress G1 point (code " + std::to_string(flags[1]) + ", in an earlier device-resident call)");
        }
    } catch (const CudaError& e) {
        cudaGetLastError();
        return ctx_fail(c, FB_ERR_CUDA, e.msg);
    }
    return FB_OK;
}

int fb_ctx_join(fb_ctx* c) {
    if (!c) return FB_ERR_ARG;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    ctx_join_locked(c);
    return FB_OK;
}

int fb_ctx_set_lanes(fb_ctx* c, int lanes) {
    if (!c || lanes < 1 || lanes > kMaxLanes) return FB_ERR_ARG;
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane*) -> int {
        c->async_lanes = lanes;
        return FB_OK;
    });
}

int fb_ctx_set_window(fb_ctx* c, int w) {
    if (!c || w < -1 || (w > 0 && w < 4) || w > 24) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->window_pref = w;
    return FB_OK;
}

int fb_ctx_get_window(fb_ctx* c, int* window_bits, int* num_windows) {
    if (!c) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->srs_ready) return ctx_fail(c, FB_ERR_STATE, "setup not loaded");
    if (window_bits) *window_bits = c->u_bases.c;
    if (num_windows) *num_windows = c->u_bases.c > 0 ? c->u_bases.nwin : 0;
    return FB_OK;
}

int fb_ctx_set_open_mode(fb_ctx* c, int mode) {
    if (!c || mode < 0 || mode > 1) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->open_mode = mode;
    return FB_OK;
}

int fb_ctx_set_msm_split(fb_ctx* c, int parts) {
    if (!c || parts < 1 || parts > kMaxSplit) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->msm_split = parts;
    return FB_OK;
}

int fb_ctx_set_pair_mode(fb_ctx* c, int mode) {
    if (!c || mode < 0 || mode > 2) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->pair_mode = mode;
    return FB_OK;
}

int fb_ctx_set_staging(fb_ctx* c, int mode) {
    if (!c || mode < 0 || mode > 1) return FB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->staging = mode;
    return FB_OK;
}

uint64_t fb_launch_count(const fb_ctx*) { return g_launches.load(); }

static fb_lane* all_lane(fb_ctx* c, int k) { return k < kMaxLanes ? c->lanes[k] : c->ulane; }

int fb_prof_enable(fb_ctx* c, int on) {
    if (!c) return FB_ERR_ARG;
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane*) -> int {
        for (int k = 0; k <= kMaxLanes; k++) all_lane(c, k)->prof.on = on != 0;
        return FB_OK;
    });
}

// JSON object {"phase": [total_ms, count], ...} of everything marked since the last read (all lanes; a phase's
// time is measured between consecutive marks on ITS lane's stream, so with lanes overlapping the sum of the
// phases exceeds the elapsed time)
int fb_prof_read(fb_ctx* c, char* buf, size_t cap) {
    if (!c || !buf || cap < 3) return FB_ERR_ARG;
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane*) -> int {
        std::map<std::string, std::pair<double, int>> acc;
        for (int k = 0; k <= kMaxLanes; k++) {
            auto& pr = all_lane(c, k)->prof;
            auto& m = pr.marks;
            for (size_t i = 0; i + 1 < m.size(); i++) {
                float ms = 0;
                if (cudaEventElapsedTime(&ms, m[i].second, m[i + 1].second) != cudaSuccess) ms = 0;
                auto& a = acc[m[i].first];
                a.first += ms;
                a.second += 1;
            }
            for (auto& kv : m) pr.pool.push_back(kv.second);
            m.clear();
        }
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
        if (out.size() + 1 > cap) return ctx_fail(c, FB_ERR_ARG, "profile buffer too small");
        memcpy(buf, out.c_str(), out.size() + 1);
        return FB_OK;
    });
}

// work counters of the MSMs run while profiling was on, summed over the lanes and reset:
// out[0] MSMs, [1] sorted entries, [2] affine-round output slots, [3] survivors through the XYZZ kernel, [4] buckets
int fb_msm_stats(fb_ctx* c, uint64_t out[8]) {
    if (!c || !out) return FB_ERR_ARG;
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane*) -> int {
        for (int i = 0; i < 8; i++) out[i] = 0;
        for (int k = 0; k <= kMaxLanes; k++) {
            fb_lane* L = all_lane(c, k);
            if (!L->msm.stats_zeroed) continue;  // (profiling runs unsplit: the sub-MSM scratch never counts)
            unsigned long long h[8];
            FB_CUDA(cudaMemcpy(h, L->msm.stats.p, sizeof h, cudaMemcpyDeviceToHost));
            FB_CUDA(cudaMemset(L->msm.stats.p, 0, sizeof h));
            for (int i = 0; i < 8; i++) out[i] += h[i];
        }
        return FB_OK;
    });
}

// ------------------------------------------------------------ field-side calls
static int do_ntt_host(fb_ctx* c, fb_lane* L, int log_n, bool inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    ctx_reset_err(c, L);
    const NttTables& tab = ctx_ntt_tables(c, L, log_n);
    Fr* d_in = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, in_be, len);
    Fr* d_out = L->fr_b.as<Fr>(len);
    Fr* d_tmp = L->fr_c.as<Fr>(len);
    ntt_run(tab, c->tile_tw, d_in, d_out, d_tmp, inverse, L->stream);
    uint8_t* d_ob = L->out_bytes.as<uint8_t>(len * 32);
    fr_to_be(d_out, d_ob, len, L->stream);
    FB_CUDA(cudaMemcpyAsync(out_be, d_ob, len * 32, cudaMemcpyDeviceToHost, L->stream));
    return ctx_check_err(c, L);
}

int fb_ntt(fb_ctx* c, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        int lg = ilog2_exact(len);
        if (lg < 0 || lg > 30) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        return do_ntt_host(c, L, lg, inverse != 0, in_be, len, out_be);
    });
}

int fb_fft(fb_ctx* c, int left, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        size_t width = left ? c->T : c->M;
        if (len > width) return ctx_fail(c, FB_ERR_ARG, "Supplied list is longer than the available max width");
        int lg = ilog2_exact(len);
        if (lg < 0) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        return do_ntt_host(c, L, lg, inverse != 0, in_be, len, out_be);
    });
}

int fb_evaluate(fb_ctx* c, const uint8_t* poly_be, uint64_t len, const uint8_t x_be[32], uint8_t out_y_be[32]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        ctx_reset_err(c, L);
        Fr* d_p = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, poly_be, len);
        Fr* d_x = ctx_upload_fr(c, L, L->small, L->fr_c, x_be, 1);
        Fr* d_y = d_x + 1;  // fr_c holds count+1 elements
        horner_quotient(d_p, len, d_x, d_y, nullptr, L->scan, L->stream);
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(32);
        fr_to_be(d_y, d_ob, 1, L->stream);
        FB_CUDA(cudaMemcpyAsync(out_y_be, d_ob, 32, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

// ----------------------------------------------------------------- raw kernels
static int bases_finish(fb_ctx*, G1Affine* pts, uint64_t count, fb_bases** out) {
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
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!out || (format != 48 && format != 96) || count == 0) return ctx_fail(c, FB_ERR_ARG, "bad base-set arguments");
        ctx_reset_err(c, L);
        uint8_t* d_b = L->in_bytes.as<uint8_t>(count * format);
        lane_h2d(c, L, d_b, points, count * format);
        G1Affine* pts;
        FB_CUDA(cudaMalloc(&pts, count * sizeof(G1Affine)));
        if (format == 48) g1_decompress_batch(d_b, pts, count, L->d_err + 1, L->stream);
        else g1_deserialize_batch(d_b, pts, count, L->d_err + 1, L->stream);
        int rc = ctx_check_err(c, L);
        if (rc) {
            cudaFree(pts);
            return rc;
        }
        return bases_finish(c, pts, count, out);
    });
}

int fb_bases_from_scalars(fb_ctx* c, const uint8_t* scalars_be, uint64_t count, fb_bases** out) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!out || count == 0) return ctx_fail(c, FB_ERR_ARG, "bad base-set arguments");
        ctx_reset_err(c, L);
        Fr* d_s = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, scalars_be, count);
        G1Affine* pts;
        FB_CUDA(cudaMalloc(&pts, count * sizeof(G1Affine)));
        g1_fixed_base_mul(d_s, count, pts, L->pt_a, L->stream);
        int rc = ctx_check_err(c, L);
        if (rc) {
            cudaFree(pts);
            return rc;
        }
        return bases_finish(c, pts, count, out);
    });
}

int fb_bases_precompute(fb_ctx* c, fb_bases* b, int window_bits) {
    return ctx_call(c, FB_CALL_EXCL, [&](fb_lane* L) -> int {
        if (!b) return ctx_fail(c, FB_ERR_ARG, "null base set");
        int cc = window_bits > 0 ? window_bits : msm_pick_window(b->b.n, true);
        if (cc < 4 || cc > 24) return ctx_fail(c, FB_ERR_ARG, "window bits out of range");
        int nwin = msm_num_windows(cc);
        void* tab;
        FB_CUDA(cudaMalloc(&tab, (size_t)nwin * b->b.n * kTableStride));
        msm_build_table(b->raw, b->b.n, 1, cc, nwin, tab, L->pt_a, L->stream);
        FB_CUDA(cudaStreamSynchronize(L->stream));
        if (b->b.pts != b->raw) cudaFree(b->b.pts);
        if (b->b.xs) cudaFree((void*)b->b.xs);
        b->b.xs = msm_alloc_xtable(tab, (size_t)nwin * b->b.n, L->stream);
        FB_CUDA(cudaStreamSynchronize(L->stream));
        b->b.pts = tab;
        b->b.c = cc;
        b->b.nwin = nwin;
        return FB_OK;
    });
}

int fb_bases_get_window(fb_ctx* c, const fb_bases* b, int* window_bits, int* num_windows) {
    if (!c || !b) return FB_ERR_ARG;
    if (window_bits) *window_bits = b->b.c;
    if (num_windows) *num_windows = b->b.c > 0 ? b->b.nwin : 0;
    return FB_OK;
}

void fb_bases_free(fb_ctx* c, fb_bases* b) {
    if (!b) return;
    if (c) cudaSetDevice(c->device);
    if (b->b.pts && b->b.pts != b->raw) cudaFree(b->b.pts);
    if (b->b.xs) cudaFree((void*)b->b.xs);
    if (b->raw) cudaFree(b->raw);
    delete b;
}

int fb_bases_get(fb_ctx* c, const fb_bases* b, uint64_t index, uint8_t out48[48]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (!b || index >= b->b.n) return ctx_fail(c, FB_ERR_ARG, "base index out of range");
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(48);
        g1_compress_batch(b->raw + index, d_ob, 1, L->stream);
        FB_CUDA(cudaMemcpyAsync(out48, d_ob, 48, cudaMemcpyDeviceToHost, L->stream));
        FB_CUDA(cudaStreamSynchronize(L->stream));
        return FB_OK;
    });
}

static int msm_host(fb_ctx* c, fb_lane* L, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t* out, int fmt) {
    if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
    ctx_reset_err(c, L);
    Fr* d_s = ctx_upload_fr(c, L, L->in_bytes, L->fr_a, scalars_be, count);
    G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
    lane_msm(c, L, b->b, d_s, count, 1, d_r);
    uint8_t* d_ob = L->out_bytes.as<uint8_t>(96);
    G1Affine* aff = L->pt_b.as<G1Affine>(2);
    g1_xyzz_to_affine_batch(d_r, aff, 1, L->stream);
    if (fmt == 48) g1_compress_batch(aff, d_ob, 1, L->stream);
    else g1_serialize_batch(aff, d_ob, 1, L->stream);
    FB_CUDA(cudaMemcpyAsync(out, d_ob, fmt, cudaMemcpyDeviceToHost, L->stream));
    return ctx_check_err(c, L);
}

int fb_msm(fb_ctx* c, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out48[48]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int { return msm_host(c, L, b, scalars_be, count, out48, 48); });
}
int fb_msm_partial(fb_ctx* c, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out96[96]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int { return msm_host(c, L, b, scalars_be, count, out96, 96); });
}

int fb_g1_sum(fb_ctx* c, const uint8_t* points, uint64_t count, int format, uint8_t out48[48]) {
    return ctx_call(c, FB_CALL_HOST, [&](fb_lane* L) -> int {
        if (format != 48 && format != 96) return ctx_fail(c, FB_ERR_ARG, "format must be 48 or 96");
        ctx_reset_err(c, L);
        uint8_t* d_b = L->in_bytes.as<uint8_t>(count * format + 16);
        G1Affine* pts = L->pt_a.as<G1Affine>(count + 1);
        if (count) {
            lane_h2d(c, L, d_b, points, count * format);
            if (format == 48) g1_decompress_batch(d_b, pts, count, L->d_err + 1, L->stream);
            else g1_deserialize_batch(d_b, pts, count, L->d_err + 1, L->stream);
        }
        G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
        g1_sum_affine(pts, count, d_r, L->scan, L->stream);
        uint8_t* d_ob = L->out_bytes.as<uint8_t>(48);
        ctx_points_out(c, L, d_r, 1, d_ob);
        FB_CUDA(cudaMemcpyAsync(out48, d_ob, 48, cudaMemcpyDeviceToHost, L->stream));
        return ctx_check_err(c, L);
    });
}

// ------------------------------------------------------- device-resident calls
int fb_dev_fr_from_be(fb_ctx* c, const void* d_in, void* d_out, uint64_t count) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        fr_from_be((const uint8_t*)d_in, (Fr*)d_out, count, L->d_err, L->stream);
        return FB_OK;
    });
}
int fb_dev_fr_to_be(fb_ctx* c, const void* d_in, void* d_out, uint64_t count) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        fr_to_be((const Fr*)d_in, (uint8_t*)d_out, count, L->stream);
        return FB_OK;
    });
}
int fb_dev_ntt(fb_ctx* c, int inverse, const void* d_in, void* d_out, uint64_t len) {
    return ctx_call(c, FB_CALL_UTIL, [&](fb_lane* L) -> int {
        int lg = ilog2_exact(len);
        if (lg < 0 || lg > 30) return ctx_fail(c, FB_ERR_ARG, "A list with power-of-two length is expected");
        const NttTables& tab = ctx_ntt_tables(c, L, lg);
        Fr* d_tmp = L->fr_c.as<Fr>(len);
        ntt_run(tab, c->tile_tw, (const Fr*)d_in, (Fr*)d_out, d_tmp, inverse != 0, L->stream);
        return FB_OK;
    });
}
int fb_dev_msm(fb_ctx* c, const fb_bases* b, const void* d_scalars, uint64_t count, void* d_out48) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
        G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
        lane_msm(c, L, b->b, (const Fr*)d_scalars, count, 1, d_r);
        ctx_points_out(c, L, d_r, 1, (uint8_t*)d_out48);
        return FB_OK;
    });
}
int fb_dev_msm_partial(fb_ctx* c, const fb_bases* b, const void* d_scalars, uint64_t count, void* d_out96) {
    return ctx_call(c, FB_CALL_ASYNC, [&](fb_lane* L) -> int {
        if (!b || count > b->b.n) return ctx_fail(c, FB_ERR_ARG, "scalar count exceeds the base set");
        G1Xyzz* d_r = L->small.as<G1Xyzz>(2);
        lane_msm(c, L, b->b, (const Fr*)d_scalars, count, 1, d_r);
        G1Affine* aff = L->pt_b.as<G1Affine>(2);
        prof_mark("out.affine_compress", L->stream);
        g1_xyzz_to_affine_batch(d_r, aff, 1, L->stream);
        g1_serialize_batch(aff, (uint8_t*)d_out96, 1, L->stream);
        return FB_OK;
    });
}

}  // extern "C"
