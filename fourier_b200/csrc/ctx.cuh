// ctx.cuh - the context object behind the C ABI (include/fourier_b200.h).
//
// Threading model (SURVEY 8b "Threading"; reference src/rpc.rs:165, 414-416, 502: the backend sits in an
// Arc and every connection's handler calls it from its own runtime thread).  The SRS, its window tables
// and the twiddle tables are read-only after setup.  Everything a call writes - scratch in HBM, the error
// flags, the pinned result / staging buffers, the CUDA stream it launches on - belongs to a LANE:
//
//   lanes[0..kMaxLanes)   own streams.  A host call (operands and results in host memory) takes the
//                         lowest free lane, runs there and synchronises that lane only, so calls from several
//                         host threads overlap on the GPU.  With fb_ctx_set_lanes(k >= 2) the heavy
//                         device-resident calls (commit / open / msm / rows) rotate over lanes[0..k): the lane
//                         forks from the caller's stream, the call returns at once, and fb_ctx_join makes the
//                         caller's stream wait for everything in flight - a commit and an open of one step then
//                         run side by side (one MSM's latency-bound phases under the other's multiplier-bound ones).
//   ulane                 runs on the CALLER's stream S (fb_ctx_set_stream, else a stream of the context): the
//                         light device-resident utilities (codecs, NTT) and, by default (lanes = 1), the heavy
//                         device-resident calls too: plain stream order, no join needed.
//
// Setup calls (generate / load / free of the SRS) are exclusive: they wait for every call in flight, drain the
// device and run alone.
#pragma once
#include <condition_variable>
#include <memory>

#include "../../include/fourier_b200.h"
#include "common.cuh"

struct fb_bases {
    fb::MsmBases b;        // pts owned
    fb::G1Affine* raw = nullptr;  // untabulated copy (== b.pts when no table)
};

struct fb_lane {
    int index = 0;
    cudaStream_t stream = nullptr;   // ulane: the caller's stream
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_done = nullptr;
    bool busy = false;   // a host thread is inside a call on this lane
    bool dirty = false;  // asynchronous work enqueued since the last join
    fb::MsmScratch msm;
    // sub-MSMs of a split MSM (msm_run_split): sub-job 0 uses `msm` on `stream`, the others these
    fb::MsmScratch msm_sub[fb::kMaxSplit - 1];
    cudaStream_t sub_stream[fb::kMaxSplit - 1] = {nullptr, nullptr, nullptr};
    cudaEvent_t sub_fork = nullptr;
    cudaEvent_t sub_join[fb::kMaxSplit - 1] = {nullptr, nullptr, nullptr};
    fb::DevBuf msm_parts;
    fb::DevBuf in_bytes, fr_a, fr_b, fr_c, pt_a, pt_b, out_bytes, scan, small, eng_a, eng_b;
    fb::Prof prof;
    int* d_err = nullptr;
    uint8_t* h_pinned = nullptr;  // small pinned result buffer
    size_t h_pinned_cap = 0;
    // pinned staging ring for uploads from pageable host memory (a Rust Vec<u8>, Python bytes)
    uint8_t* h_stage[2] = {nullptr, nullptr};
    cudaEvent_t ev_stage[2] = {nullptr, nullptr};
};

constexpr int kMaxLanes = 4;
constexpr size_t kStageChunk = (size_t)4 << 20;

enum fb_call_kind { FB_CALL_HOST = 0, FB_CALL_ASYNC = 1, FB_CALL_UTIL = 2, FB_CALL_EXCL = 3 };

struct fb_ctx {
    int device = 0;
    uint32_t n = 0, m = 0, t = 0;
    size_t T = 1, M = 1;
    cudaStream_t own_stream = nullptr;   // S when the caller gave none
    cudaStream_t user_stream = nullptr;  // S as given through fb_ctx_set_stream
    std::mutex mu;                       // lane table, call accounting
    std::condition_variable cv;
    fb_lane* lanes[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
    fb_lane* ulane = nullptr;
    int async_lanes = 1;  // fb_ctx_set_lanes
    unsigned rr = 0;      // rotation of the asynchronous heavy calls
    int active = 0;       // calls in flight
    bool excl = false;
    int window_pref = 0;
    int open_mode = 0;    // 0: evaluation form (default); 1: the reference's literal iNTT -> eval -> divide -> NTT
    int msm_split = 1;    // sub-MSMs per MSM (fb_ctx_set_msm_split)
    int pair_mode = 0;    // commit + open of one row: 0 by size, 1 one batch-of-two pass, 2 two MSMs on two streams (fb_ctx_set_pair_mode)
    int staging = 0;      // 0: pinned ring for pageable host input; 1: hand pageable pointers to cudaMemcpyAsync

    std::mutex tab_mu;  // twiddle tables are built on first use, by whichever lane needs them
    fb::TileTwiddles tile_tw;
    bool tile_tw_ready = false;
    std::map<int, fb::NttTables> ntt;

    // SRS (device, affine Montgomery): read-only between setup calls
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

    // sharded commit + open (fb_*_commit_open_shard_begin / _finish): what lives between the two halves.  One sharded
    // open in flight per context.
    fb::DevBuf shard_buf;     // 1/(w^j - alpha) over the slice, partial sums, constants
    fb::DevBuf shard_coeffs;  // host form only: the slice's coefficients (Montgomery)
    uint64_t shard_len = 0, shard_cnt = 0;

    cudaStream_t caller_stream() const { return user_stream ? user_stream : own_stream; }
};

namespace fb {

extern std::atomic<unsigned long long> g_launches;

const NttTables& ctx_ntt_tables(fb_ctx* c, fb_lane* L, int log_n);
int ctx_fail(fb_ctx* c, int code, const std::string& msg);
fb_lane* ctx_acquire(fb_ctx* c, fb_call_kind kind);
void ctx_release(fb_ctx* c, fb_lane* L, fb_call_kind kind);
// host -> device copy on the lane's stream; pageable sources go through the lane's pinned ring
void lane_h2d(fb_ctx* c, fb_lane* L, void* d_dst, const void* h_src, size_t bytes);
// msm_run on the lane, split into the context's number of sub-MSMs when the job is large enough
void lane_msm(fb_ctx* c, fb_lane* L, const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, G1Xyzz* d_out,
              const Fr* d_scalars2 = nullptr);

// The two MSMs of a commit + open over one base row: d_r[0] = sum a_k P_k, d_r[1] = sum b_k P_k.  Rows of >= 2^18 points run
// as two plain MSMs on two streams of the lane - lane_pair_fork starts the commit MSM on a sub-stream BEFORE the caller
// enqueues the field half that produces b, so that it runs under the field half and its latency-bound phases under the
// other MSM's multiplier-bound ones - smaller rows as one batch-of-two pass (one sort, one reduction).  lane_pair_join runs
// the second MSM (or the batch of two) on the lane's stream and joins.
bool lane_pair_fork(fb_ctx* c, fb_lane* L, const MsmBases& bases, const Fr* d_a, size_t cnt, G1Xyzz* d_r);
void lane_pair_join(fb_ctx* c, fb_lane* L, bool forked, const MsmBases& bases, const Fr* d_a, const Fr* d_b, size_t cnt, G1Xyzz* d_r);

// run f(lane), mapping exceptions to status codes
template <class F>
int ctx_call(fb_ctx* c, fb_call_kind kind, F&& f) {
    if (!c) return FB_ERR_ARG;
    cudaSetDevice(c->device);
    fb_lane* L = ctx_acquire(c, kind);
    g_prof = &L->prof;
    int rc;
    try {
        rc = f(L);
    } catch (const CudaError& e) {
        cudaGetLastError();
        rc = ctx_fail(c, FB_ERR_CUDA, e.msg);
    } catch (const std::exception& e) {
        rc = ctx_fail(c, FB_ERR_CUDA, e.what());
    }
    prof_mark("idle", L->stream);
    g_prof = nullptr;
    ctx_release(c, L, kind);
    return rc;
}

inline int ilog2_exact(uint64_t v) {  // -1 unless v is a power of two
    if (v == 0 || (v & (v - 1))) return -1;
    int l = 0;
    while ((1ull << l) < v) l++;
    return l;
}

}  // namespace fb
