// common.cuh - shared declarations for the kernels and the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "g1.cuh"

namespace fb {

constexpr int kNumSMs = 148;  // B200
extern std::atomic<unsigned long long> g_launches;  // kernel launches issued by this library (api.cu); contexts on several host threads count together

struct CudaError {
    std::string msg;
};

#define FB_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess) {                                                                        \
            char _b[512];                                                                               \
            snprintf(_b, sizeof _b, "CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                     cudaGetErrorString(_e));                                                           \
            throw fb::CudaError{_b};                                                                    \
        }                                                                                               \
    } while (0)

// Optional per-phase timing with CUDA events on the launching stream (bench.py's
// roofline numbers).  prof_mark(name) says "phase `name` starts here"; the time to
// the next mark is charged to it.  A no-op unless enabled through fb_prof_enable.
struct Prof {
    bool on = false;
    std::vector<std::pair<const char*, cudaEvent_t>> marks;
    std::vector<cudaEvent_t> pool;
};
extern thread_local Prof* g_prof;
inline void prof_mark(const char* name, cudaStream_t st) {
    Prof* p = g_prof;
    if (!p || !p->on) return;
    cudaEvent_t e;
    if (!p->pool.empty()) {
        e = p->pool.back();
        p->pool.pop_back();
    } else {
        cudaEventCreate(&e);
    }
    cudaEventRecord(e, st);
    p->marks.emplace_back(name, e);
}

// grow-only device scratch buffer
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* get(size_t bytes) {
        if (bytes > cap) {
            if (p) FB_CUDA(cudaFree(p));
            p = nullptr;
            cap = 0;
            size_t want = bytes + bytes / 8 + 256;
            FB_CUDA(cudaMalloc(&p, want));
            cap = want;
        }
        return p;
    }
    template <class T>
    T* as(size_t count) { return (T*)get(count * sizeof(T)); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// ---------------------------------------------------------------- NTT (ntt.cu)
// Twiddle tables for one transform size N = 2^log_n (Montgomery form, device).
struct NttTables {
    int log_n = -1;
    int lo_bits = 0;           // w^e = hi[e >> lo_bits] * lo[e & (2^lo_bits - 1)]
    Fr* lo[2] = {nullptr, nullptr};   // [0] forward, [1] inverse
    Fr* hi[2] = {nullptr, nullptr};   // inverse hi table is pre-scaled by 1/N
    Fr* hi_unscaled_inv = nullptr;    // inverse hi table without the 1/N factor
    Fr* consts = nullptr;             // device: [0] w_N, [1] w_N^-1, [2] 1/N
    // full[d][j]: single-multiply inter-pass twiddles of pass j, w_N^(y << log_q_j) for y < N >> log_q_j (the inverse
    // table of pass 0 carries 1/N); nullptr where the table would exceed kNttFullTableLog entries (then hi*lo)
    Fr* full[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
};
constexpr int kNttFullTableLog = 24;  // 2^24 entries = 512 MiB per direction, built only for sizes that are used

struct NttPlan {
    int npass = 0;
    int log_r[4] = {0, 0, 0, 0};
};

void ntt_build_tables(NttTables& t, int log_n, cudaStream_t st);
void ntt_free_tables(NttTables& t);
// tile twiddles w_R^j (j < R/2) for every radix 2^1..2^10, forward and inverse
struct TileTwiddles {
    Fr* base[2] = {nullptr, nullptr};  // entry for radix log_r starts at offset (1<<(log_r-1)) - 1
};
void ntt_build_tile_twiddles(TileTwiddles& t, cudaStream_t st);
NttPlan ntt_make_plan(int log_n);
// out may alias in; scratch must hold batch*N elements when the plan has >= 2 passes.
// batch > 1: `batch` transforms of the same size, back to back in in/out/scratch.
void ntt_run(const NttTables& tab, const TileTwiddles& tt, const Fr* in, Fr* out, Fr* scratch, bool inverse,
             cudaStream_t st, size_t batch = 1);

// -------------------------------------------------------------- poly (poly.cu)
// wire <-> device conversions and the Horner/quotient scan
void fr_from_be(const uint8_t* d_in, Fr* d_out, size_t n, int* d_err, cudaStream_t st);  // sets *d_err=1 on >= r
void fr_to_be(const Fr* d_in, uint8_t* d_out, size_t n, cudaStream_t st);
void fr_fill_zero(Fr* d, size_t n, cudaStream_t st);
// s_k = a_k + x*s_{k+1} (suffix Horner scan), s_n = 0.  Writes y = s_0 to d_y and,
// when d_q != nullptr, q_k = s_{k+1} for k < n-1 and q_{n-1} = 0 (the zero-padded
// quotient (p - y)/(X - x)).  d_x, d_y are single device Fr (Montgomery).
// batch > 1: `batch` polynomials of n coefficients back to back (same x); d_y holds batch values,
// d_q batch*n coefficients.
void horner_quotient(const Fr* d_a, size_t n, const Fr* d_x, Fr* d_y, Fr* d_q, DevBuf& scratch, cudaStream_t st,
                     size_t batch = 1);
// evalform.cu: y[b] = p_b(alpha) and q_hat[b][j] = ((p_b - y_b)/(X - alpha))(w^j) straight from the
// evaluations evals[b][j] = p_b(w^j), n = 2^tab.log_n, any alpha (domain points included)
void evalform_open(const NttTables& tab, const Fr* d_evals, size_t n, size_t batch, const Fr* d_alpha, Fr* d_y, Fr* d_qhat,
                   DevBuf& scratch, cudaStream_t st);

// the same for one column shard (elements j0 .. j0+cnt-1 of the row): begin -> 4 Fr of state per shard, exchanged between
// the shards; finish -> y and the shard's slice of q_hat.  `session` must be the same buffer in both calls.
void evalform_shard_begin(const NttTables& tab, const Fr* d_evals_slice, size_t n, size_t j0, size_t cnt, const Fr* d_alpha,
                          Fr* d_state, DevBuf& session, cudaStream_t st);
void evalform_shard_finish(const NttTables& tab, const Fr* d_evals_slice, size_t n, size_t j0, size_t cnt, const Fr* d_alpha,
                           const Fr* d_states, int nparts, Fr* d_y, Fr* d_qhat_slice, DevBuf& session, cudaStream_t st);

// ---------------------------------------------------------------- MSM (msm.cu)
// Window-table entries are padded to one 128-byte line (x in the first 64-byte DRAM atom):
// a gathered point is exactly one line, and the batched-affine rounds' first pass, which
// needs x only, touches one atom.  Untabulated bases stay packed (96 bytes).
constexpr uint32_t kTableStride = 128;
struct MsmBases {
    void* pts = nullptr;      // [rows][nwin][n] x 128 B when tabulated (window w holds 2^(c*w) * P), else [rows][n] x 96 B
    size_t n = 0;             // row length
    int c = 0;                // window bits of the table (0: no table)
    int nwin = 0;
    const char* xs = nullptr;  // optional: the x coordinates alone, one 64-byte slot per table entry (same indexing)
    uint32_t stride() const { return c > 0 ? kTableStride : (uint32_t)sizeof(G1Affine); }
};

struct MsmScratch {
    DevBuf counts, offsets, cursor, sorted, buckets, block_sums, seg_a, seg_t, order, misc, pts_a, pts_b, pref, where, stage, inv;
    // while profiling: [0] MSMs, [1] sorted entries, [2] output slots of all affine rounds, [3] survivors that go
    // through the XYZZ accumulation, [4] buckets  (u64 each; fb_msm_stats)
    DevBuf stats;
    bool stats_zeroed = false;
    // set by the caller when another MSM runs next to this one (the two streams of a commit + open pair, calls rotating over
    // several lanes): one more affine round then pays - the rounds fill the multiplier better than the XYZZ pass they shorten,
    // which only shows when something else can use the gaps (profiles/r2_v7_rounds_sweep.txt, r2_v9_ab_rounds_overlap.txt)
    bool overlapped = false;
};

// d_out[b] (device XYZZ) = sum_k scalars[b*len + k] * bases_b[k] for b < batch, where
// row b's bases start at pts + b*(nwin or 1)*bases.n.  Scalars are device Fr either
// in Montgomery form (scalars_mont) or canonical little-endian limbs.
// d_scalars2 != nullptr: batch = 2 over the SAME bases (row 0 = d_scalars, row 1 = d_scalars2, `len` each): the commit
// and the opening proof of one row in one pass (sort bookkeeping, rounds, reduction and their latency-bound tails once).
void msm_run(const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, bool scalars_mont, G1Xyzz* d_out,
             MsmScratch& s, cudaStream_t st, const Fr* d_scalars2 = nullptr);
// One MSM as S independent sub-MSMs on S streams: over point (column) ranges when there is one output (or the
// shared-bases pair), over row ranges when the batch has many rows.  A sub-MSM's latency-bound phases (sort
// bookkeeping, the gather-bound first pass, the per-launch inversion, the reduction tail) then run under another's
// multiplier-bound ones.  stream[0] is the caller's stream; the others fork from it and join it again.
constexpr int kMaxSplit = 4;
struct MsmExec {
    int S = 1;
    MsmScratch* scratch[kMaxSplit] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t stream[kMaxSplit] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t fork = nullptr;
    cudaEvent_t join[kMaxSplit] = {nullptr, nullptr, nullptr, nullptr};
    DevBuf* parts = nullptr;  // S x 2 XYZZ partial sums (column split)
};
void msm_run_split(const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, bool scalars_mont, G1Xyzz* d_out,
                   MsmExec& ex, const Fr* d_scalars2 = nullptr);
// table construction: out[row][w][k] = 2^(c*w) * in[row][k], w < nwin, kTableStride bytes per entry
void msm_build_table(const G1Affine* d_in, size_t n, size_t rows, int c, int nwin, void* d_out, DevBuf& scratch,
                     cudaStream_t st);
// a copy of the table's x coordinates in 64-byte slots for MsmBases::xs (cudaMalloc'ed; nullptr if disabled or out of memory)
const char* msm_alloc_xtable(const void* d_table, size_t entries, cudaStream_t st);
int msm_pick_window(size_t n, bool tabulated);
int msm_num_windows(int c);  // ceil(255 / c)

// -------------------------------------------------------------- points (g1ops.cu)
void g1_decompress_batch(const uint8_t* d_in48, G1Affine* d_out, size_t n, int* d_err, cudaStream_t st);
void g1_deserialize_batch(const uint8_t* d_in96, G1Affine* d_out, size_t n, int* d_err, cudaStream_t st);
void g1_compress_batch(const G1Affine* d_in, uint8_t* d_out48, size_t n, cudaStream_t st);
void g1_serialize_batch(const G1Affine* d_in, uint8_t* d_out96, size_t n, cudaStream_t st);
void g1_xyzz_to_affine_batch(const G1Xyzz* d_in, G1Affine* d_out, size_t n, cudaStream_t st);
// out[k] = scalars[k] * G (fixed-base comb on the generator), scalars Montgomery Fr
void g1_fixed_base_mul(const Fr* d_scalars, size_t n, G1Affine* d_out, DevBuf& scratch, cudaStream_t st);
// sum of the shards' partial commitments / proofs (nparts x (96 | 32 | 96) bytes) -> 48 | 32 | 48 bytes, on the device
void g1_combine_parts(const uint8_t* d_parts224, size_t nparts, uint8_t* d_out128, int* d_err, cudaStream_t st);
// d_out (1 XYZZ) = sum of n affine points (tree sum)
void g1_sum_affine(const G1Affine* d_in, size_t n, G1Xyzz* d_out, DevBuf& scratch, cudaStream_t st);


// ------------------------------------------------------------------ SRS (srs.cu)
void srs_secret_to_fr(const uint8_t* d_secret32, Fr* d_out, cudaStream_t st);
void srs_lagrange_at(const Fr* d_tau, int log_n, Fr* d_out, Fr* d_hdr3, cudaStream_t st);
void srs_powers(const Fr* d_tau, size_t n, Fr* d_out, cudaStream_t st);
void srs_outer(const Fr* d_r, size_t nrows, const Fr* d_l, size_t T, Fr* d_out, cudaStream_t st);

}  // namespace fb
