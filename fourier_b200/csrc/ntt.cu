// ntt.cu - Fr number-theoretic transform, natural order in and out.
//
// Replaces FsFFTSettings::fft_fr as called through PianoFFTSettings::fft_left /
// fft_right (reference src/engine/piano.rs:1096-1103): out[k] = sum_j in[j] w^(jk)
// with w = 7^((r-1)/N); the inverse uses w^-1 and scales by 1/N.
//
// Decomposition.  N = R1*R2*...*Rp (each Rj <= 2^10).  Pass j transforms the most
// significant remaining input digit: the buffer is viewed as R rows by C = N/R
// columns, a CTA owns all R rows of B adjacent columns (R*B elements staged in
// shared memory, limb-plane layout), runs log2(R) radix-2 DIF stages there, then
// multiplies element (k, c) by the inter-pass twiddle w_N^(k * (c/Q) * Q) and
// scatters it to  (c/Q)*(R*Q) + k*Q + (c mod Q),  Q = R1*...*R(j-1).  After the
// last pass the digits are fully reversed, i.e. the output is in natural order.
// Global traffic per pass is one 32-byte read and one 32-byte write per element,
// in segments of B*32 bytes (reads) and B*32 or R*32 bytes (writes).
#include "common.cuh"

namespace fb {

namespace {

__constant__ uint32_t kRoot32[2][8] = {
    // 7^((r-1)/2^32) and its inverse, Montgomery form
    {0x5f0e466au, 0xb9b58d8cu, 0x1819d7ecu, 0x5b1b4c80u, 0x52a31e64u, 0x0af53ae3u, 0x19e9b27bu, 0x5bf3addau},
    {0xdcf3219au, 0x4256481au, 0x96b6cad3u, 0x45f37b7fu, 0x5f7a3b27u, 0xf9c3f1d7u, 0x658afd43u, 0x2d2fc049u}};

__device__ __forceinline__ Fr fr_pow_u64(Fr base, uint64_t e) {
    Fr r = Fr::one();
    while (e) {
        if (e & 1) r = fe_mul<FrParams>(r, base);
        e >>= 1;
        if (e) base = fe_sqr<FrParams>(base);
    }
    return r;
}

// consts[0] = w_N, consts[1] = w_N^-1, consts[2] = 1/N
__global__ void k_ntt_consts(Fr* consts, int log_n) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    for (int d = 0; d < 2; d++) {
        Fr w;
        for (int i = 0; i < 8; i++) w.v[i] = kRoot32[d][i];
        for (int i = log_n; i < 32; i++) w = fe_sqr<FrParams>(w);
        consts[d] = w;
    }
    // 1/2 = (r+1)/2 -> Montgomery
    Fr half = Fr::modulus();
    // (r+1)/2: r is odd, so (r+1)/2 = (r>>1) + 1
    for (int i = 0; i < 8; i++) half.v[i] = (half.v[i] >> 1) | (i < 7 ? (Fr::modulus().v[i + 1] << 31) : 0u);
    half.v[0] += 1;  // r>>1 ends in ...0000 (r = ...00000001), no carry
    half = fe_to_mont<FrParams>(half);
    Fr ninv = Fr::one();
    for (int i = 0; i < log_n; i++) ninv = fe_mul<FrParams>(ninv, half);
    consts[2] = ninv;
}

// out[x] = base^(x << shift) [* scale]
__global__ void k_pow_table(Fr* out, uint32_t count, const Fr* base, int shift, const Fr* scale) {
    uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= count) return;
    Fr r = fr_pow_u64(*base, (uint64_t)x << shift);
    if (scale) r = fe_mul<FrParams>(r, *scale);
    out[x] = r;
}

// tile twiddles: for every log_r in 1..10, w_R^j for j < R/2 at offset (R/2 - 1)
__global__ void k_tile_twiddles(Fr* out, int dir) {
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;  // 0 .. 1022
    if (idx >= 1023) return;
    int log_half = 31 - __clz(idx + 1);  // R/2 = 2^log_half
    uint32_t j = idx + 1 - (1u << log_half);
    int log_r = log_half + 1;
    Fr w;
    for (int i = 0; i < 8; i++) w.v[i] = kRoot32[dir][i];
    for (int i = log_r; i < 32; i++) w = fe_sqr<FrParams>(w);
    out[idx] = fr_pow_u64(w, j);
}

struct PassArgs {
    const Fr* in;
    Fr* out;
    int log_r, log_b, log_q, lo_bits;
    unsigned long long C;
    int twiddle, scale;
    const Fr* tw_lo;
    const Fr* tw_hi;
    const Fr* tw_full;  // single-multiply table of this pass boundary, or nullptr (then tw_hi * tw_lo)
    const Fr* tile_tw;
    const Fr* n_inv;
    unsigned long long batch_stride;  // elements between consecutive transforms of a batch (blockIdx.y)
    int pipe;  // 1: the tile is loaded with cp.async (small transforms: latency-bound, see ntt_run)
};

// 16-byte asynchronous copy global -> shared (LDGSTS): the whole tile is in flight at once, with no register staging
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ Fr lds_fr(const uint4* p0, const uint4* p1, uint32_t e) {
    Fr r;
    uint4 a = p0[e], b = p1[e];
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void sts_fr(uint4* p0, uint4* p1, uint32_t e, const Fr& r) {
    p0[e] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
    p1[e] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}
__device__ __forceinline__ Fr ldg_fr(const Fr* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    Fr r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}

// One group of G DIF stages (s_hi .. s_hi-G+1) over the staged tile.  A work item is the set of
// 2^G rows  base | (m << s_lo)  (m < 2^G; base has zero bits in [s_lo, s_hi]) of one column.
// LAST: the group ends at stage 0 (s_lo == 0), where the twiddle exponent depends on m alone, so the
// butterflies whose twiddle is 1 (all of stage 0, half of stage 1, a quarter of stage 2) are plain add/sub.
template <int G, bool LAST>
__device__ __forceinline__ void ntt_group(uint4* s0, uint4* s1, const uint4* t0, const uint4* t1, int s_hi, int log_r, int log_b,
                                          uint32_t tile, uint32_t tid, uint32_t nt) {
    constexpr int M = 1 << G;
    const int s_lo = LAST ? 0 : s_hi - G + 1;
    const uint32_t B = 1u << log_b;
    for (uint32_t q = tid; q < (tile >> G); q += nt) {
        const uint32_t cb = q & (B - 1), qr = q >> log_b;
        const uint32_t low = LAST ? 0u : (qr & ((1u << s_lo) - 1u));
        const uint32_t base = LAST ? (qr << G) : (((qr >> s_lo) << (s_hi + 1)) | low);
        Fr x[M];
#pragma unroll
        for (int m = 0; m < M; m++) x[m] = lds_fr(s0, s1, ((base | ((uint32_t)m << s_lo)) << log_b) | cb);
#pragma unroll
        for (int t = G - 1; t >= 0; t--) {
            const int st = s_lo + t;  // stage: rows m and m + 2^t differ in bit st
#pragma unroll
            for (int m = 0; m < M; m++) {
                if (m & (1 << t)) continue;
                Fr u = x[m], v = x[m | (1 << t)];
                x[m] = fe_add<FrParams>(u, v);
                Fr dif = fe_sub<FrParams>(u, v);
                // twiddle w_R^(j * 2^(log_r-1-st)),  j = row index mod 2^st
                if (LAST) {
                    const uint32_t j = (uint32_t)(m & ((1 << t) - 1));
                    if (j != 0) dif = fe_mul<FrParams>(dif, lds_fr(t0, t1, j << (log_r - 1 - st)));
                } else {
                    const uint32_t j = low | ((uint32_t)(m & ((1 << t) - 1)) << s_lo);
                    dif = fe_mul<FrParams>(dif, lds_fr(t0, t1, j << (log_r - 1 - st)));
                }
                x[m | (1 << t)] = dif;
            }
        }
#pragma unroll
        for (int m = 0; m < M; m++) sts_fr(s0, s1, ((base | ((uint32_t)m << s_lo)) << log_b) | cb, x[m]);
    }
}

// MAXG: stages per register-resident group.  3: 8 elements per thread, 128 registers, 16 warps/SM;
// 2: 4 elements per thread, 80 registers, 24 warps/SM (one more shared-memory round trip per 6 stages).
// (A 64-register build of the latter, 32 warps/SM, measured no faster: profiles/r1_v7_ntt_sweep.txt.)
template <int MAXG, int MINB>
__global__ void __launch_bounds__(256, MINB) k_ntt_pass(PassArgs a) {
    extern __shared__ uint4 smem[];
    const uint32_t R = 1u << a.log_r, B = 1u << a.log_b;
    const uint32_t tile = R << a.log_b;
    uint4* s0 = smem;
    uint4* s1 = smem + tile;
    uint4* t0 = smem + 2 * tile;
    uint4* t1 = t0 + (R >> 1);
    const uint32_t tid = threadIdx.x, nt = blockDim.x;
    const unsigned long long c0 = (unsigned long long)blockIdx.x << a.log_b;
    const Fr* in = a.in + (unsigned long long)blockIdx.y * a.batch_stride;
    Fr* out = a.out + (unsigned long long)blockIdx.y * a.batch_stride;

    if (a.pipe) {
        // every 16-byte half of every element of the tile (and the tile twiddles) as one asynchronous copy: a thread's
        // loads no longer wait for one another (LDG -> STS chains serialise on the scoreboard, one DRAM latency each)
        for (uint32_t e = tid; e < tile; e += nt) {
            uint32_t cb = e & (B - 1), r = e >> a.log_b;
            const uint4* src = reinterpret_cast<const uint4*>(in + ((unsigned long long)r * a.C + c0 + cb));
            cp_async16(s0 + e, src);
            cp_async16(s1 + e, src + 1);
        }
        for (uint32_t j = tid; j < (R >> 1); j += nt) {
            const uint4* src = reinterpret_cast<const uint4*>(a.tile_tw + j);
            cp_async16(t0 + j, src);
            cp_async16(t1 + j, src + 1);
        }
        cp_async_wait_all();
    } else {
        for (uint32_t e = tid; e < tile; e += nt) {
            uint32_t cb = e & (B - 1), r = e >> a.log_b;
            const uint4* src = reinterpret_cast<const uint4*>(in + ((unsigned long long)r * a.C + c0 + cb));
            s0[e] = src[0];
            s1[e] = src[1];
        }
        for (uint32_t j = tid; j < (R >> 1); j += nt) {
            const uint4* src = reinterpret_cast<const uint4*>(a.tile_tw + j);
            t0[j] = __ldg(src);
            t1[j] = __ldg(src + 1);
        }
    }
    __syncthreads();

    // radix-2 decimation-in-frequency stages (natural in, bit-reversed out), taken in groups of up
    // to 3: a thread pulls the 2^g rows that interact over the group's stages into registers,
    // runs the g stages there and writes them back - one shared-memory round trip and one
    // barrier per group instead of per stage.  The short groups come first so that the last group,
    // the one with the multiplier-free butterflies, is a full one: 10 = 2+2+3+3, 8 = 2+3+3, 7 = 2+2+3.
    for (int s_hi = a.log_r - 1; s_hi >= 0;) {
        const int left = s_hi + 1;
        int g;
        if (MAXG == 3) {
            if (left <= 3) g = left;
            else if (left % 3 == 0) g = 3;
            else if (left % 3 == 2) g = 2;
            else g = left == 4 ? 1 : 2;  // left % 3 == 1: 2 + 2 + 3k, except 4 = 1 + 3
        } else {
            g = (left <= 2) ? left : ((left & 1) ? 1 : 2);  // odd: 1 + 2k
        }
        const bool last = (g == left);
        if (last) {
            if (MAXG == 3 && g == 3) ntt_group<3, true>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
            else if (g == 2) ntt_group<2, true>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
            else ntt_group<1, true>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
        } else {
            if (MAXG == 3 && g == 3) ntt_group<3, false>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
            else if (g == 2) ntt_group<2, false>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
            else ntt_group<1, false>(s0, s1, t0, t1, s_hi, a.log_r, a.log_b, tile, tid, nt);
        }
        __syncthreads();
        s_hi -= g;
    }

    const uint32_t lo_mask = (1u << a.lo_bits) - 1;
    const unsigned long long qmask = (1ull << a.log_q) - 1;
    Fr ninv;
    if (a.scale) ninv = *a.n_inv;
    auto coords = [&](uint32_t e, uint32_t& k, uint32_t& cb) {
        if (a.log_q == 0) {  // transposing pass: rows are contiguous in the output
            k = e & (R - 1);
            cb = e >> a.log_r;
        } else {
            cb = e & (B - 1);
            k = e >> a.log_b;
        }
    };
    auto emit = [&](uint32_t k, uint32_t cb) {
        uint32_t pos = __brev(k) >> (32 - a.log_r);
        if (a.log_r == 0) pos = 0;
        Fr val = lds_fr(s0, s1, (pos << a.log_b) | cb);
        const unsigned long long c = c0 + cb;
        const unsigned long long rest = c >> a.log_q;
        if (a.twiddle) {
            if (a.tw_full) {
                val = fe_mul<FrParams>(val, ldg_fr(a.tw_full + (unsigned long long)k * rest));
            } else {
                unsigned long long x = ((unsigned long long)k * rest) << a.log_q;
                Fr t = fe_mul<FrParams>(ldg_fr(a.tw_hi + (x >> a.lo_bits)), ldg_fr(a.tw_lo + (x & lo_mask)));
                val = fe_mul<FrParams>(val, t);
            }
        }
        if (a.scale) val = fe_mul<FrParams>(val, ninv);
        unsigned long long o = (rest << (a.log_q + a.log_r)) | ((unsigned long long)k << a.log_q) | (c & qmask);
        uint4* dst = reinterpret_cast<uint4*>(out + o);
        dst[0] = make_uint4(val.v[0], val.v[1], val.v[2], val.v[3]);
        dst[1] = make_uint4(val.v[4], val.v[5], val.v[6], val.v[7]);
    };
    for (uint32_t e = tid; e < tile; e += nt) {
        uint32_t k, cb;
        coords(e, k, cb);
        emit(k, cb);
    }
}

constexpr int kTileLog = 11;  // elements staged per CTA = 2^11 (64 KiB of shared memory)

}  // namespace

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

// Tuning knobs (defaults are the measured optimum, profiles/r1_v7_ntt_sweep.txt): largest radix per pass,
// elements staged per CTA, columns per tile, stages per register group, elements per thread.
static int ntt_max_log_r() { static int v = env_int("FB_NTT_MAXR", 10); return v < 1 ? 1 : (v > 10 ? 10 : v); }
static int ntt_tile_log() { static int v = env_int("FB_NTT_TILE", kTileLog); return v < 6 ? 6 : (v > 12 ? 12 : v); }
static int ntt_group_stages() { static int v = env_int("FB_NTT_G", 2); return v == 3 ? 3 : 2; }
static int ntt_threads_div() { static int v = env_int("FB_NTT_TDIV", 4); return v < 2 ? 2 : (v > 16 ? 16 : v); }
static int ntt_full_table_log() { static int v = env_int("FB_NTT_FULL", kNttFullTableLog); return v; }
static int ntt_pipe() { static int v = env_int("FB_NTT_PIPE", -1); return v; }  // -1: by size
static int ntt_max_log_b() { static int v = env_int("FB_NTT_MAXB", 3); return v < 0 ? 0 : v; }

NttPlan ntt_make_plan(int log_n) {
    NttPlan p;
    const int maxr = ntt_max_log_r();
    int np = (log_n + maxr - 1) / maxr;
    if (np < 1) np = 1;
    if (np > 4) np = 4;
    p.npass = np;
    int left = log_n;
    for (int j = 0; j < np; j++) {  // as even as possible, larger radices first
        p.log_r[j] = (left + (np - j) - 1) / (np - j);
        left -= p.log_r[j];
    }
    return p;
}

void ntt_build_tile_twiddles(TileTwiddles& t, cudaStream_t st) {
    for (int d = 0; d < 2; d++) {
        FB_CUDA(cudaMalloc(&t.base[d], 1023 * sizeof(Fr)));
        (++fb::g_launches, k_tile_twiddles<<<8, 128, 0, st>>>(t.base[d], d));
    }
    FB_CUDA(cudaGetLastError());
}

void ntt_build_tables(NttTables& t, int log_n, cudaStream_t st) {
    t.log_n = log_n;
    t.lo_bits = (log_n + 1) / 2;
    int hi_bits = log_n - t.lo_bits;
    uint32_t nlo = 1u << t.lo_bits, nhi = 1u << hi_bits;
    Fr* consts;
    FB_CUDA(cudaMalloc(&consts, 3 * sizeof(Fr)));
    (++fb::g_launches, k_ntt_consts<<<1, 1, 0, st>>>(consts, log_n));
    for (int d = 0; d < 2; d++) {
        FB_CUDA(cudaMalloc(&t.lo[d], nlo * sizeof(Fr)));
        FB_CUDA(cudaMalloc(&t.hi[d], nhi * sizeof(Fr)));
        (++fb::g_launches, k_pow_table<<<(nlo + 127) / 128, 128, 0, st>>>(t.lo[d], nlo, consts + d, 0, nullptr));
        (++fb::g_launches, k_pow_table<<<(nhi + 127) / 128, 128, 0, st>>>(t.hi[d], nhi, consts + d, t.lo_bits, d == 1 ? consts + 2 : nullptr));
    }
    FB_CUDA(cudaMalloc(&t.hi_unscaled_inv, nhi * sizeof(Fr)));
    (++fb::g_launches, k_pow_table<<<(nhi + 127) / 128, 128, 0, st>>>(t.hi_unscaled_inv, nhi, consts + 1, t.lo_bits, nullptr));
    t.consts = consts;
    // single-multiply tables per pass boundary (pass j spans N >> log_q_j exponents)
    NttPlan plan = ntt_make_plan(log_n);
    int log_q = 0;
    for (int j = 0; j + 1 < plan.npass; j++) {
        const int log_cnt = log_n - log_q;
        if (log_cnt <= ntt_full_table_log()) {
            const uint32_t cnt = 1u << log_cnt;
            for (int d = 0; d < 2; d++) {
                FB_CUDA(cudaMalloc(&t.full[d][j], (size_t)cnt * sizeof(Fr)));
                (++fb::g_launches, k_pow_table<<<(cnt + 127) / 128, 128, 0, st>>>(t.full[d][j], cnt, consts + d, log_q,
                                                                                 (d == 1 && j == 0) ? consts + 2 : nullptr));
            }
        }
        log_q += plan.log_r[j];
    }
    FB_CUDA(cudaGetLastError());
}

void ntt_free_tables(NttTables& t) {
    for (int d = 0; d < 2; d++) {
        if (t.lo[d]) cudaFree(t.lo[d]);
        if (t.hi[d]) cudaFree(t.hi[d]);
        t.lo[d] = t.hi[d] = nullptr;
    }
    for (int d = 0; d < 2; d++)
        for (int j = 0; j < 3; j++) {
            if (t.full[d][j]) cudaFree(t.full[d][j]);
            t.full[d][j] = nullptr;
        }
    if (t.hi_unscaled_inv) cudaFree(t.hi_unscaled_inv);
    if (t.consts) cudaFree(t.consts);
    t.hi_unscaled_inv = nullptr;
    t.consts = nullptr;
    t.log_n = -1;
}

void ntt_run(const NttTables& tab, const TileTwiddles& tt, const Fr* in, Fr* out, Fr* scratch, bool inverse,
             cudaStream_t st, size_t batch) {
    if (batch == 0) return;
    if (batch > 65535) throw CudaError{"ntt: batch too large"};
    // the shared-memory opt-in is per device: a process may hold contexts on several GPUs
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    FB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load()) {
        FB_CUDA(cudaFuncSetAttribute(k_ntt_pass<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        FB_CUDA(cudaFuncSetAttribute(k_ntt_pass<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true);
    }
    const int log_n = tab.log_n;
    const int d = inverse ? 1 : 0;
    NttPlan plan = ntt_make_plan(log_n);
    const Fr* src = in;
    int log_q = 0;
    for (int j = 0; j < plan.npass; j++) {
        PassArgs a;
        a.log_r = plan.log_r[j];
        int log_c = log_n - a.log_r;
        a.C = 1ull << log_c;
        a.log_b = ntt_tile_log() - a.log_r;
        if (a.log_b < 0) a.log_b = 0;
        if (a.log_b > log_c) a.log_b = log_c;
        if (a.log_b > ntt_max_log_b()) a.log_b = ntt_max_log_b();
        // enough tiles per SM that the last, partial wave is short (2^20 = 1024 x 1024: 512 tiles of 2 columns
        // leave 4 tiles on some SMs and 3 on others; 1024 tiles of one column balance to 7 vs 6.9)
        while (a.log_b > 0 && ((a.C * batch) >> a.log_b) < 8ull * kNumSMs) a.log_b--;
        a.log_q = log_q;
        a.lo_bits = tab.lo_bits;
        bool last = (j == plan.npass - 1);
        a.twiddle = last ? 0 : 1;
        a.scale = (inverse && plan.npass == 1) ? 1 : 0;
        a.tw_lo = tab.lo[d];
        // the inverse transform's 1/N rides on the first pass's twiddle table
        a.tw_hi = inverse ? (j == 0 ? tab.hi[1] : tab.hi_unscaled_inv) : tab.hi[0];
        a.tw_full = (j < 3) ? tab.full[d][j] : nullptr;
        a.tile_tw = a.log_r >= 1 ? tt.base[d] + ((1u << (a.log_r - 1)) - 1) : tt.base[d];
        a.n_inv = tab.consts + 2;
        a.in = src;
        Fr* dst;
        // the last pass may run in place (a CTA writes exactly the positions it read); the others ping-pong
        if (last) dst = out;
        else dst = (j % 2 == 0) ? scratch : out;
        a.out = dst;
        uint32_t tile = 1u << (a.log_r + a.log_b);
        const uint32_t tdiv = (uint32_t)ntt_threads_div();
        uint32_t threads = tile / tdiv < 32 ? 32 : (tile / tdiv > 256 ? 256 : tile / tdiv);
        size_t smem = (size_t)tile * 32 + ((size_t)1 << a.log_r) * 16;
        unsigned long long blocks = a.C >> a.log_b;
        static const char* const names[4] = {"ntt.pass0", "ntt.pass1", "ntt.pass2", "ntt.pass3"};
        prof_mark(names[j], st);
        a.batch_stride = 1ull << log_n;
        // Measured on B200 (profiles/r2_v4_rejected_ntt_pipe_msm_ring.txt): with cp.async a thread's tile loads no longer wait for one
        // another (LDG -> STS pairs serialise on the scoreboard), which helps where nothing else hides the latency -
        // 2^16: 0.0325 -> 0.0312 ms - and costs 3-10 % at 2^18 / 2^20 where three resident CTAs per SM already overlap
        // one another's phases; fetching the inter-pass twiddles four elements ahead in the output phase cost 6 % at 2^22+
        // (registers) and is not kept.
        a.pipe = ntt_pipe() >= 0 ? (ntt_pipe() & 1) : ((batch << log_n) <= ((size_t)1 << 16) ? 1 : 0);
        const dim3 grid((unsigned)blocks, (unsigned)batch);
        if (ntt_group_stages() == 3) (++fb::g_launches, k_ntt_pass<3, 2><<<grid, threads, smem, st>>>(a));
        else (++fb::g_launches, k_ntt_pass<2, 3><<<grid, threads, smem, st>>>(a));
        src = dst;
        log_q += a.log_r;
    }
    prof_mark("post", st);
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
