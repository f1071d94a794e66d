// msm.cu - G1 multi-scalar multiplication (Pippenger, signed digits, sort-based
// bucket accumulation).
//
// Replaces rust_kzg_blst::kzg_proofs::g1_linear_combination (reference
// src/engine/piano.rs:150-158, 228-236, 286-295) together with the fixed-base
// tables of PianoPrecomputation (piano.rs:469-504): a resident base set may carry
// per-window multiples 2^(c*w)*P, in which case every window of every scalar
// falls into ONE set of 2^(c-1) buckets and no doubling ladder is needed at the end.
//
// Pipeline (all on one stream):
//   1. k_digits<false>: scalar -> W signed c-bit digits; histogram of bucket sizes
//   2. exclusive scan of the histogram                       (bucket offsets)
//   3. k_digits<true>: same digits again; scatter (point index | sign) into its
//      bucket's slice (a counting sort; order inside a bucket is irrelevant)
//   4. k_accumulate: one thread per bucket, XYZZ += affine (8M+2S), points gathered
//      by index from the resident table
//   5. k_seg_reduce levels: sum_d d*B_d by blocked running sums (segments of 8)
//   6. k_finish: add the per-window results (doubling ladder when not tabulated)
#include "common.cuh"

namespace fb {

namespace {

constexpr int kSegLog = 3;  // bucket-reduction segment = 8 buckets per thread per level
constexpr int kSeg = 1 << kSegLog;

struct DigitArgs {
    const Fr* scalars;
    uint32_t n;       // total scalars = batch * len
    uint32_t len;     // scalars per row
    uint32_t stride;  // row length of the resident table (index = (row*nwin + w)*stride + col)
    int c, nwin;
    int mont;       // scalars are in Montgomery form
    int tabulated;  // single bucket set; entry = w*n + j
    uint32_t buckets_per_win;  // 2^(c-1) + 1 (index = |digit|, index 0 unused)
    uint32_t* counts;   // histogram (pass 0) / cursors (pass 1)
    uint32_t* sorted;
};

template <bool SCATTER>
__global__ void __launch_bounds__(256) k_digits(DigitArgs a) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const uint32_t row = j / a.len, col = j - row * a.len;
    Fr k = a.scalars[j];
    if (a.mont) k = fe_from_mont<FrParams>(k);
    uint32_t kk[9];
#pragma unroll
    for (int i = 0; i < 8; i++) kk[i] = k.v[i];
    kk[8] = 0;
    const uint32_t mask = (1u << a.c) - 1, half = 1u << (a.c - 1);
    uint32_t carry = 0;
    for (int w = 0; w < a.nwin; w++) {
        uint32_t pos = (uint32_t)w * a.c;
        uint32_t limb = pos >> 5, off = pos & 31;
        uint32_t raw = 0;
        if (limb < 8) {
            uint64_t two = ((uint64_t)kk[limb + 1] << 32) | kk[limb];
            raw = (uint32_t)(two >> off) & mask;
        }
        raw += carry;
        uint32_t neg = raw > half ? 1u : 0u;
        carry = neg;
        uint32_t mag = neg ? (1u << a.c) - raw : raw;
        if (mag == 0) continue;
        uint32_t bucket = (a.tabulated ? row : row * a.nwin + (uint32_t)w) * a.buckets_per_win + mag;
        if (!SCATTER) {
            atomicAdd(a.counts + bucket, 1u);
        } else {
            uint32_t slot = atomicAdd(a.counts + bucket, 1u);
            uint32_t idx = a.tabulated ? (row * a.nwin + (uint32_t)w) * a.stride + col : row * a.stride + col;
            a.sorted[slot] = idx | (neg << 31);
        }
    }
}

// ---- exclusive scan of uint32 (three small kernels, 4096 items per block)
__global__ void __launch_bounds__(1024) k_scan_blocks(const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* block_sums) {
    __shared__ uint32_t warp_tot[32];
    uint32_t base = blockIdx.x * 4096u + threadIdx.x * 4u;
    uint32_t v[4], s = 0;
    for (int i = 0; i < 4; i++) {
        v[i] = (base + i < n) ? in[base + i] : 0u;
        s += v[i];
    }
    uint32_t incl = s;
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if ((threadIdx.x & 31) >= d) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t w = warp_tot[threadIdx.x], wi = w;
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, wi, d);
            if (threadIdx.x >= d) wi += t;
        }
        warp_tot[threadIdx.x] = wi - w;  // exclusive
        if (threadIdx.x == 31 && block_sums) block_sums[blockIdx.x] = wi;
    }
    __syncthreads();
    uint32_t excl = warp_tot[threadIdx.x >> 5] + incl - s;
    for (int i = 0; i < 4; i++) {
        if (base + i < n) out[base + i] = excl;
        excl += v[i];
    }
}
__global__ void k_scan_add(uint32_t* out, uint32_t n, const uint32_t* block_offs) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] += block_offs[i / 4096u];
}

void exclusive_scan(const uint32_t* in, uint32_t* out, uint32_t n, DevBuf& tmp, cudaStream_t st) {
    uint32_t nb = (n + 4095) / 4096;
    if (nb > 4096) throw CudaError{"exclusive_scan: input too large"};
    uint32_t* bs = tmp.as<uint32_t>(2 * 4096);
    (++fb::g_launches, k_scan_blocks<<<nb, 1024, 0, st>>>(in, out, n, bs));
    if (nb > 1) {
        (++fb::g_launches, k_scan_blocks<<<1, 1024, 0, st>>>(bs, bs + 4096, nb, nullptr));
        (++fb::g_launches, k_scan_add<<<(n + 255) / 256, 256, 0, st>>>(out, n, bs + 4096));
    }
}

__device__ __forceinline__ G1Affine ldg_affine(const G1Affine* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    G1Affine r;
    uint4 t;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        t = __ldg(q + i);
        r.x.v[4 * i] = t.x; r.x.v[4 * i + 1] = t.y; r.x.v[4 * i + 2] = t.z; r.x.v[4 * i + 3] = t.w;
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        t = __ldg(q + 3 + i);
        r.y.v[4 * i] = t.x; r.y.v[4 * i + 1] = t.y; r.y.v[4 * i + 2] = t.z; r.y.v[4 * i + 3] = t.w;
    }
    return r;
}

// ---- balanced bucket accumulation.  A bucket with cnt entries is cut into
// nt = max(ceil(cnt/cap), isqrt(cnt)) tasks of equal length (1 task when cnt <= cap),
// so that no thread walks more than ~cap entries however skewed the digits are (the
// top window of a 255-bit scalar, or structured inputs such as all-equal scalars,
// put thousands of points into a few buckets).  One thread per task produces a
// partial sum; one thread per bucket then adds its partials.
__device__ __forceinline__ uint32_t tasks_for(uint32_t cnt, uint32_t cap) {
    if (cnt <= cap) return cnt ? 1u : 0u;
    uint32_t a = (cnt + cap - 1) / cap;
    uint32_t r = (uint32_t)sqrtf((float)cnt);
    return a > r ? a : r;
}

__global__ void __launch_bounds__(256) k_task_count(const uint32_t* __restrict__ counts, uint32_t nbuckets, uint32_t cap,
                                                    uint32_t* __restrict__ ntasks) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nbuckets) ntasks[b] = tasks_for(counts[b], cap);
}

// task t of bucket b covers entries [start, start+len)
__global__ void __launch_bounds__(256) k_task_fill(const uint32_t* __restrict__ counts, const uint32_t* __restrict__ offsets,
                                                   const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ toff,
                                                   uint32_t nbuckets, uint2* __restrict__ tasks, uint32_t* __restrict__ total) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    uint32_t nt = ntasks[b], cnt = counts[b], base = offsets[b], t0 = toff[b];
    if (b == nbuckets - 1) *total = t0 + nt;
    if (!nt) return;
    uint32_t len = (cnt + nt - 1) / nt;
    for (uint32_t k = 0; k < nt; k++) {
        uint32_t lo = k * len;
        uint32_t l = lo >= cnt ? 0u : (cnt - lo < len ? cnt - lo : len);
        tasks[t0 + k] = make_uint2(base + lo, l);
    }
}

__global__ void __launch_bounds__(128) k_accumulate(const G1Affine* __restrict__ pts, const uint32_t* __restrict__ sorted,
                                                    const uint2* __restrict__ tasks, const uint32_t* __restrict__ total,
                                                    G1Xyzz* __restrict__ partial) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *total) return;
    uint2 tk = tasks[t];
    G1Xyzz acc = G1Xyzz::inf();
    for (uint32_t e = tk.x; e < tk.x + tk.y; e++) {
        uint32_t ent = sorted[e];
        G1Affine p = ldg_affine(pts + (ent & 0x7fffffffu));
        if (ent >> 31) p.y = fe_neg<FpParams>(p.y);
        g1_add_mixed(acc, p);
    }
    partial[t] = acc;
}

__global__ void __launch_bounds__(128) k_combine(const G1Xyzz* __restrict__ partial, const uint32_t* __restrict__ ntasks,
                                                 const uint32_t* __restrict__ toff, uint32_t nbuckets,
                                                 G1Xyzz* __restrict__ buckets) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    uint32_t nt = ntasks[b], t0 = toff[b];
    G1Xyzz acc = G1Xyzz::inf();
    if (nt) acc = partial[t0];
    for (uint32_t k = 1; k < nt; k++) g1_add(acc, partial[t0 + k]);
    buckets[b] = acc;
}

// Blocked running sums.  For segment s of X (length m, zero-padded):
//   A = sum_i X[sL+i],  T = sum_i i*X[sL+i]
// outputs  X'[s] = 2^l * A  (so that S(X) = sum_s T_s + S(X')), and
//          P'[s] = T + sum_i P[sL+i]   (pending plain sums carried along).
// Several independent windows are laid out back to back with stride `stride_in`.
__global__ void __launch_bounds__(128) k_seg_reduce(const G1Xyzz* __restrict__ X, const G1Xyzz* __restrict__ P, uint32_t m,
                                                    uint32_t stride_in, G1Xyzz* __restrict__ Xo, G1Xyzz* __restrict__ Po,
                                                    uint32_t mo, uint32_t nwin) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= mo * nwin) return;
    uint32_t w = t / mo, s = t % mo;
    const G1Xyzz* x = X + (size_t)w * stride_in;
    const G1Xyzz* p = P ? P + (size_t)w * stride_in : nullptr;
    uint32_t lo = s << kSegLog;
    uint32_t hi = lo + kSeg < m ? lo + kSeg : m;
    G1Xyzz run = G1Xyzz::inf(), acc = G1Xyzz::inf(), pend = G1Xyzz::inf();
    for (uint32_t i = hi; i > lo; i--) {
        g1_add(run, x[i - 1]);
        if (i - 1 > lo) g1_add(acc, run);
        if (p) g1_add(pend, p[i - 1]);
    }
    g1_add(pend, acc);
    for (int d = 0; d < kSegLog; d++) run = g1_dbl(run);
    Xo[(size_t)w * mo + s] = run;
    Po[(size_t)w * mo + s] = pend;
}

// top of the reduction + window combine.  Thread w: S_w = sum_i i*X_w[i] + sum_i P_w[i],
// scaled by 2^(c*w); thread 0 then adds the nwin results.
__global__ void k_finish(const G1Xyzz* X, const G1Xyzz* P, uint32_t m, uint32_t nwin, int c, G1Xyzz* out) {
    __shared__ G1Xyzz part[64];
    uint32_t w = threadIdx.x;
    X += (size_t)blockIdx.x * nwin * m;  // one block per output (batch row)
    if (P) P += (size_t)blockIdx.x * nwin * m;
    out += blockIdx.x;
    if (w < nwin) {
        const G1Xyzz* x = X + (size_t)w * m;
        G1Xyzz run = G1Xyzz::inf(), acc = G1Xyzz::inf();
        for (uint32_t i = m; i > 1; i--) {
            g1_add(run, x[i - 1]);
            g1_add(acc, run);
        }
        if (P)
            for (uint32_t i = 0; i < m; i++) g1_add(acc, P[(size_t)w * m + i]);
        for (uint32_t d = 0; d < (uint32_t)c * w; d++) acc = g1_dbl(acc);
        part[w] = acc;
    }
    __syncthreads();
    if (w == 0) {
        G1Xyzz total = part[0];
        for (uint32_t i = 1; i < nwin; i++) g1_add(total, part[i]);
        *out = total;
    }
}

// table[row][w][k] = 2^(c*w) * P[row][k].  One thread per point: the W-1 multiples
// are produced by repeated doubling in XYZZ, parked in the output slots, and brought
// back to affine with ONE field inversion per point (Montgomery's trick across the
// windows of that point).
__global__ void __launch_bounds__(128) k_build_table(const G1Affine* __restrict__ in, G1Affine* __restrict__ out, size_t n,
                                                     size_t rows, int c, int nwin, G1Xyzz* __restrict__ park) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * rows) return;
    size_t row = t / n, k = t - row * n;
    G1Affine p = in[t];
    G1Affine* o = out + row * (size_t)nwin * n + k;  // window stride n
    o[0] = p;
    if (p.is_inf()) {
        for (int w = 1; w < nwin; w++) o[(size_t)w * n] = p;
        return;
    }
    G1Xyzz* pk = park + t;  // window stride n*rows
    const size_t ps = n * rows;
    G1Xyzz q = g1_dbl_affine(p);
    Fp pre[64];  // pre[w] = zzz_1 * ... * zzz_w (local memory; nwin <= 64)
    pre[0] = Fp::one();
    for (int w = 1; w < nwin; w++) {
        if (w > 1) q = g1_dbl(q);
        for (int d = 1; d < c; d++) q = g1_dbl(q);
        pk[(size_t)(w - 1) * ps] = q;
        pre[w] = fe_mul<FpParams>(pre[w - 1], q.zzz);
    }
    Fp inv = fe_inv<FpParams>(pre[nwin - 1]);  // 1 / (zzz_1 * ... * zzz_(nwin-1))
    for (int w = nwin - 1; w >= 1; w--) {
        G1Xyzz s = pk[(size_t)(w - 1) * ps];
        Fp izzz = fe_mul<FpParams>(inv, pre[w - 1]);
        o[(size_t)w * n] = g1_to_affine_with_inv(s, izzz);
        inv = fe_mul<FpParams>(inv, s.zzz);
    }
}

}  // namespace

int msm_pick_window(size_t n, bool tabulated) {
    int lg = 0;
    while (((size_t)1 << (lg + 1)) <= n) lg++;
    int c;
    if (tabulated) c = lg - 1;  // one bucket set shared by all windows
    else c = lg - 4;
    if (c < 4) c = 4;
    if (c > 20) c = 20;
    return c;
}

void msm_build_table(const G1Affine* d_in, size_t n, size_t rows, int c, int nwin, G1Affine* d_out, DevBuf& scratch,
                     cudaStream_t st) {
    if (!n || !rows) return;
    // chunk the rows so the XYZZ parking area stays below ~2 GiB
    size_t per_row = (size_t)(nwin - 1) * n * sizeof(G1Xyzz);
    size_t chunk = per_row ? ((size_t)2 << 30) / per_row : rows;
    if (chunk < 1) chunk = 1;
    if (chunk > rows) chunk = rows;
    G1Xyzz* park = scratch.as<G1Xyzz>((size_t)(nwin - 1) * n * chunk + 1);
    for (size_t r0 = 0; r0 < rows; r0 += chunk) {
        size_t rc = rows - r0 < chunk ? rows - r0 : chunk;
        size_t threads = n * rc;
        (++fb::g_launches, k_build_table<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(d_in + r0 * n, d_out + r0 * (size_t)nwin * n, n, rc, c,
                                                                        nwin, park));
    }
    FB_CUDA(cudaGetLastError());
}

void msm_run(const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, bool scalars_mont, G1Xyzz* d_out,
             MsmScratch& s, cudaStream_t st) {
    if (batch == 0) return;
    if (len == 0) {
        FB_CUDA(cudaMemsetAsync(d_out, 0, batch * sizeof(G1Xyzz), st));
        return;
    }
    if (len > bases.n) throw CudaError{"msm: more scalars than resident bases"};
    const size_t n = len * batch;
    const bool tab = bases.c > 0;
    const int c = tab ? bases.c : msm_pick_window(len, false);
    const int nwin = (256 + c - 1) / c;
    if (tab && nwin != bases.nwin) throw CudaError{"msm: table window count mismatch"};
    // with a table the entries index [w][k] over the FULL resident row length
    if ((uint64_t)(tab ? nwin : 1) * bases.n * batch >= (1ull << 31) || n >= (1ull << 32)) throw CudaError{"msm: index space exceeds 31 bits"};
    const uint32_t bpw = (1u << (c - 1)) + 1;         // bucket index = |digit|, 0 unused
    const uint32_t bpw_pad = (bpw + kSeg - 1) / kSeg * kSeg;
    const uint32_t per_out = tab ? 1u : (uint32_t)nwin;  // bucket sets per output
    const uint32_t nsets = per_out * (uint32_t)batch;
    const uint32_t nbuckets = nsets * bpw_pad;

    uint32_t* counts = s.counts.as<uint32_t>(nbuckets);
    uint32_t* offsets = s.offsets.as<uint32_t>(nbuckets);
    uint32_t* cursor = s.cursor.as<uint32_t>(nbuckets);
    uint32_t* sorted = s.sorted.as<uint32_t>((size_t)n * nwin);
    G1Xyzz* buckets = s.buckets.as<G1Xyzz>(nbuckets);

    prof_mark("msm.digits_count", st);
    FB_CUDA(cudaMemsetAsync(counts, 0, nbuckets * sizeof(uint32_t), st));
    DigitArgs a;
    a.scalars = d_scalars;
    a.n = (uint32_t)n;
    a.len = (uint32_t)len;
    a.stride = (uint32_t)bases.n;
    a.c = c;
    a.nwin = nwin;
    a.mont = scalars_mont ? 1 : 0;
    a.tabulated = tab ? 1 : 0;
    a.buckets_per_win = bpw_pad;
    a.counts = counts;
    a.sorted = sorted;
    DigitArgs a2 = a;
    const unsigned gb = (unsigned)((n + 255) / 256);
    (++fb::g_launches, k_digits<false><<<gb, 256, 0, st>>>(a));
    prof_mark("msm.scan", st);
    exclusive_scan(counts, offsets, nbuckets, s.block_sums, st);
    FB_CUDA(cudaMemcpyAsync(cursor, offsets, nbuckets * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    a2.counts = cursor;
    prof_mark("msm.scatter", st);
    (++fb::g_launches, k_digits<true><<<gb, 256, 0, st>>>(a2));
    prof_mark("msm.tasks", st);
    const uint64_t entries = (uint64_t)n * nwin;
    uint32_t cap = (uint32_t)(2 * entries / nbuckets);
    if (cap < 32) cap = 32;
    if (cap > 512) cap = 512;
    const uint32_t max_tasks = nbuckets + (uint32_t)(entries / cap) + 1;
    uint32_t* ntasks = s.order.as<uint32_t>((size_t)2 * nbuckets + 4);
    uint32_t* toff = ntasks + nbuckets;
    uint32_t* d_total = toff + nbuckets;
    uint2* tasks = s.misc.as<uint2>(max_tasks);
    G1Xyzz* partial = s.seg_t.as<G1Xyzz>(max_tasks);
    (++fb::g_launches, k_task_count<<<(nbuckets + 255) / 256, 256, 0, st>>>(counts, nbuckets, cap, ntasks));
    exclusive_scan(ntasks, toff, nbuckets, s.block_sums, st);
    (++fb::g_launches, k_task_fill<<<(nbuckets + 255) / 256, 256, 0, st>>>(counts, offsets, ntasks, toff, nbuckets, tasks, d_total));
    prof_mark("msm.accumulate", st);
    (++fb::g_launches, k_accumulate<<<(max_tasks + 127) / 128, 128, 0, st>>>(bases.pts, sorted, tasks, d_total, partial));
    prof_mark("msm.combine", st);
    (++fb::g_launches, k_combine<<<(nbuckets + 127) / 128, 128, 0, st>>>(partial, ntasks, toff, nbuckets, buckets));

    // bucket reduction
    prof_mark("msm.reduce", st);
    const G1Xyzz* X = buckets;
    const G1Xyzz* P = nullptr;
    uint32_t m = bpw_pad;
    G1Xyzz* seg = s.seg_a.as<G1Xyzz>((size_t)4 * nsets * ((bpw_pad >> kSegLog) + kSeg));
    size_t seg_off = 0;
    while (m > kSeg) {
        uint32_t mo = (m + kSeg - 1) >> kSegLog;
        G1Xyzz* Xo = seg + seg_off;
        G1Xyzz* Po = Xo + (size_t)nsets * mo;
        seg_off += (size_t)2 * nsets * mo;
        uint32_t threads = mo * nsets;
        (++fb::g_launches, k_seg_reduce<<<(threads + 127) / 128, 128, 0, st>>>(X, P, m, m, Xo, Po, mo, nsets));
        X = Xo;
        P = Po;
        m = mo;
    }
    prof_mark("msm.finish", st);
    (++fb::g_launches, k_finish<<<(unsigned)batch, 64, 0, st>>>(X, P, m, per_out, c, d_out));
    prof_mark("post", st);
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
