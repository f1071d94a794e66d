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
//   1. k_digits<0>: scalar -> W signed c-bit digits; histogram of bucket sizes
//   2. exclusive scan of the histogram                       (bucket offsets)
//   3. k_digits<1|2>: same digits again; scatter (point index | sign), or the point, into its
//      bucket's slice (a counting sort; order inside a bucket is irrelevant)
//   4. k_accumulate: the sorted entries cut into equal chunks, one per resident thread,
//      XYZZ += affine (8M+2S), points gathered by index from the resident table;
//      k_combine adds the partial sums of the buckets that straddle chunks
//   5. k_fold / k_bits / k_bit_combine: sum_i (i+1)*B_i as row + column sums of the
//      bucket matrix and bit-sliced weighted sums (plain sums only, short chains)
//   6. k_window_combine: per-window results (doubling ladder) when not tabulated
#include <stdlib.h>

#include "common.cuh"
#include "batchinv.cuh"

namespace fb {

namespace {

struct DigitArgs {
    const Fr* scalars;
    const Fr* scalars2;  // shared bases: the second scalar vector (row 1), else nullptr
    uint32_t row_bases;  // 1: row b has its own bases (b-th row of the table); 0: every row uses the same bases
    uint32_t n;       // total scalars = batch * len
    uint32_t len;     // scalars per row
    uint32_t stride;  // row length of the resident table (index = (row*nwin + w)*stride + col)
    int c, nwin;
    int mont;       // scalars are in Montgomery form
    int tabulated;  // single bucket set; entry = w*n + j
    uint32_t buckets_per_win;  // 2^(c-1) (index = |digit| - 1)
    uint32_t* counts;   // histogram (pass 0) / cursors (pass 1, 2)
    uint32_t* sorted;
    const uint32_t* off_out;  // MODE 2: offsets of the first round's output slots (ceil(count / 2) per bucket)
    uint32_t* where2;         // MODE 2: [slot][2] the slot's two table entries
};

__device__ __forceinline__ G1Affine ldg_affine(const G1Affine* p);
__device__ __forceinline__ void st_affine(G1Affine* dst, const G1Affine& r) {
    uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = make_uint4(r.x.v[4 * i], r.x.v[4 * i + 1], r.x.v[4 * i + 2], r.x.v[4 * i + 3]);
#pragma unroll
    for (int i = 0; i < 3; i++) d[3 + i] = make_uint4(r.y.v[4 * i], r.y.v[4 * i + 1], r.y.v[4 * i + 2], r.y.v[4 * i + 3]);
}

// MODE 0: histogram of bucket sizes; 1: scatter (table index | sign) into bucket order; 2: scatter straight into
// the slot map of the first batched-affine round (the k-th entry of a bucket is input k & 1 of the bucket's output
// slot k >> 1; cursors start at 0), which then needs neither the sorted list nor a slot-map kernel.
// (Scattering the points themselves, so that later kernels stream instead of gather, was
// measured: 1.5 ms of random 96-byte writes per 2^20 MSM, slower than the gathers it saves.)
template <int MODE>
__global__ void __launch_bounds__(256) k_digits(DigitArgs a) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const uint32_t row = j / a.len, col = j - row * a.len;
    Fr k = (a.scalars2 && row) ? a.scalars2[col] : a.scalars[j];
    if (a.mont) k = fe_from_mont<FrParams>(k);
    // k > (r-1)/2: use -(r - k).  The recoded scalar is < 2^254, so ceil(255/c) windows
    // hold every signed digit and the top window never carries out (without this a
    // window count that divides 255 exactly would put ~40% of the scalars' carries
    // into bucket 1 of an extra window).
    uint32_t sneg = 0;
    {
        Fr hr;
#pragma unroll
        for (int i = 0; i < 8; i++) hr.v[i] = (FrParams::mod(i) >> 1) | (i < 7 ? FrParams::mod(i + 1) << 31 : 0u);
        if (!fe_geq<FrParams>(hr.v, k.v)) {
            k = fe_neg<FrParams>(k);
            sneg = 1;
        }
    }
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
        uint32_t bucket = (a.tabulated ? row : row * a.nwin + (uint32_t)w) * a.buckets_per_win + mag - 1;
        if (MODE == 0) {
            atomicAdd(a.counts + bucket, 1u);
        } else if (MODE == 2) {
            const uint32_t k = atomicAdd(a.counts + bucket, 1u);  // cursors start at 0: the entry's rank in its bucket
            const uint32_t brow = row * a.row_bases;
            const uint32_t idx = a.tabulated ? (brow * a.nwin + (uint32_t)w) * a.stride + col : brow * a.stride + col;
            a.where2[2u * (a.off_out[bucket] + (k >> 1)) + (k & 1u)] = idx | ((neg ^ sneg) << 31);
        } else {
            uint32_t slot = atomicAdd(a.counts + bucket, 1u);
            const uint32_t brow = row * a.row_bases;
            uint32_t idx = a.tabulated ? (brow * a.nwin + (uint32_t)w) * a.stride + col : brow * a.stride + col;
            a.sorted[slot] = idx | ((neg ^ sneg) << 31);
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

// point i (bit 31 = negate) of an array with `stride` bytes per point
__device__ __forceinline__ const G1Affine* point_at(const char* pts, uint32_t stride, uint32_t i) {
    return reinterpret_cast<const G1Affine*>(pts + (size_t)(i & 0x7fffffffu) * stride);
}
__device__ __forceinline__ G1Affine load_point(const char* pts, uint32_t stride, uint32_t i) {
    G1Affine p = ldg_affine(point_at(pts, stride, i));
    if (i >> 31) p.y = fe_neg<FpParams>(p.y);
    return p;
}

// ---- batched-affine halving rounds.  A mixed XYZZ addition costs 10 Fp multiplies;
// an AFFINE addition costs 3 once 1/(x2 - x1) is known, and Montgomery's trick turns
// P inversions into one inversion plus 3(P-1) multiplies.  One round replaces every
// bucket's points (p0, p1, p2, ...) by (p0+p1, p2+p3, ...): 6 multiplies per addition.
// Each thread owns P consecutive OUTPUT slots of the round:
//   pass 1  walks its slots, forms the denominator of each pair (x2 - x1; 2*y for a
//           doubling; 1 when nothing is to be inverted) and parks the running product
//           in scratch (slot-major, so a warp writes consecutive elements);
//   warp    the 32 per-thread products are multiplied together with shuffles, ONE
//           lane inverts the product (binary GCD, no multiplies) and every lane
//           recovers the inverse of its own product from the prefix/suffix products;
//   pass 2  walks the slots backwards, peels off each pair's inverse and writes the sum.
// After a few rounds the survivors (a few per bucket) go through the XYZZ kernel below.
struct RoundArgs {
    const char* xs;           // GATHER, optional: x coordinates at 64-byte stride (pass 1 reads these instead of the lines)
    const char* pts;          // GATHER: the resident table; else the previous round's output (bucket order)
    uint32_t stride;          // bytes per input point (128 in a table, 96 otherwise)
    const uint32_t* off_out;  // [nbuckets + 1] bucket offsets of the output, off_out[b+1]-off_out[b] = ceil(count_in/2)
    uint32_t nbuckets;
    G1Affine* out;
    Fp* pref;                 // [step][thread] running products
    const uint32_t* where;    // streamed: [slot] first input index | (has a second point) << 31
    const uint2* where2;      // GATHER:   [slot] (table index | sign << 31) of both points, .y = kNone if single
    uint32_t p_max;           // steps per thread the scratch was sized for
    G1Affine* stage;          // GATHER: [step][thread][2] the two points of each slot, parked by pass 1
    Fp* inv;                  // [thread] inverse of the thread's running product (pass 1 -> pass 2)
    uint32_t waves;           // pass 2 runs `waves` times the threads of pass 1, each with 1/waves of the steps: a thread of
                              // pass 1 covers the slots of `waves` consecutive pass-2 warps as that many separate chains
};

constexpr uint32_t kMinSlots = 16;
constexpr uint32_t kNone = 0xffffffffu;

// Slot map: for every output slot o (one thread per slot, binary search of its bucket) the
// input(s) it adds.  GATHER (first round): the two table entries, so that the round kernel
// needs no dependent index load and can prefetch one step ahead.
template <bool GATHER>
__global__ void __launch_bounds__(256) k_slot_map(const uint32_t* __restrict__ off_in, const uint32_t* __restrict__ off_out,
                                                  uint32_t nbuckets, const uint32_t* __restrict__ sorted,
                                                  uint32_t* __restrict__ where, uint2* __restrict__ where2) {
    const uint32_t O = off_out[nbuckets];
    for (uint32_t o = blockIdx.x * blockDim.x + threadIdx.x; o < O; o += gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = nbuckets;  // off_out[lo] <= o < off_out[hi]
        while (hi - lo > 1) {
            uint32_t mid = (lo + hi) >> 1;
            if (off_out[mid] <= o) lo = mid;
            else hi = mid;
        }
        const uint32_t e = off_in[lo] + 2 * (o - off_out[lo]);
        const bool has2 = e + 1 < off_in[lo + 1];
        if (GATHER) where2[o] = make_uint2(sorted[e], has2 ? sorted[e + 1] : kNone);
        else where[o] = e | (has2 ? 0x80000000u : 0u);
    }
}

// streamed rounds: one warp per bucket writes the bucket's slots (no search): slot i of bucket b adds the
// inputs off_in[b] + 2i and + 2i + 1 (the last slot of an odd bucket has one input)
__global__ void __launch_bounds__(256) k_slot_fill(const uint32_t* __restrict__ off_in, const uint32_t* __restrict__ off_out,
                                                   uint32_t nbuckets, uint32_t* __restrict__ where) {
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < nbuckets; b += (gridDim.x * blockDim.x) >> 5) {
        const uint32_t i0 = off_in[b], cnt = off_in[b + 1] - i0, o0 = off_out[b];
        const uint32_t nout = (cnt + 1) >> 1;
        for (uint32_t i = lane; i < nout; i += 32u) where[o0 + i] = (i0 + 2u * i) | ((2u * i + 1u < cnt) ? 0x80000000u : 0u);
    }
}

// first round with the direct scatter (k_digits<2>): a bucket with an odd number of entries leaves the second
// input of its last slot unwritten - mark it (runs BEFORE the scatter, which never writes that word)
__global__ void __launch_bounds__(256) k_mark_single(const uint32_t* __restrict__ off_in, const uint32_t* __restrict__ off_out,
                                                     uint32_t nbuckets, uint32_t* __restrict__ where2) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    const uint32_t cnt = off_in[b + 1] - off_in[b];
    if (cnt & 1u) where2[2u * (off_out[b + 1] - 1u) + 1u] = kNone;
}

template <bool GATHER>
__device__ __forceinline__ uint2 slot_desc(const RoundArgs& a, uint32_t o) {
    if (GATHER) return a.where2[o];
    const uint32_t w = a.where[o], e = w & 0x7fffffffu;
    return make_uint2(e, (w >> 31) ? e + 1 : kNone);
}



__device__ __forceinline__ Fp ldg_fp(const Fp* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    Fp r;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint4 t = __ldg(q + i);
        r.v[4 * i] = t.x; r.v[4 * i + 1] = t.y; r.v[4 * i + 2] = t.z; r.v[4 * i + 3] = t.w;
    }
    return r;
}

// kind of a slot: 0 result = p1, 1 result = p2, 2 generic sum, 3 doubling, 4 identity
__device__ __forceinline__ int pair_kind(const G1Affine& p1, const G1Affine& p2, bool has2, Fp& den) {
    using F = FpParams;
    den = Fp::one();
    if (!has2 || p2.is_inf()) return 0;
    if (p1.is_inf()) return 1;
    Fp dx = fe_sub<F>(p2.x, p1.x);
    if (!dx.is_zero()) {
        den = dx;
        return 2;
    }
    if (p1.y == p2.y) {  // y != 0: the curve has no point of order 2
        den = fe_dbl<F>(p1.y);
        return 3;
    }
    return 4;
}



// A warp owns 32*P consecutive output slots; at step j lane l works on slot
// base + 32*j + l, so the slot map, the streamed points (2 x 96 B per lane, neighbouring
// lanes neighbouring pairs) and the results are read and written as contiguous runs.
struct WarpSlots {
    uint32_t base, cnt;  // first slot of the warp; steps of this lane
};
__device__ __forceinline__ bool warp_slots_of(const RoundArgs& a, uint32_t O, uint32_t nt, uint32_t t, WarpSlots& ws);
__device__ __forceinline__ bool warp_slots(const RoundArgs& a, uint32_t O, WarpSlots& ws) {
    return warp_slots_of(a, O, gridDim.x * blockDim.x, blockIdx.x * blockDim.x + threadIdx.x, ws);
}
// the slots of thread t of nt (the grid of pass 2)
__device__ __forceinline__ bool warp_slots_of(const RoundArgs& a, uint32_t O, uint32_t nt, uint32_t t, WarpSlots& ws) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t P = (O + nt - 1) / nt;
    if (P < kMinSlots) P = kMinSlots;
    if (P > a.p_max) P = a.p_max;  // cannot happen: p_max comes from an upper bound on O
    const uint64_t base64 = (uint64_t)(t >> 5) * 32u * P;
    if (base64 >= O) return false;  // whole warp idle
    // steps of this lane: slots base + 32*j + lane < end = min(O, base + 32*P).
    // NB: written as min(O - base, 32*P), ptxas 12.9 (sm_100a) fuses the subtraction and the min into
    // VIADDMNMX.U32 with the negation of `base` dropped (seen on B200 as out-of-range slots); the
    // add-then-min form below is what that instruction really computes.
    const uint64_t end64 = base64 + 32ull * P;
    const uint32_t end = end64 < O ? (uint32_t)end64 : O;
    ws.base = (uint32_t)base64;
    const uint32_t avail = end - ws.base;  // >= 1
    ws.cnt = avail > lane ? (avail - lane + 31u) / 32u : 0u;
    return true;
}

// Pass 1 (+ the warp's inversion).  Few live registers, so it runs at a higher occupancy than
// pass 2: it is the latency-bound half (GATHER: two random 128-byte lines per slot).  With
// GATHER it also parks both points (sign applied; identity when the slot has one input) in
// `stage`, [step][thread], so that pass 2 streams them instead of gathering them again.
// x coordinates of a slot's two inputs, issued early so that they are in flight while the
// previous slot's product is formed (a thread's running product is one dependent chain)
struct SlotX {
    uint2 d;
    Fp x1, x2;
};
__device__ __forceinline__ void slot_x_load(const char* pts, uint32_t stride, uint2 d, SlotX& s, const char* xs = nullptr) {
    s.d = d;
    if (d.y == kNone) return;
    if (xs) {
        s.x1 = ldg_fp(reinterpret_cast<const Fp*>(xs + (size_t)(d.x & 0x7fffffffu) * 64));
        s.x2 = ldg_fp(reinterpret_cast<const Fp*>(xs + (size_t)(d.y & 0x7fffffffu) * 64));
        return;
    }
    s.x1 = ldg_fp(&point_at(pts, stride, d.x)->x);
    s.x2 = ldg_fp(&point_at(pts, stride, d.y)->x);
}
__device__ __forceinline__ Fp slot_x_den(const char* pts, uint32_t stride, const SlotX& s) {
    using F = FpParams;
    if (s.d.y == kNone) return Fp::one();
    Fp dx = fe_sub<F>(s.x2, s.x1);
    if (s.x1.is_zero() || s.x2.is_zero() || dx.is_zero()) {  // identity, doubling or cancellation: needs y
        G1Affine p1 = load_point(pts, stride, s.d.x), p2 = load_point(pts, stride, s.d.y);
        Fp den;
        (void)pair_kind(p1, p2, true, den);
        return den;
    }
    return dx;
}

// MODE 0: streamed input (rounds >= 1); 1: gather + park the points for pass 2; 2: gather the
// x coordinates only (pass 2 gathers the points itself).  Launched on pass 2's grid (the running
// products are per thread), i.e. at 4 CTAs per SM: the loads of the next slot are issued before
// the current product to keep two slots in flight per thread.
template <int MODE, int NW>
__device__ __forceinline__ void round_pass1_body(const RoundArgs& a) {
    using F = FpParams;
    const uint32_t nt = gridDim.x * blockDim.x;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t O = a.off_out[a.nbuckets];
    if (MODE == 1) {  // (waves == 1 on this path)
        WarpSlots ws;
        const bool active = warp_slots(a, O, ws);
        if (!active) ws.base = ws.cnt = 0;  // an idle warp still meets the CTA's barriers below
        Fp run = Fp::one();
        for (uint32_t j = 0; j < ws.cnt; j++) {
            const uint2 d = slot_desc<true>(a, ws.base + 32u * j + lane);
            Fp den;
            G1Affine p1 = load_point(a.pts, a.stride, d.x);
            G1Affine p2 = d.y != kNone ? load_point(a.pts, a.stride, d.y) : G1Affine::inf();
            (void)pair_kind(p1, p2, true, den);
            G1Affine* st = a.stage + ((size_t)j * nt + t) * 2;
            st_affine(st, p1);
            st_affine(st + 1, p2);
            run = fe_mul<F>(run, den);
            a.pref[(size_t)j * nt + t] = run;
        }
        const Fp r = block_batch_inverse<F, NW>(run);
        if (active) a.inv[t] = r;
        return;
    }
    // The thread covers the slots of `waves` consecutive warps of pass 2's grid (ntv threads), as that many chains: chain k
    // restarts its running product (pass 2 peels every chain on its own), leaves its total T_k in inv[tv] and the product
    // G_k of the totals before it in the spare row of pref; after the CTA's one inversion of G = T_0 ... T_(w-1), a walk
    // back turns them into 1 / T_k (Montgomery's trick once more, two multiplications per chain).
    const uint32_t w = a.waves, ntv = nt * w;
    const uint32_t tv0 = ((t >> 5) * w) * 32u + lane;
    Fp* gpre = a.pref + (size_t)a.p_max * ntv;
    const char* xs = MODE == 2 ? a.xs : nullptr;
    Fp grand = Fp::one();
    uint32_t live = 0;  // bit k: chain k has steps
    for (uint32_t k = 0; k < w; k++) {
        const uint32_t tv = tv0 + 32u * k;
        WarpSlots ws;
        if (!warp_slots_of(a, O, ntv, tv, ws) || ws.cnt == 0) continue;
        live |= 1u << k;
        Fp run = Fp::one();
        SlotX nxt;
        slot_x_load(a.pts, a.stride, slot_desc<MODE != 0>(a, ws.base + lane), nxt, xs);
        for (uint32_t j = 0; j < ws.cnt; j++) {
            const SlotX cur = nxt;
            if (j + 1 < ws.cnt) slot_x_load(a.pts, a.stride, slot_desc<MODE != 0>(a, ws.base + 32u * (j + 1) + lane), nxt, xs);
            run = fe_mul<F>(run, slot_x_den(a.pts, a.stride, cur));
            a.pref[(size_t)j * ntv + tv] = run;
        }
        if (w > 1) {
            gpre[tv] = grand;
            a.inv[tv] = run;
            grand = fe_mul<F>(grand, run);
        } else {
            grand = run;
        }
    }
    // one inversion per CTA (4 warps): 176 -> ~80 us per launch against one per warp (batchinv.cuh)
    Fp r = block_batch_inverse<F, NW>(grand);
    if (w == 1) {
        if (live) a.inv[tv0] = r;
        return;
    }
    for (uint32_t k = w; k-- > 0;) {
        if (!(live >> k & 1u)) continue;
        const uint32_t tv = tv0 + 32u * k;
        const Fp tk = a.inv[tv];
        a.inv[tv] = fe_mul<F>(r, gpre[tv]);
        r = fe_mul<F>(r, tk);
    }
}

// MODE 0: streamed input; 1: the points parked by pass 1; 2: gather through the slot map
template <int MODE, int NW>
__device__ __forceinline__ void round_pass2_body(const RoundArgs& a) {
    using F = FpParams;
    const uint32_t nt = gridDim.x * blockDim.x;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    WarpSlots ws;
    if (!warp_slots(a, a.off_out[a.nbuckets], ws)) return;
    Fp inv = a.inv[t];
    for (uint32_t j = ws.cnt; j-- > 0;) {
        const uint32_t o = ws.base + 32u * j + lane;
        G1Affine p1, p2;
        if (MODE == 1) {
            const G1Affine* st = a.stage + ((size_t)j * nt + t) * 2;
            p1 = ldg_affine(st);
            p2 = ldg_affine(st + 1);
        } else {
            const uint2 d = slot_desc<MODE == 2>(a, o);
            p1 = load_point(a.pts, a.stride, d.x);
            p2 = d.y != kNone ? load_point(a.pts, a.stride, d.y) : G1Affine::inf();
        }
        Fp den;
        const int kind = pair_kind(p1, p2, true, den);
        Fp dinv = inv;
        if (j) dinv = fe_mul<F>(inv, a.pref[(size_t)(j - 1) * nt + t]);
        inv = fe_mul<F>(inv, den);
        G1Affine r;
        if (kind == 0) r = p1;
        else if (kind == 1) r = p2;
        else if (kind == 4) r = G1Affine::inf();
        else {
            Fp num;
            if (kind == 2) num = fe_sub<F>(p2.y, p1.y);
            else {
                Fp xx = fe_sqr<F>(p1.x);
                num = fe_add<F>(fe_dbl<F>(xx), xx);
            }
            Fp lam = fe_mul<F>(num, dinv);
            Fp x3 = fe_sub<F>(fe_sub<F>(fe_sqr<F>(lam), p1.x), p2.x);
            Fp y3 = fe_sub<F>(fe_mul<F>(lam, fe_sub<F>(p1.x, x3)), p1.y);
            r.x = x3;
            r.y = y3;
        }
        st_affine(a.out + o, r);
    }
}

// CTA shapes of the round kernels: 4 warps x 4 CTAs/SM (124 registers; default) or 6 warps x 3 CTAs/SM at 112 registers
// (18 instead of 16 warps per SM; FB_MSM_CTA=6)
template <int MODE>
__global__ void __launch_bounds__(128, 4) k_round_pass1(RoundArgs a) { round_pass1_body<MODE, 4>(a); }
template <int MODE>
__global__ void __launch_bounds__(128, 4) k_round_pass2(RoundArgs a) { round_pass2_body<MODE, 4>(a); }
template <int MODE>
__global__ void __maxnreg__(112) k_round_pass1_w6(RoundArgs a) { round_pass1_body<MODE, 6>(a); }
template <int MODE>
__global__ void __maxnreg__(112) k_round_pass2_w6(RoundArgs a) { round_pass2_body<MODE, 6>(a); }

// work counters for the bench's real-multiply accounting (launched only while profiling is on)
__global__ void k_stat_add(unsigned long long* stats, int slot, const uint32_t* __restrict__ val, unsigned long long add) {
    stats[slot] += (val ? (unsigned long long)*val : 0ull) + add;
}

// counts of the next round: ceil(count/2) per bucket (entry nbuckets = 0 closes the scan)
__global__ void __launch_bounds__(256) k_halve_counts(const uint32_t* __restrict__ off_in, uint32_t nbuckets,
                                                      uint32_t* __restrict__ cnt_out) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nbuckets) return;
    cnt_out[b] = b < nbuckets ? (off_in[b + 1] - off_in[b] + 1) >> 1 : 0u;
}

// ---- balanced bucket accumulation.  The sorted entry list (E entries, grouped by
// bucket) is cut into NT equal chunks of K = ceil(E / NT) entries, NT = one wave of
// resident threads (a multiple of the SM count), whatever the bucket boundaries are:
// every thread does exactly K mixed additions, so there is no tail and no lane
// idles while its neighbours walk a longer bucket (the top window of a 255-bit
// scalar, or structured inputs such as all-equal scalars, put thousands of points
// into a few buckets).  When the bucket changes inside a chunk the running sum is
// flushed to partial[bucket + chunk] - (bucket, chunk) pairs are visited in
// lexicographic order, so that slot is unique - and k_combine adds the partials
// of each bucket (typically one or two).
constexpr uint32_t kMinChunk = 8;

__device__ __forceinline__ uint32_t chunk_len(uint32_t E, uint32_t nt) {
    uint32_t k = (E + nt - 1) / nt;
    return k < kMinChunk ? kMinChunk : k;
}

template <int MINB, bool DIRECT>  // DIRECT: pts[e] is entry e itself (output of the affine rounds), no gather, no sign
__global__ void __launch_bounds__(128, MINB) k_accumulate(const char* __restrict__ pts, uint32_t stride,
                                                          const uint32_t* __restrict__ sorted,
                                                          const uint32_t* __restrict__ offsets, uint32_t nbuckets,
                                                          G1Xyzz* __restrict__ partial) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t E = offsets[nbuckets];
    const uint32_t K = chunk_len(E, gridDim.x * blockDim.x);
    uint64_t e0 = (uint64_t)t * K;
    if (e0 >= E) return;
    uint32_t e = (uint32_t)e0;
    const uint32_t end = E - e < K ? E : e + K;
    // bucket of the first entry: largest b with offsets[b] <= e
    uint32_t lo = 0, hi = nbuckets;  // invariant: offsets[lo] <= e < offsets[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (offsets[mid] <= e) lo = mid;
        else hi = mid;
    }
    uint32_t cb = lo;
    uint32_t next = offsets[cb + 1];
    G1Xyzz acc = G1Xyzz::inf();
    for (; e < end; e++) {
        if (e == next) {
            partial[cb + t] = acc;
            acc = G1Xyzz::inf();
            do {
                cb++;
                next = offsets[cb + 1];
            } while (next <= e);
        }
        G1Affine p;
        if (DIRECT) p = ldg_affine(point_at(pts, stride, e));
        else p = load_point(pts, stride, sorted[e]);
        g1_add_mixed(acc, p);
    }
    partial[cb + t] = acc;
}

// bucket b = sum of its partials: chunks first..last that its entry range meets.
// Buckets met by more than kLongPartials chunks (skewed digits: equal or tiny scalars)
// are queued for k_combine_long, where a whole CTA adds the partials as a tree.
constexpr uint32_t kLongPartials = 6;
constexpr int kLongThreads = 128;

__global__ void __launch_bounds__(128) k_combine(const G1Xyzz* __restrict__ partial, const uint32_t* __restrict__ offsets,
                                                 uint32_t nbuckets, uint32_t nt, G1Xyzz* __restrict__ buckets,
                                                 uint32_t* __restrict__ long_list /* [0] = count */) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    const uint32_t K = chunk_len(offsets[nbuckets], nt);
    uint32_t lo = offsets[b], hi = offsets[b + 1];
    G1Xyzz acc = G1Xyzz::inf();
    if (hi > lo) {
        uint32_t t0 = lo / K, t1 = (hi - 1) / K;
        if (t1 - t0 >= kLongPartials) {
            long_list[1 + atomicAdd(long_list, 1u)] = b;
            return;
        }
        acc = partial[b + t0];
        for (uint32_t t = t0 + 1; t <= t1; t++) g1_add(acc, partial[b + t]);
    }
    buckets[b] = acc;
}

__global__ void __launch_bounds__(kLongThreads) k_combine_long(const G1Xyzz* __restrict__ partial,
                                                               const uint32_t* __restrict__ offsets, uint32_t nbuckets,
                                                               uint32_t nt, G1Xyzz* __restrict__ buckets,
                                                               const uint32_t* __restrict__ long_list) {
    __shared__ G1Xyzz sm[kLongThreads];
    const uint32_t K = chunk_len(offsets[nbuckets], nt);
    const uint32_t nlong = long_list[0];
    for (uint32_t i = blockIdx.x; i < nlong; i += gridDim.x) {
        const uint32_t b = long_list[1 + i];
        const uint32_t t0 = offsets[b] / K, t1 = (offsets[b + 1] - 1) / K;
        G1Xyzz acc = G1Xyzz::inf();
        for (uint32_t t = t0 + threadIdx.x; t <= t1; t += kLongThreads) g1_add(acc, partial[b + t]);
        __syncthreads();  // previous round's readers of sm[] are done
        sm[threadIdx.x] = acc;
        for (uint32_t s = kLongThreads / 2; s > 0; s >>= 1) {
            __syncthreads();
            if (threadIdx.x < s) {
                G1Xyzz x = sm[threadIdx.x];
                g1_add(x, sm[threadIdx.x + s]);
                sm[threadIdx.x] = x;
            }
        }
        if (threadIdx.x == 0) buckets[b] = sm[0];
    }
}

// ---- bucket reduction  S = sum_i (i+1) * X[i]  over 2^k buckets per set (bucket
// index = |digit| - 1), organised so that every step is a PLAIN sum with a short
// dependent chain (an XYZZ addition is ~14 dependent Fp multiplies, ~14 us on one
// thread, so running-sum schemes are latency-bound):
//   * k > kDirectBits: write i = hi*2^l + lo.  With row sums R[hi] and column sums
//     C[lo] of the 2^h x 2^l bucket matrix,
//         S = sum_lo lo*C[lo] + 2^l * sum_hi hi*R[hi] + sum_hi R[hi]
//     (2 additions per bucket, done by k_fold passes of <= 3 additions per thread);
//   * a weighted sum over n <= 2^kDirectBits.. elements is taken bit by bit:
//         sum_i i*Y[i] = sum_j 2^j * (sum of the Y[i] whose index has bit j set)
//     one block per (set, bit): strided partial sums + a shared-memory tree (k_bits);
//   * k_bit_combine scales the per-bit sums by 2^j (j doublings, in parallel) and adds
//     them with a tree.
constexpr int kDirectBits = 8;

struct FoldDir {
    const G1Xyzz* in;
    G1Xyzz* out;
    uint32_t n_out;  // outputs of this pass (0: direction idle)
    uint32_t G;      // inputs per output
    uint32_t L;      // stride between the inputs of one output (1: contiguous)
};
struct FoldArgs {
    FoldDir d[2];  // [0] row sums, [1] column sums
};

// out[o] = sum_{g<G} in[(o / L)*G*L + (o % L) + g*L]
__global__ void __launch_bounds__(128, 3) k_fold(FoldArgs a) {
    const FoldDir& d = a.d[blockIdx.y];
    uint32_t o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= d.n_out) return;
    size_t base = (size_t)(o / d.L) * d.G * d.L + (o % d.L);
    G1Xyzz acc = d.in[base];
    for (uint32_t g = 1; g < d.G; g++) g1_add(acc, d.in[base + (size_t)g * d.L]);
    d.out[o] = acc;
}

struct BitsArgs {
    const G1Xyzz* X;  // direct: [set][2^k]
    const G1Xyzz* C;  // split: [set][2^l] column sums
    const G1Xyzz* R;  // split: [set][2^h] row sums
    int k, l, h;      // l == 0: direct
    G1Xyzz* Q;        // [set][k+1]: per-bit sums by overall weight, Q[k] = plain total
    uint32_t njobs;   // sets * (k + 1)
};

constexpr int kBitsThreads = 128;

// One job per (set, bit), G threads per job and kBitsThreads / G jobs per CTA: G = 32 (64) when a job has at most
// 32 (64) addends (the row MSMs of the bivariate path: 2^10 sets of 2^11 buckets, 12288 jobs of 16-32 addends,
// where a 128-thread CTA per job left three warps of four idle at every barrier), else the whole CTA.
template <int G>
__global__ void __launch_bounds__(kBitsThreads) k_bits(BitsArgs a) {
    __shared__ G1Xyzz sm[kBitsThreads];
    const uint32_t job = blockIdx.x * (kBitsThreads / G) + threadIdx.x / G;
    const uint32_t tj = threadIdx.x % G;
    const bool live = job < a.njobs;
    const uint32_t set = job / (a.k + 1), q = job % (a.k + 1);
    const G1Xyzz* src;
    uint32_t n;
    int bit;  // -1: plain total
    if (a.l == 0) {
        src = a.X + ((size_t)set << a.k);
        n = 1u << a.k;
        bit = (int)q < a.k ? (int)q : -1;
    } else if ((int)q < a.l) {
        src = a.C + ((size_t)set << a.l);
        n = 1u << a.l;
        bit = (int)q;
    } else {
        src = a.R + ((size_t)set << a.h);
        n = 1u << a.h;
        bit = (int)q < a.k ? (int)q - a.l : -1;
    }
    const uint32_t cnt = live ? (bit < 0 ? n : n >> 1) : 0u;
    G1Xyzz acc = G1Xyzz::inf();
    for (uint32_t i = tj; i < cnt; i += G) {
        uint32_t idx = bit < 0 ? i : (((i >> bit) << (bit + 1)) | (1u << bit) | (i & ((1u << bit) - 1u)));
        g1_add(acc, src[idx]);
    }
    sm[threadIdx.x] = acc;
    for (uint32_t s = G / 2; s > 0; s >>= 1) {
        __syncthreads();
        if (tj < s && tj + s < cnt) {
            G1Xyzz x = sm[threadIdx.x];
            g1_add(x, sm[threadIdx.x + s]);
            sm[threadIdx.x] = x;
        }
    }
    if (live && tj == 0) a.Q[job] = sm[threadIdx.x];
}

// S[set] = sum_{j<k} 2^j * Q[set][j] + Q[set][k].  One block of 32 threads per set.
__global__ void __launch_bounds__(32) k_bit_combine(const G1Xyzz* __restrict__ Q, int k, G1Xyzz* __restrict__ S) {
    __shared__ G1Xyzz sm[32];
    const uint32_t set = blockIdx.x, j = threadIdx.x;
    G1Xyzz acc = G1Xyzz::inf();
    if ((int)j <= k) {
        acc = Q[(size_t)set * (k + 1) + j];
        if ((int)j < k)
            for (uint32_t d = 0; d < j; d++) acc = g1_dbl(acc);
    }
    sm[j] = acc;
    for (uint32_t s = 16; s > 0; s >>= 1) {
        __syncwarp();
        if (j < s) {
            G1Xyzz x = sm[j];
            g1_add(x, sm[j + s]);
            sm[j] = x;
        }
    }
    if (j == 0) S[set] = sm[0];
}

// untabulated bases: out = sum_w 2^(c*w) * S[w] (thread w doubles c*w times, thread 0 adds)
__global__ void __launch_bounds__(64) k_window_combine(const G1Xyzz* __restrict__ S, uint32_t nwin, int c, G1Xyzz* __restrict__ out) {
    __shared__ G1Xyzz part[64];
    uint32_t w = threadIdx.x;
    S += (size_t)blockIdx.x * nwin;  // one block per output (batch row)
    if (w < nwin) {
        G1Xyzz acc = S[w];
        for (uint32_t d = 0; d < (uint32_t)c * w; d++) acc = g1_dbl(acc);
        part[w] = acc;
    }
    __syncthreads();
    if (w == 0) {
        G1Xyzz total = part[0];
        for (uint32_t i = 1; i < nwin; i++) g1_add(total, part[i]);
        out[blockIdx.x] = total;
    }
}

// table[row][w][k] = 2^(c*w) * P[row][k].  One thread per point: the W-1 multiples
// are produced by repeated doubling in XYZZ, parked in the output slots, and brought
// back to affine with ONE field inversion per point (Montgomery's trick across the
// windows of that point).
struct alignas(128) G1TableEntry {  // kTableStride bytes: x | y | 32 bytes of padding
    G1Affine p;
    uint32_t pad[8];
};
static_assert(sizeof(G1TableEntry) == kTableStride, "table entries are one 128-byte line");

__global__ void __launch_bounds__(128) k_build_table(const G1Affine* __restrict__ in, G1TableEntry* __restrict__ out, size_t n,
                                                     size_t rows, int c, int nwin, G1Xyzz* __restrict__ park) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * rows) return;
    size_t row = t / n, k = t - row * n;
    G1Affine p = in[t];
    G1TableEntry* o = out + row * (size_t)nwin * n + k;  // window stride n
    o[0].p = p;
    if (p.is_inf()) {
        for (int w = 1; w < nwin; w++) o[(size_t)w * n].p = p;
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
        o[(size_t)w * n].p = g1_to_affine_with_inv(s, izzz);
        inv = fe_mul<FpParams>(inv, s.zzz);
    }
}

// out[b] = sum_q parts[b * P + q]  (b < nout outputs, P partial sums each)
__global__ void __launch_bounds__(32) k_sum_parts(const G1Xyzz* __restrict__ parts, int P, int nout, G1Xyzz* __restrict__ out) {
    const int b = threadIdx.x;
    if (b >= nout) return;
    G1Xyzz acc = parts[b * P];
    for (int q = 1; q < P; q++) g1_add(acc, parts[b * P + q]);
    out[b] = acc;
}

}  // namespace

__global__ void __launch_bounds__(256) k_build_xtable(const char* __restrict__ table, size_t entries, uint4* __restrict__ xs) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // one 16-byte chunk per thread, 4 per entry
    if (i >= entries * 4) return;
    const size_t e = i >> 2, c = i & 3;
    xs[i] = c < 3 ? reinterpret_cast<const uint4*>(table + e * kTableStride)[c] : make_uint4(0, 0, 0, 0);
}
// The x coordinates of a window table once more, alone, in 64-byte slots.  The first round's pass 1 needs x only; out of the
// 128-byte table lines the whole line travels (2.07 GB per 2^20-point MSM), out of this copy one 64-byte DRAM atom does:
// pass 1 2.71 -> 2.46 ms per step, step 10.89 -> 10.68 ms (profiles/r2_v6_ab_xtable.txt), for +50 % table memory.
// Tables above 4 GiB (the 2^24-point MSM: 25.8 GB; the 2^10 rows: 11.8 GB) measured no gain (pass 1 16.1 ms / 11.2 ms either
// way), so the copy is built for tables of <= 2^25 entries only.  FB_MSM_XTAB=0 turns it off, =2 forces it; nullptr when off
// or when the memory is not there (the MSM then reads the lines).
const char* msm_alloc_xtable(const void* d_table, size_t entries, cudaStream_t st) {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("FB_MSM_XTAB");
        on = e ? atoi(e) : 1;
    }
    if (!on || !entries || (on == 1 && entries > ((size_t)1 << 25))) return nullptr;
    void* xs = nullptr;
    if (cudaMalloc(&xs, entries * 64) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    (++fb::g_launches, k_build_xtable<<<(unsigned)((entries * 4 + 255) / 256), 256, 0, st>>>((const char*)d_table, entries, (uint4*)xs));
    FB_CUDA(cudaGetLastError());
    return (const char*)xs;
}

int msm_num_windows(int c) { return (255 + c - 1) / c; }  // scalars are recoded below 2^254 (k_digits)

// Window width by a cost model in Fp multiplications: n*W mixed additions (10 each) plus
// 60 per bucket (2 full additions of 14, doubled: the reduction kernels are short and
// latency-bound; calibrated on the measured window sweep, profiles/r1_v2_window_sweep.jsonl);
// without a table every window has its own bucket set and a doubling ladder joins them.
int msm_pick_window(size_t n, bool tabulated) {
    int best = 4;
    double best_cost = 1e300;
    for (int c = 4; c <= 22; c++) {
        const double W = (double)msm_num_windows(c);
        const double buckets = (double)((size_t)1 << (c - 1)) * (tabulated ? 1.0 : W);
        double cost = (double)n * W * 10.0 + buckets * 100.0;  // a bucket costs two full XYZZ additions (measured, profiles/)
        if (!tabulated) cost += 255.0 * 9.0;
        if (cost < best_cost) {
            best_cost = cost;
            best = c;
        }
    }
    return best;
}

void msm_build_table(const G1Affine* d_in, size_t n, size_t rows, int c, int nwin, void* d_out_v, DevBuf& scratch,
                     cudaStream_t st) {
    if (!n || !rows) return;
    G1TableEntry* d_out = static_cast<G1TableEntry*>(d_out_v);
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
             MsmScratch& s, cudaStream_t st, const Fr* d_scalars2) {
    if (batch == 0) return;
    if (d_scalars2 && batch != 2) throw CudaError{"msm: a second scalar vector means batch = 2"};
    const bool shared = d_scalars2 != nullptr;
    if (len == 0) {
        FB_CUDA(cudaMemsetAsync(d_out, 0, batch * sizeof(G1Xyzz), st));
        return;
    }
    if (len > bases.n) throw CudaError{"msm: more scalars than resident bases"};
    const size_t n = len * batch;
    const bool tab = bases.c > 0;
    const int c = tab ? bases.c : msm_pick_window(len, false);
    const int nwin = msm_num_windows(c);
    if (tab && nwin != bases.nwin) throw CudaError{"msm: table window count mismatch"};
    // with a table the entries index [w][k] over the FULL resident row length
    if ((uint64_t)(tab ? nwin : 1) * bases.n * (shared ? 1 : batch) >= (1ull << 31) || n >= (1ull << 32)) throw CudaError{"msm: index space exceeds 31 bits"};
    const int k = c - 1;                                // buckets per set = 2^k, index = |digit| - 1
    const uint32_t bpw = 1u << k;
    const uint32_t per_out = tab ? 1u : (uint32_t)nwin;  // bucket sets per output
    const uint32_t nsets = per_out * (uint32_t)batch;
    if ((uint64_t)nsets * bpw >= (1ull << 31)) throw CudaError{"msm: too many buckets"};
    const uint32_t nbuckets = nsets * bpw;

    uint32_t* counts = s.counts.as<uint32_t>(nbuckets + 1);
    uint32_t* offsets = s.offsets.as<uint32_t>(nbuckets + 1);  // offsets[nbuckets] = number of entries
    uint32_t* cursor = s.cursor.as<uint32_t>(nbuckets + 1);
    G1Xyzz* buckets = s.buckets.as<G1Xyzz>(nbuckets);

    // number of batched-affine rounds (0: gather + XYZZ only)
    const uint64_t entries = (uint64_t)n * nwin;
    static int rounds_env = -2, round_blocks_per_sm = 0, round_warps = 4;
    if (rounds_env == -2) {
        const char* e = getenv("FB_MSM_ROUNDS");  // tuning knob: number of affine rounds (default: by size)
        const char* w = getenv("FB_MSM_CTA");     // tuning knob: warps per CTA of the round kernels (4 or 6)
        round_warps = (w && atoi(w) == 6) ? 6 : 4;
        if (round_warps == 6) FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&round_blocks_per_sm, k_round_pass2_w6<1>, 192, 0));
        else FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&round_blocks_per_sm, k_round_pass2<1>, 128, 0));
        if (round_blocks_per_sm < 1) round_blocks_per_sm = 1;
        rounds_env = e ? atoi(e) : -1;
    }
    const uint32_t rthreads = 32u * (uint32_t)round_warps;
    // Measured on B200 (profiles/): three rounds ahead of the XYZZ kernel are ~8% faster than the
    // XYZZ kernel alone at 2^20 points (pass 2 runs at the multiplier's rate, pass 1 of the first
    // round is bound by the gather); small inputs do not amortise the per-round inversion.
    // Jobs of >= 64 M entries (the 2^24-point MSM, the 2^10 batched rows) amortise two more rounds when the buckets
    // are deep enough: 2^24 points 217.5 -> 222.1 Mpts/s, bivariate rows 69.1 -> 67.8 ms; at 2^20 points (16 M
    // entries) a fourth round costs 0.1 ms (profiles/r1_v7_sweeps.txt).  Round 2 (profiles/r2_v7_rounds_sweep.txt): at 2^24 points
    // (96 entries per bucket) 5 / 6 / 7 / 8 rounds give 232.1 / 231.3 / 235.8 / 233.8 Mpts/s, three runs each.
    int rounds = 0;
    if (entries >= (4u << 20) && entries / nbuckets >= 8) {
        rounds = (s.overlapped && entries >= (8u << 20)) ? 4 : 3;
        if (entries >= (64u << 20) && entries / nbuckets >= 64) rounds = 7;  // one survivor per bucket: the XYZZ pass and its combine all but vanish
        else if (entries >= (64u << 20) && entries / nbuckets >= 32) rounds = 4;
    }
    if (rounds_env >= 0) rounds = rounds_env;
    if (rounds > 8) rounds = 8;
    uint32_t* sorted = s.sorted.as<uint32_t>(entries);

    prof_mark("msm.digits_count", st);
    FB_CUDA(cudaMemsetAsync(counts, 0, (nbuckets + 1) * sizeof(uint32_t), st));
    DigitArgs a;
    a.scalars = d_scalars;
    a.scalars2 = d_scalars2;
    a.row_bases = shared ? 0u : 1u;
    a.n = (uint32_t)n;
    a.len = (uint32_t)len;
    a.stride = (uint32_t)bases.n;
    a.c = c;
    a.nwin = nwin;
    a.mont = scalars_mont ? 1 : 0;
    a.tabulated = tab ? 1 : 0;
    a.buckets_per_win = bpw;
    a.counts = counts;
    a.sorted = sorted;
    DigitArgs a2 = a;
    const unsigned gb = (unsigned)((n + 255) / 256);
    (++fb::g_launches, k_digits<0><<<gb, 256, 0, st>>>(a));
    prof_mark("msm.scan", st);
    exclusive_scan(counts, offsets, nbuckets + 1, s.block_sums, st);
    a2.counts = cursor;
    // With batched-affine rounds the scatter writes the first round's slot map directly (k_digits<2>): the output
    // offsets of that round (ceil(count / 2) per bucket, scanned) are formed first, into `counts` (the histogram has
    // been consumed by the scan).  FB_MSM_DIRECT=0 keeps the sorted list + slot-map kernel.
    static int direct_env = -1;
    if (direct_env < 0) {
        const char* e = getenv("FB_MSM_DIRECT");
        direct_env = e ? atoi(e) : 1;
    }
    const bool direct = rounds > 0 && direct_env != 0;
    uint32_t* wbuf0 = nullptr;
    if (direct) {
        const uint64_t ub_out0 = (entries + nbuckets) / 2 + 1;
        uint32_t* cnt_tmp = s.order.as<uint32_t>(nbuckets + 1);
        (++fb::g_launches, k_halve_counts<<<(nbuckets + 256) / 256, 256, 0, st>>>(offsets, nbuckets, cnt_tmp));
        exclusive_scan(cnt_tmp, counts, nbuckets + 1, s.block_sums, st);
        wbuf0 = s.where.as<uint32_t>(2 * ub_out0);
        (++fb::g_launches, k_mark_single<<<(nbuckets + 255) / 256, 256, 0, st>>>(offsets, counts, nbuckets, wbuf0));
        FB_CUDA(cudaMemsetAsync(cursor, 0, nbuckets * sizeof(uint32_t), st));
        a2.off_out = counts;
        a2.where2 = wbuf0;
        prof_mark("msm.scatter", st);
        (++fb::g_launches, k_digits<2><<<gb, 256, 0, st>>>(a2));
    } else {
        FB_CUDA(cudaMemcpyAsync(cursor, offsets, nbuckets * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        prof_mark("msm.scatter", st);
        (++fb::g_launches, k_digits<1><<<gb, 256, 0, st>>>(a2));
    }
    const bool stats_on = g_prof && g_prof->on;
    unsigned long long* stats = s.stats.as<unsigned long long>(8);
    if (stats_on) {
        if (!s.stats_zeroed) {
            FB_CUDA(cudaMemsetAsync(stats, 0, 8 * sizeof(unsigned long long), st));
            s.stats_zeroed = true;
        }
        k_stat_add<<<1, 1, 0, st>>>(stats, 0, nullptr, 1);
        k_stat_add<<<1, 1, 0, st>>>(stats, 1, offsets + nbuckets, 0);
        k_stat_add<<<1, 1, 0, st>>>(stats, 4, nullptr, nbuckets);
    }

    // batched-affine halving rounds, then XYZZ accumulation of the survivors
    const uint32_t* off_cur = offsets;
    const char* pts_cur = static_cast<const char*>(bases.pts);  // round 0 gathers from the table through `sorted`
    uint32_t stride_cur = bases.stride();
    uint64_t ub = entries;                 // upper bound on the number of live points
    for (int r = 0; r < rounds; r++) {
        prof_mark("msm.round_map", st);
        const uint64_t ub_out = (ub + nbuckets) / 2 + 1;
        uint32_t* off_out = (r & 1) ? cursor : counts;
        if (!(direct && r == 0)) {  // (the direct scatter formed round 0's offsets already)
            uint32_t* cnt_tmp = s.order.as<uint32_t>(nbuckets + 1);
            (++fb::g_launches, k_halve_counts<<<(nbuckets + 256) / 256, 256, 0, st>>>(off_cur, nbuckets, cnt_tmp));
            exclusive_scan(cnt_tmp, off_out, nbuckets + 1, s.block_sums, st);
        }
        // Pass 1 runs one wave of CTAs (each pays one inversion on one lane while its other warps wait); pass 2 runs `waves`
        // times as many CTAs with 1 / waves of the steps each: one wave gives every warp the same number of steps, but the
        // warps of a scheduler do not advance evenly and the last ones run alone (ncu: 12.8 of 16 warps resident on
        // average); with several waves a finished CTA is replaced.  Every chain costs pass 1 three more multiplications
        // and a restart of its load pipeline, so chains keep a minimum length: up to 6 waves (profiles/r2_v8_ab_waves.txt: 2^24-point
        // MSM 235.9 -> 243.9 Mpts/s at 6, the 2^10 x 2^12 rows 58.4 -> 57.7 ms, the headline step 12.46 -> 12.34 ms on one
        // stream and unchanged when its two MSMs overlap).  FB_MSM_WAVES=k fixes the number (1: the single-wave form).
        static const uint32_t waves_env = [] {
            const char* e = getenv("FB_MSM_WAVES");
            int v = e ? atoi(e) : 0;
            return (uint32_t)(v < 0 ? 0 : (v > 16 ? 16 : v));
        }();
        static const int regather = [] {
            // First round: pass 1 reads the x coordinates only (two sectors per point) and pass 2, which is
            // multiplier-bound, hides a second gather of the full points; FB_MSM_REGATHER=0 parks the
            // points in pass 1 instead (measured 4% slower at 2^20; L2 prefetches made both slower).
            const char* e = getenv("FB_MSM_REGATHER");
            return e ? atoi(e) : 1;
        }();
        uint32_t blocks = (uint32_t)(kNumSMs * round_blocks_per_sm);
        uint32_t waves = waves_env;
        {
            uint64_t need = (ub_out / kMinSlots + rthreads - 1) / rthreads + 1;  // CTAs of pass 2 with >= kMinSlots steps per thread
            if (need < blocks) blocks = (uint32_t)need;
            if (!waves) {  // by size: chains of >= 24 steps in the first round, whose pass 1 waits for its gather anyway, of >= 64
                           // in the streamed rounds, where the chains' extra multiplications show (ncu, profiles/r2_v9_*: two
                           // chains of 24 cost the second round's pass 1 more than its pass 2 gained)
                const uint64_t steps = ub_out / ((uint64_t)blocks * rthreads), per = r == 0 ? 24 : 64;
                waves = (uint32_t)(steps / per < 1 ? 1 : (steps / per > 6 ? 6 : steps / per));
            }
            if (r == 0 && !regather) waves = 1;
            while (waves > 1 && (uint64_t)blocks * waves > need) waves--;
        }
        const uint32_t blocks2 = blocks * waves;
        const uint32_t rnt = blocks2 * rthreads;  // threads of pass 2
        RoundArgs ra;
        ra.pts = pts_cur;
        ra.xs = (r == 0 && tab) ? bases.xs : nullptr;
        ra.stride = stride_cur;
        ra.off_out = off_out;
        ra.nbuckets = nbuckets;
        ra.out = ((r & 1) ? s.pts_b : s.pts_a).as<G1Affine>(ub_out);
        uint64_t p_max = (ub_out + rnt - 1) / rnt;
        if (p_max < kMinSlots) p_max = kMinSlots;
        ra.p_max = (uint32_t)p_max;
        ra.pref = s.pref.as<Fp>((p_max + 1) * rnt);  // + one row: the products of the chain totals (pass 1)
        ra.waves = waves;
        uint32_t* wbuf = s.where.as<uint32_t>(2 * ub_out);
        ra.where = wbuf;
        ra.where2 = reinterpret_cast<const uint2*>(wbuf);
        uint64_t mb = (ub_out + 255) / 256;
        if (mb > 4u * kNumSMs * 8u) mb = 4u * kNumSMs * 8u;
        ra.inv = s.inv.as<Fp>(rnt);
        ra.stage = nullptr;
        if (r == 0) {
            if (!direct) (++fb::g_launches, k_slot_map<true><<<(unsigned)mb, 256, 0, st>>>(off_cur, off_out, nbuckets, sorted, wbuf, reinterpret_cast<uint2*>(wbuf)));
            prof_mark("msm.round_pass1", st);
            if (regather) {
                (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass1_w6<2><<<blocks, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass1<2><<<blocks, 128, 0, st>>>(ra)));
                prof_mark("msm.round_pass2", st);
                (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass2_w6<2><<<blocks2, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass2<2><<<blocks2, 128, 0, st>>>(ra)));
            } else {
                ra.stage = s.stage.as<G1Affine>(2 * p_max * rnt);
                (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass1_w6<1><<<blocks, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass1<1><<<blocks, 128, 0, st>>>(ra)));
                prof_mark("msm.round_pass2", st);
                (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass2_w6<1><<<blocks2, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass2<1><<<blocks2, 128, 0, st>>>(ra)));
            }
        } else {
            {
                uint32_t fb_blocks = (nbuckets + 7) / 8;  // 8 warps per CTA, one bucket per warp
                if (fb_blocks > 16u * kNumSMs) fb_blocks = 16u * kNumSMs;
                (++fb::g_launches, k_slot_fill<<<fb_blocks, 256, 0, st>>>(off_cur, off_out, nbuckets, wbuf));
            }
            prof_mark("msm.round_pass1", st);
            (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass1_w6<0><<<blocks, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass1<0><<<blocks, 128, 0, st>>>(ra)));
            prof_mark("msm.round_pass2", st);
            (round_warps == 6 ? (void)(++fb::g_launches, k_round_pass2_w6<0><<<blocks2, 192, 0, st>>>(ra)) : (void)(++fb::g_launches, k_round_pass2<0><<<blocks2, 128, 0, st>>>(ra)));
        }
        if (stats_on) k_stat_add<<<1, 1, 0, st>>>(stats, 2, off_out + nbuckets, 0);
        pts_cur = reinterpret_cast<const char*>(ra.out);
        stride_cur = (uint32_t)sizeof(G1Affine);
        off_cur = off_out;
        ub = ub_out;
    }

    if (stats_on) k_stat_add<<<1, 1, 0, st>>>(stats, 3, off_cur + nbuckets, 0);
    // accumulate: one wave of resident threads, equal chunks of the (remaining) entries
    prof_mark("msm.accumulate", st);
    static int acc_blocks_per_sm = 0;
    if (!acc_blocks_per_sm) {
        FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&acc_blocks_per_sm, k_accumulate<4, false>, 128, 0));
        if (acc_blocks_per_sm < 1) acc_blocks_per_sm = 1;
    }
    uint32_t acc_blocks = (uint32_t)(kNumSMs * acc_blocks_per_sm);
    {   // small inputs: no more threads than chunks of kMinChunk entries
        uint64_t need = (ub / kMinChunk + 127) / 128 + 1;
        if (need < acc_blocks) acc_blocks = (uint32_t)need;
    }
    const uint32_t nt = acc_blocks * 128;
    G1Xyzz* partial = s.seg_t.as<G1Xyzz>((size_t)nbuckets + nt);
    if (rounds > 0) (++fb::g_launches, k_accumulate<4, true><<<acc_blocks, 128, 0, st>>>(pts_cur, stride_cur, nullptr, off_cur, nbuckets, partial));
    else (++fb::g_launches, k_accumulate<4, false><<<acc_blocks, 128, 0, st>>>(pts_cur, stride_cur, sorted, offsets, nbuckets, partial));
    const uint32_t* offsets_final = off_cur;
    prof_mark("msm.combine", st);
    uint32_t* long_list = s.order.as<uint32_t>((size_t)nt + 2);
    FB_CUDA(cudaMemsetAsync(long_list, 0, sizeof(uint32_t), st));
    (++fb::g_launches, k_combine<<<(nbuckets + 127) / 128, 128, 0, st>>>(partial, offsets_final, nbuckets, nt, buckets, long_list));
    (++fb::g_launches, k_combine_long<<<2 * kNumSMs, kLongThreads, 0, st>>>(partial, offsets_final, nbuckets, nt, buckets, long_list));

    // bucket reduction
    prof_mark("msm.reduce", st);
    BitsArgs ba;
    ba.X = buckets;
    ba.C = ba.R = nullptr;
    ba.k = k;
    ba.l = ba.h = 0;
    if (k > kDirectBits) {
        const int l = (k + 1) / 2, h = k - l;
        // ping-pong arenas: every pass shrinks its input by >= 2
        G1Xyzz* arena = s.seg_a.as<G1Xyzz>((size_t)2 * nbuckets);
        G1Xyzz* row_buf[2] = {arena, arena + nbuckets / 2};
        G1Xyzz* col_buf[2] = {arena + nbuckets, arena + nbuckets + nbuckets / 2};
        const G1Xyzz* rin = buckets;
        const G1Xyzz* cin = buckets;
        static const uint64_t fold_wide = [] {  // FB_MSM_FOLD_WIDE: outputs of a pass from which it folds four inputs per thread
            const char* e = getenv("FB_MSM_FOLD_WIDE");
            return e ? (uint64_t)atoll(e) : (uint64_t)kNumSMs * 4 * 32 * 2;  // two warps per scheduler
        }();
        uint32_t rl = 1u << l;  // remaining row length (row sums shrink it to 1)
        uint32_t ch = 1u << h;  // remaining column height
        int pp = 0;
        while (rl > 1 || ch > 1) {
            FoldArgs fa;
            fa.d[0] = FoldDir{rin, row_buf[pp], 0, 1, 1};
            fa.d[1] = FoldDir{cin, col_buf[pp], 0, 1, 1};
            uint32_t mx = 0;
            // four inputs per output (three dependent additions per thread) while the pass has enough warps to keep the
            // multipliers busy; two per output (one addition, twice the threads, one more pass) once it does not - a
            // thread's additions are sequential, so the narrow passes are shorter as a tree
            if (rl > 1) {
                uint32_t G = (rl >= 4 && (uint64_t)nsets * (1u << h) * (rl / 4) >= fold_wide) ? 4 : 2;
                rl /= G;
                fa.d[0].G = G;
                fa.d[0].L = 1;
                fa.d[0].n_out = nsets * (1u << h) * rl;
                rin = row_buf[pp];
                mx = fa.d[0].n_out;
            }
            if (ch > 1) {
                uint32_t G = (ch >= 4 && (uint64_t)nsets * (ch / 4) * (1u << l) >= fold_wide) ? 4 : 2;
                ch /= G;
                fa.d[1].G = G;
                fa.d[1].L = 1u << l;
                fa.d[1].n_out = nsets * ch * (1u << l);
                cin = col_buf[pp];
                if (fa.d[1].n_out > mx) mx = fa.d[1].n_out;
            }
            (++fb::g_launches, k_fold<<<dim3((mx + 127) / 128, 2), 128, 0, st>>>(fa));
            pp ^= 1;
        }
        ba.R = rin;
        ba.C = cin;
        ba.l = l;
        ba.h = h;
    }
    G1Xyzz* Q = s.misc.as<G1Xyzz>((size_t)nsets * (k + 1) + nsets);
    G1Xyzz* S = Q + (size_t)nsets * (k + 1);
    ba.Q = Q;
    ba.njobs = (uint32_t)(nsets * (k + 1));
    {
        // most addends of a job, log2: column bits have 2^(l-1), row bits 2^(h-1), the plain total 2^h (2^k when direct)
        const int widest = ba.l ? (ba.l - 1 > ba.h ? ba.l - 1 : ba.h) : k;
        if (widest <= 5) (++fb::g_launches, k_bits<32><<<(ba.njobs + kBitsThreads / 32 - 1) / (kBitsThreads / 32), kBitsThreads, 0, st>>>(ba));
        else if (widest == 6) (++fb::g_launches, k_bits<64><<<(ba.njobs + kBitsThreads / 64 - 1) / (kBitsThreads / 64), kBitsThreads, 0, st>>>(ba));
        else (++fb::g_launches, k_bits<kBitsThreads><<<ba.njobs, kBitsThreads, 0, st>>>(ba));
    }
    prof_mark("msm.finish", st);
    (++fb::g_launches, k_bit_combine<<<nsets, 32, 0, st>>>(Q, k, tab ? d_out : S));
    if (!tab) (++fb::g_launches, k_window_combine<<<(unsigned)batch, 64, 0, st>>>(S, (uint32_t)nwin, c, d_out));
    prof_mark("post", st);
    FB_CUDA(cudaGetLastError());
}

void msm_run_split(const MsmBases& bases, const Fr* d_scalars, size_t len, size_t batch, bool scalars_mont, G1Xyzz* d_out,
                   MsmExec& ex, const Fr* d_scalars2) {
    const bool shared = d_scalars2 != nullptr;
    const bool by_rows = !shared && batch >= 2;
    int S = ex.S;
    if (S > kMaxSplit) S = kMaxSplit;
    // The shared-bases pair (commit + open of one row) splits into its two rows first - two plain MSMs over the same
    // table on two streams - and only then into column halves.  Every sub-MSM keeps at least 2^17 points (columns) /
    // 64 rows: below that the per-launch costs dominate.
    int col_parts = shared ? S / 2 : S;  // column ranges per scalar vector
    if (col_parts < 1) col_parts = 1;
    if (by_rows) {
        while (S > 1 && batch / S < 64) S--;
    } else {
        while (col_parts > 1 && len / col_parts < ((size_t)1 << 17)) col_parts--;
        if (shared && len < ((size_t)1 << 17)) S = 1;       // small pair: one fused pass
        else S = shared ? 2 * col_parts : col_parts;
    }
    if (S <= 1) {
        msm_run(bases, d_scalars, len, batch, scalars_mont, d_out, *ex.scratch[0], ex.stream[0], d_scalars2);
        return;
    }
    cudaStream_t st0 = ex.stream[0];
    FB_CUDA(cudaEventRecord(ex.fork, st0));
    const size_t per_row = (size_t)(bases.c > 0 ? bases.nwin : 1) * bases.n;  // entries of one row of the (tabulated) bases
    const int nout = shared ? 2 : 1;
    // column split: parts[k][b]; a sub-MSM of the pair writes one output (vector v, column range q: k = v * col_parts + q)
    G1Xyzz* parts = by_rows ? nullptr : ex.parts->as<G1Xyzz>((size_t)S);
    for (int k = 0; k < S; k++) {
        cudaStream_t st = ex.stream[k];
        if (k) FB_CUDA(cudaStreamWaitEvent(st, ex.fork, 0));
        if (by_rows) {
            const size_t r0 = batch * k / S, r1 = batch * (k + 1) / S;
            MsmBases b = bases;
            b.pts = (char*)bases.pts + r0 * per_row * bases.stride();
            if (bases.xs) b.xs = bases.xs + r0 * per_row * 64;
            msm_run(b, d_scalars + r0 * len, len, r1 - r0, scalars_mont, d_out + r0, *ex.scratch[k], st);
        } else {
            const int v = k / col_parts, q = k % col_parts;
            const size_t c0 = len * q / col_parts, c1 = len * (q + 1) / col_parts;
            MsmBases b = bases;
            b.pts = (char*)bases.pts + c0 * bases.stride();  // index = w * bases.n + column: shifting the base shifts the column
            if (bases.xs) b.xs = bases.xs + c0 * 64;
            msm_run(b, (v ? d_scalars2 : d_scalars) + c0, c1 - c0, 1, scalars_mont, parts + k, *ex.scratch[k], st);
        }
        if (k) {
            FB_CUDA(cudaEventRecord(ex.join[k], st));
            FB_CUDA(cudaStreamWaitEvent(st0, ex.join[k], 0));
        }
    }
    if (!by_rows) (++fb::g_launches, k_sum_parts<<<1, 32, 0, st0>>>(parts, col_parts, nout, d_out));
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
