// evalform.cu - the field half of worker_open without transforms.
//
// The reference (src/engine/piano.rs:193-236) takes the evaluations c_j = p(w^j),
// interpolates (iNTT), evaluates y = p(alpha), divides q = (p - y)/(X - alpha) and
// transforms back: q_hat_j = q(w^j).  The same field elements follow directly from
// the evaluations:
//     alpha outside the domain:  y = (alpha^N - 1)/N * sum_j c_j w^j / (alpha - w^j)
//                                q_hat_j = (c_j - y) / (w^j - alpha)
//     alpha = w^k:               y = c_k,  q_hat_j as above for j != k, and
//                                q_hat_k = sum_{j != k} (c_j - c_k) w^j / (w^k (w^k - w^j))
// (exact arithmetic in Fr, so the bytes are identical).  With inv_j = 1/(w^j - alpha)
// (inv_k = 0), h_j = w^j inv_j, S1 = sum_j c_j h_j and S0 = sum_j h_j:
//     y = -(alpha^N - 1)/N * S1            or c_k
//     q_hat_k = -w^(-k) (S1 - c_k S0)
// inv and h depend on alpha only, so a batch of rows opened at the same alpha (the
// bivariate row loop) shares them: 2 multiplies per coefficient and row instead of two
// transforms and a scan.  One binary-GCD inversion per CTA (Montgomery's trick, batchinv.cuh).
#include "common.cuh"
#include "batchinv.cuh"

namespace fb {

namespace {

using F = FrParams;
constexpr int kPer = 4;          // fewest elements per thread in the inversion kernel (small inputs)
constexpr int kInvBlocksPerSM = 4;  // its grid is capped at one resident wave
constexpr int kThreads = 128;
constexpr int kDotPer = 8;       // elements per thread in the dot-product kernel
constexpr uint32_t kNoIndex = 0xffffffffu;

struct EfArgs {
    const Fr* lo;     // w^e = hi[e >> lo_bits] * lo[e & mask]
    const Fr* hi;
    int lo_bits;
    int log_n;
    const Fr* inv_n;  // 1/N
    const Fr* alpha;
    const Fr* evals;  // [batch][cnt]
    size_t n, batch;  // n = domain size N
    size_t j0, cnt;   // the elements at hand are j0 .. j0 + cnt - 1 of the domain (whole rows: 0, N; a column shard: its slice)
    Fr* inv;          // [cnt]
    Fr* h;            // [cnt]
    Fr* partial;      // [batch + 1][nblk]
    Fr* sums;         // [batch + 1]: S1 per row, then S0
    Fr* neg_s;        // -(alpha^N - 1)/N
    uint32_t* kidx;   // index of alpha in the domain, or kNoIndex
    uint32_t nblk;    // dot-product blocks per row
    Fr* y;            // [batch]
    Fr* qhat;         // [batch][cnt]
    // column shards (batch = 1): the partial sums of every shard, 4 Fr each: S1 | S0 | c_k (if alpha = w^k lies in the
    // shard's slice) | k or kNoIndex in the first word
    Fr* state_out;
    const Fr* states_in;
    int nparts;
};

__device__ __forceinline__ Fr omega_pow(const EfArgs& a, size_t e) {
    return fe_mul<F>(a.hi[e >> a.lo_bits], a.lo[e & (((size_t)1 << a.lo_bits) - 1)]);
}

__global__ void k_ef_consts(EfArgs a) {
    Fr p = *a.alpha;
    for (int i = 0; i < a.log_n; i++) p = fe_sqr<F>(p);
    Fr s = fe_mul<F>(fe_sub<F>(p, Fr::one()), *a.inv_n);
    *a.neg_s = fe_neg<F>(s);
    *a.kidx = kNoIndex;
}

// inv_j = 1/(w^j - alpha) (0 where w^j = alpha, whose index goes to kidx) and h_j = w^j inv_j.
// ONE resident wave: thread t owns elements i*nt + t (a warp touches contiguous runs), walks them once
// forwards keeping the running product of its denominators (the product before element j is parked in
// inv[j], w^j in h[j]), the CTA shares one inversion, and the walk back peels the inverses off.  The
// inversion is a ~50-80 us dependent chain on one lane: it is paid once per launch, not once per 512
// elements in each of several waves (that form took 373 us at 2^20).
__global__ void __launch_bounds__(kThreads) k_ef_inverse(EfArgs a) {
    static_assert(kThreads == 128, "block_batch_inverse<F, 4> expects 4 warps");
    const size_t nt = (size_t)gridDim.x * kThreads;
    const size_t t = (size_t)blockIdx.x * kThreads + threadIdx.x;
    const Fr alpha = *a.alpha;
    Fr run = Fr::one();
    for (size_t j = t; j < a.cnt; j += nt) {
        const Fr w = omega_pow(a, a.j0 + j);
        const Fr d = fe_sub<F>(w, alpha);
        a.h[j] = w;
        a.inv[j] = run;
        if (d.is_zero()) *a.kidx = (uint32_t)(a.j0 + j);
        else run = fe_mul<F>(run, d);
    }
    Fr inv = block_batch_inverse<F, 4>(run);  // 1 / (product of all the thread's denominators)
    if (t >= a.cnt) return;
    for (size_t j = t + ((a.cnt - 1 - t) / nt) * nt;; j -= nt) {  // the thread's last element, then backwards
        const Fr w = a.h[j];
        const Fr d = fe_sub<F>(w, alpha);
        Fr v = Fr::zero();
        if (!d.is_zero()) {
            v = fe_mul<F>(inv, a.inv[j]);
            inv = fe_mul<F>(inv, d);
        }
        a.inv[j] = v;
        a.h[j] = fe_mul<F>(w, v);
        if (j < nt) break;
    }
}

__device__ __forceinline__ Fr block_sum(Fr acc, Fr* sm) {
    sm[threadIdx.x] = acc;
    for (uint32_t s = kThreads / 2; s > 0; s >>= 1) {
        __syncthreads();
        if (threadIdx.x < s) sm[threadIdx.x] = fe_add<F>(sm[threadIdx.x], sm[threadIdx.x + s]);
    }
    __syncthreads();
    return sm[0];
}

// partial[row][blk] = sum over the block's slice of c_row,j * h_j; row == batch: sum of h_j
__global__ void __launch_bounds__(kThreads) k_ef_dot(EfArgs a) {
    __shared__ Fr sm[kThreads];
    const size_t row = blockIdx.y;
    const Fr* c = row < a.batch ? a.evals + row * a.cnt : nullptr;
    const size_t base = (size_t)blockIdx.x * (kThreads * kDotPer) + threadIdx.x;
    Fr acc = Fr::zero();
#pragma unroll
    for (int i = 0; i < kDotPer; i++) {
        const size_t j = base + (size_t)i * kThreads;
        if (j < a.cnt) acc = fe_add<F>(acc, c ? fe_mul<F>(c[j], a.h[j]) : a.h[j]);
    }
    Fr total = block_sum(acc, sm);
    if (threadIdx.x == 0) a.partial[row * a.nblk + blockIdx.x] = total;
}

// sums[row] and, for the rows proper, y
__global__ void __launch_bounds__(kThreads) k_ef_finish(EfArgs a) {
    __shared__ Fr sm[kThreads];
    const size_t row = blockIdx.x;
    Fr acc = Fr::zero();
    for (uint32_t b = threadIdx.x; b < a.nblk; b += kThreads) acc = fe_add<F>(acc, a.partial[row * a.nblk + b]);
    Fr total = block_sum(acc, sm);
    if (threadIdx.x == 0) {
        a.sums[row] = total;
        if (row < a.batch) {
            const uint32_t k = *a.kidx;
            a.y[row] = k == kNoIndex ? fe_mul<F>(*a.neg_s, total) : a.evals[row * a.cnt + (k - a.j0)];
        }
    }
}

__global__ void __launch_bounds__(kThreads) k_ef_quotient(EfArgs a) {
    const size_t j = (size_t)blockIdx.x * kThreads + threadIdx.x;
    const size_t row = blockIdx.y;
    if (j >= a.cnt) return;
    const Fr y = a.y[row];
    Fr q = fe_mul<F>(fe_sub<F>(a.evals[row * a.cnt + j], y), a.inv[j]);
    if (a.j0 + j == *a.kidx) {  // alpha = w^k: q_hat_k = -w^(-k) (S1 - c_k S0)
        Fr t = fe_sub<F>(a.sums[row], fe_mul<F>(y, a.sums[a.batch]));
        Fr wk_inv = omega_pow(a, (a.n - (a.j0 + j)) & (a.n - 1));
        q = fe_neg<F>(fe_mul<F>(wk_inv, t));
    }
    a.qhat[row * a.cnt + j] = q;
}

// column shard, first half: this shard's partial sums (rows 0 = S1, 1 = S0 of `partial`) -> state_out
__global__ void __launch_bounds__(kThreads) k_ef_shard_state(EfArgs a) {
    __shared__ Fr sm[kThreads];
    const size_t row = blockIdx.x;  // 0: S1, 1: S0
    Fr acc = Fr::zero();
    for (uint32_t b = threadIdx.x; b < a.nblk; b += kThreads) acc = fe_add<F>(acc, a.partial[row * a.nblk + b]);
    Fr total = block_sum(acc, sm);
    if (threadIdx.x == 0) {
        a.state_out[row] = total;
        if (row == 0) {
            const uint32_t k = *a.kidx;
            a.state_out[2] = k == kNoIndex ? Fr::zero() : a.evals[k - a.j0];
            Fr tag = Fr::zero();
            tag.v[0] = k;
            a.state_out[3] = tag;
        }
    }
}

// column shard, second half: all shards' partial sums -> S1, S0, the index of alpha in the domain and y
__global__ void k_ef_shard_reduce(EfArgs a) {
    if (threadIdx.x || blockIdx.x) return;
    Fr s1 = Fr::zero(), s0 = Fr::zero(), ck = Fr::zero();
    uint32_t k = kNoIndex;
    for (int p = 0; p < a.nparts; p++) {
        const Fr* st = a.states_in + 4 * p;
        s1 = fe_add<F>(s1, st[0]);
        s0 = fe_add<F>(s0, st[1]);
        if (st[3].v[0] != kNoIndex && k == kNoIndex) {
            k = st[3].v[0];
            ck = st[2];
        }
    }
    a.sums[0] = s1;
    a.sums[1] = s0;
    *a.kidx = k;
    a.y[0] = k == kNoIndex ? fe_mul<F>(*a.neg_s, s1) : ck;
}

}  // namespace

// arguments + scratch layout for `cnt` elements starting at j0 of the size-n domain, `batch` rows
static EfArgs ef_args(const NttTables& tab, const Fr* d_evals, size_t n, size_t j0, size_t cnt, size_t batch, const Fr* d_alpha,
                      DevBuf& scratch) {
    EfArgs a;
    a.lo = tab.lo[0];
    a.hi = tab.hi[0];
    a.lo_bits = tab.lo_bits;
    a.log_n = tab.log_n;
    a.inv_n = tab.consts + 2;
    a.alpha = d_alpha;
    a.evals = d_evals;
    a.n = n;
    a.j0 = j0;
    a.cnt = cnt;
    a.batch = batch;
    a.nblk = (uint32_t)((cnt + kThreads * kDotPer - 1) / (kThreads * kDotPer));
    if (a.nblk == 0) a.nblk = 1;
    const size_t need = 2 * cnt + (batch + 1) * a.nblk + (batch + 1) + 2;
    Fr* buf = scratch.as<Fr>(need);
    a.inv = buf;
    a.h = buf + cnt;
    a.partial = a.h + cnt;
    a.sums = a.partial + (batch + 1) * a.nblk;
    a.neg_s = a.sums + (batch + 1);
    a.kidx = reinterpret_cast<uint32_t*>(a.neg_s + 1);
    a.y = nullptr;
    a.qhat = nullptr;
    a.state_out = nullptr;
    a.states_in = nullptr;
    a.nparts = 0;
    return a;
}

static void ef_launch_inverse(const EfArgs& a, cudaStream_t st) {
    size_t blocks = (a.cnt + kThreads * kPer - 1) / (kThreads * kPer);
    if (blocks > (size_t)kNumSMs * kInvBlocksPerSM) blocks = (size_t)kNumSMs * kInvBlocksPerSM;  // one resident wave
    if (blocks) (++g_launches, k_ef_inverse<<<(unsigned)blocks, kThreads, 0, st>>>(a));
}

void evalform_open(const NttTables& tab, const Fr* d_evals, size_t n, size_t batch, const Fr* d_alpha, Fr* d_y, Fr* d_qhat,
                   DevBuf& scratch, cudaStream_t st) {
    if (!n || !batch) return;
    EfArgs a = ef_args(tab, d_evals, n, 0, n, batch, d_alpha, scratch);
    a.y = d_y;
    a.qhat = d_qhat;
    prof_mark("open.evalform", st);
    (++g_launches, k_ef_consts<<<1, 1, 0, st>>>(a));
    ef_launch_inverse(a, st);
    (++g_launches, k_ef_dot<<<dim3(a.nblk, (unsigned)(batch + 1)), kThreads, 0, st>>>(a));
    (++g_launches, k_ef_finish<<<(unsigned)(batch + 1), kThreads, 0, st>>>(a));
    (++g_launches, k_ef_quotient<<<dim3((unsigned)((n + kThreads - 1) / kThreads), (unsigned)batch), kThreads, 0, st>>>(a));
    FB_CUDA(cudaGetLastError());
}

// Column shards (one row split over GPUs by point ranges): the field half shards too.  Each shard runs the
// inversion and the dot products over ITS slice only and emits 128 bytes (S1 | S0 | c_k | k); the shards' states
// are exchanged (the one collective of this path besides the final gather), and every shard finishes its slice of
// the quotient's evaluations.  `session` keeps inv[] between the two halves.
void evalform_shard_begin(const NttTables& tab, const Fr* d_evals_slice, size_t n, size_t j0, size_t cnt, const Fr* d_alpha,
                          Fr* d_state, DevBuf& session, cudaStream_t st) {
    EfArgs a = ef_args(tab, d_evals_slice, n, j0, cnt, 1, d_alpha, session);
    a.state_out = d_state;
    prof_mark("open.evalform", st);
    (++g_launches, k_ef_consts<<<1, 1, 0, st>>>(a));
    FB_CUDA(cudaMemsetAsync(a.partial, 0, 2 * a.nblk * sizeof(Fr), st));  // an empty slice contributes zeros
    if (cnt) {
        ef_launch_inverse(a, st);
        (++g_launches, k_ef_dot<<<dim3(a.nblk, 2), kThreads, 0, st>>>(a));
    }
    (++g_launches, k_ef_shard_state<<<2, kThreads, 0, st>>>(a));
    FB_CUDA(cudaGetLastError());
}

void evalform_shard_finish(const NttTables& tab, const Fr* d_evals_slice, size_t n, size_t j0, size_t cnt, const Fr* d_alpha,
                           const Fr* d_states, int nparts, Fr* d_y, Fr* d_qhat_slice, DevBuf& session, cudaStream_t st) {
    EfArgs a = ef_args(tab, d_evals_slice, n, j0, cnt, 1, d_alpha, session);  // same layout: inv[] is still there
    a.states_in = d_states;
    a.nparts = nparts;
    a.y = d_y;
    a.qhat = d_qhat_slice;
    prof_mark("open.evalform", st);
    (++g_launches, k_ef_shard_reduce<<<1, 1, 0, st>>>(a));
    if (cnt) (++g_launches, k_ef_quotient<<<dim3((unsigned)((cnt + kThreads - 1) / kThreads), 1), kThreads, 0, st>>>(a));
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
