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
    const Fr* evals;  // [batch][N]
    size_t n, batch;
    Fr* inv;          // [N]
    Fr* h;            // [N]
    Fr* partial;      // [batch + 1][nblk]
    Fr* sums;         // [batch + 1]: S1 per row, then S0
    Fr* neg_s;        // -(alpha^N - 1)/N
    uint32_t* kidx;   // index of alpha in the domain, or kNoIndex
    uint32_t nblk;    // dot-product blocks per row
    Fr* y;            // [batch]
    Fr* qhat;         // [batch][N]
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
    for (size_t j = t; j < a.n; j += nt) {
        const Fr w = omega_pow(a, j);
        const Fr d = fe_sub<F>(w, alpha);
        a.h[j] = w;
        a.inv[j] = run;
        if (d.is_zero()) *a.kidx = (uint32_t)j;
        else run = fe_mul<F>(run, d);
    }
    Fr inv = block_batch_inverse<F, 4>(run);  // 1 / (product of all the thread's denominators)
    if (t >= a.n) return;
    for (size_t j = t + ((a.n - 1 - t) / nt) * nt;; j -= nt) {  // the thread's last element, then backwards
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
    const Fr* c = row < a.batch ? a.evals + row * a.n : nullptr;
    const size_t base = (size_t)blockIdx.x * (kThreads * kDotPer) + threadIdx.x;
    Fr acc = Fr::zero();
#pragma unroll
    for (int i = 0; i < kDotPer; i++) {
        const size_t j = base + (size_t)i * kThreads;
        if (j < a.n) acc = fe_add<F>(acc, c ? fe_mul<F>(c[j], a.h[j]) : a.h[j]);
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
            a.y[row] = k == kNoIndex ? fe_mul<F>(*a.neg_s, total) : a.evals[row * a.n + k];
        }
    }
}

__global__ void __launch_bounds__(kThreads) k_ef_quotient(EfArgs a) {
    const size_t j = (size_t)blockIdx.x * kThreads + threadIdx.x;
    const size_t row = blockIdx.y;
    if (j >= a.n) return;
    const Fr y = a.y[row];
    Fr q = fe_mul<F>(fe_sub<F>(a.evals[row * a.n + j], y), a.inv[j]);
    if (j == *a.kidx) {  // alpha = w^k: q_hat_k = -w^(-k) (S1 - c_k S0)
        Fr t = fe_sub<F>(a.sums[row], fe_mul<F>(y, a.sums[a.batch]));
        Fr wk_inv = omega_pow(a, (a.n - j) & (a.n - 1));
        q = fe_neg<F>(fe_mul<F>(wk_inv, t));
    }
    a.qhat[row * a.n + j] = q;
}

}  // namespace

void evalform_open(const NttTables& tab, const Fr* d_evals, size_t n, size_t batch, const Fr* d_alpha, Fr* d_y, Fr* d_qhat,
                   DevBuf& scratch, cudaStream_t st) {
    if (!n || !batch) return;
    EfArgs a;
    a.lo = tab.lo[0];
    a.hi = tab.hi[0];
    a.lo_bits = tab.lo_bits;
    a.log_n = tab.log_n;
    a.inv_n = tab.consts + 2;
    a.alpha = d_alpha;
    a.evals = d_evals;
    a.n = n;
    a.batch = batch;
    a.nblk = (uint32_t)((n + kThreads * kDotPer - 1) / (kThreads * kDotPer));
    const size_t need = 2 * n + (batch + 1) * a.nblk + (batch + 1) + 2;
    Fr* buf = scratch.as<Fr>(need);
    a.inv = buf;
    a.h = buf + n;
    a.partial = a.h + n;
    a.sums = a.partial + (batch + 1) * a.nblk;
    a.neg_s = a.sums + (batch + 1);
    a.kidx = reinterpret_cast<uint32_t*>(a.neg_s + 1);
    a.y = d_y;
    a.qhat = d_qhat;
    prof_mark("open.evalform", st);
    (++g_launches, k_ef_consts<<<1, 1, 0, st>>>(a));
    {
        size_t blocks = (n + kThreads * kPer - 1) / (kThreads * kPer);
        if (blocks > (size_t)kNumSMs * kInvBlocksPerSM) blocks = (size_t)kNumSMs * kInvBlocksPerSM;  // one resident wave
        (++g_launches, k_ef_inverse<<<(unsigned)blocks, kThreads, 0, st>>>(a));
    }
    (++g_launches, k_ef_dot<<<dim3(a.nblk, (unsigned)(batch + 1)), kThreads, 0, st>>>(a));
    (++g_launches, k_ef_finish<<<(unsigned)(batch + 1), kThreads, 0, st>>>(a));
    (++g_launches, k_ef_quotient<<<dim3((unsigned)((n + kThreads - 1) / kThreads), (unsigned)batch), kThreads, 0, st>>>(a));
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
