// poly.cu - wire codecs for Fr and the Horner / quotient scan.
//
// Replaces, for reference src/engine/piano.rs:194-225 and :268-282,
//   FsFr::from_bytes / to_bytes  (32-byte big-endian, canonical or "Invalid scalar"),
//   FsPoly::eval                 (Horner), and
//   FsPoly::div by (X - x)       (synthetic division, remainder dropped).
// eval and div are the same recurrence  s_k = a_k + x*s_(k+1):  y = s_0 and the
// quotient is q_k = s_(k+1).  On the GPU it runs as a blocked suffix scan: chunks
// of L coefficients are collapsed to one value each (up-sweep, with multiplier
// x^L), recursively, then the chunk boundary values are pushed back down.
#include "ctx.cuh"  // (common.cuh + the status codes of the C ABI, for fb_dev_evaluate_noalloc at the end)

namespace fb {

namespace {

constexpr int kChunkLog = 5;  // L = 32 coefficients per thread per level
constexpr int kChunk = 1 << kChunkLog;

__global__ void k_fr_from_be(const uint8_t* __restrict__ in, Fr* __restrict__ out, size_t n, int* err) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* src = reinterpret_cast<const uint4*>(in + 32 * i);
    uint4 a = src[0], b = src[1];  // a = bytes 0..15 (most significant), b = bytes 16..31
    Fr x;
    x.v[7] = __byte_perm(a.x, 0, 0x0123);
    x.v[6] = __byte_perm(a.y, 0, 0x0123);
    x.v[5] = __byte_perm(a.z, 0, 0x0123);
    x.v[4] = __byte_perm(a.w, 0, 0x0123);
    x.v[3] = __byte_perm(b.x, 0, 0x0123);
    x.v[2] = __byte_perm(b.y, 0, 0x0123);
    x.v[1] = __byte_perm(b.z, 0, 0x0123);
    x.v[0] = __byte_perm(b.w, 0, 0x0123);
    Fr m = Fr::modulus();
    if (fe_geq<FrParams>(x.v, m.v)) {
        *err = 1;
        x = Fr::zero();
    }
    out[i] = fe_to_mont<FrParams>(x);
}

__global__ void k_fr_to_be(const Fr* __restrict__ in, uint8_t* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr x = fe_from_mont<FrParams>(in[i]);
    uint4 a, b;
    a.x = __byte_perm(x.v[7], 0, 0x0123);
    a.y = __byte_perm(x.v[6], 0, 0x0123);
    a.z = __byte_perm(x.v[5], 0, 0x0123);
    a.w = __byte_perm(x.v[4], 0, 0x0123);
    b.x = __byte_perm(x.v[3], 0, 0x0123);
    b.y = __byte_perm(x.v[2], 0, 0x0123);
    b.z = __byte_perm(x.v[1], 0, 0x0123);
    b.w = __byte_perm(x.v[0], 0, 0x0123);
    uint4* dst = reinterpret_cast<uint4*>(out + 32 * i);
    dst[0] = a;
    dst[1] = b;
}

__global__ void k_fill_zero(uint4* p, size_t n16) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n16) p[i] = make_uint4(0, 0, 0, 0);
}

// pows[0] = x, pows[l+1] = pows[l]^L
__global__ void k_level_pows(const Fr* x, Fr* pows, int levels) {
    if (threadIdx.x || blockIdx.x) return;
    Fr p = *x;
    pows[0] = p;
    for (int l = 1; l <= levels; l++) {
        for (int i = 0; i < kChunkLog; i++) p = fe_sqr<FrParams>(p);
        pows[l] = p;
    }
}

// All scan kernels take a batch of polynomials: row r (blockIdx.y, or the thread index in
// k_horner_top) reads a + r*n and writes its outputs r*<row length> further on.

// h[j] = sum_i a[j*L + i] * beta^i
__global__ void k_horner_up(const Fr* __restrict__ a, size_t n, const Fr* beta_p, Fr* __restrict__ h, size_t nchunks) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nchunks) return;
    a += (size_t)blockIdx.y * n;
    h += (size_t)blockIdx.y * nchunks;
    const Fr beta = *beta_p;
    size_t lo = j << kChunkLog;
    size_t hi = lo + kChunk < n ? lo + kChunk : n;
    Fr acc = a[hi - 1];
    for (size_t k = hi - 1; k > lo; k--) acc = fe_add<FrParams>(fe_mul<FrParams>(acc, beta), a[k - 1]);
    h[j] = acc;
}

// sequential suffix scan of a short array: s[j] = a[j] + beta*s[j+1]; one thread per row
__global__ void k_horner_top(const Fr* a, size_t n, const Fr* beta_p, Fr* s, size_t batch) {
    size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= batch) return;
    a += r * n;
    s += r * n;
    const Fr beta = *beta_p;
    Fr acc = Fr::zero();
    for (size_t k = n; k > 0; k--) {
        acc = fe_add<FrParams>(fe_mul<FrParams>(acc, beta), a[k - 1]);
        s[k - 1] = acc;
    }
}

// Down-sweep: chunk j restarts from the suffix value of the next chunk (up[j+1])
// and regenerates s_k for its coefficients.  mode 0: write s (intermediate level);
// mode 1 (level 0): write y = s_0 and q_(k-1) = s_k, q_(n-1) = 0.
__global__ void k_horner_down(const Fr* __restrict__ a, size_t n, const Fr* beta_p, const Fr* __restrict__ up_s,
                              size_t nchunks, Fr* __restrict__ s_out, int mode, Fr* y_out) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nchunks) return;
    a += (size_t)blockIdx.y * n;
    s_out += (size_t)blockIdx.y * n;
    if (up_s) up_s += (size_t)blockIdx.y * nchunks;
    if (y_out) y_out += blockIdx.y;
    const Fr beta = *beta_p;
    size_t lo = j << kChunkLog;
    size_t hi = lo + kChunk < n ? lo + kChunk : n;
    Fr acc = (j + 1 < nchunks) ? up_s[j + 1] : Fr::zero();
    if (mode == 1 && hi == n) s_out[n - 1] = Fr::zero();
    for (size_t k = hi; k > lo; k--) {
        acc = fe_add<FrParams>(fe_mul<FrParams>(acc, beta), a[k - 1]);
        if (mode == 0) {
            s_out[k - 1] = acc;
        } else if (k - 1 > 0) {
            s_out[k - 2] = acc;
        } else {
            *y_out = acc;
        }
    }
}

// y[r] = s[r*stride]
__global__ void k_gather_first(const Fr* __restrict__ s, size_t stride, Fr* __restrict__ y, size_t batch) {
    size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < batch) y[r] = s[r * stride];
}

}  // namespace

void fr_from_be(const uint8_t* d_in, Fr* d_out, size_t n, int* d_err, cudaStream_t st) {
    if (!n) return;
    prof_mark("codec.fr_from_be", st);
    (++fb::g_launches, k_fr_from_be<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_in, d_out, n, d_err));
    FB_CUDA(cudaGetLastError());
}
void fr_to_be(const Fr* d_in, uint8_t* d_out, size_t n, cudaStream_t st) {
    if (!n) return;
    prof_mark("codec.fr_to_be", st);
    (++fb::g_launches, k_fr_to_be<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_in, d_out, n));
    FB_CUDA(cudaGetLastError());
}
void fr_fill_zero(Fr* d, size_t n, cudaStream_t st) {
    if (!n) return;
    size_t n16 = n * 2;
    (++fb::g_launches, k_fill_zero<<<(unsigned)((n16 + 255) / 256), 256, 0, st>>>((uint4*)d, n16));
    FB_CUDA(cudaGetLastError());
}

void horner_quotient(const Fr* d_a, size_t n, const Fr* d_x, Fr* d_y, Fr* d_q, DevBuf& scratch, cudaStream_t st,
                     size_t batch) {
    if (batch == 0) return;
    if (batch > 65535) throw CudaError{"horner_quotient: batch too large"};
    if (n == 0) {
        FB_CUDA(cudaMemsetAsync(d_y, 0, batch * sizeof(Fr), st));
        return;
    }
    prof_mark("poly.scan", st);
    // level sizes (per row)
    std::vector<size_t> sz{n};
    while (sz.back() > (size_t)kChunk) sz.push_back((sz.back() + kChunk - 1) >> kChunkLog);
    const int levels = (int)sz.size() - 1;  // number of up-sweeps
    size_t total = 0;
    for (int l = 1; l <= levels; l++) total += sz[l];
    // scratch: pows[levels+2] | h_1..h_levels | s_1..s_levels | s_top (levels == 0), each batch rows deep
    Fr* base = scratch.as<Fr>((size_t)(levels + 2) + batch * (2 * total + kChunk) + 8);
    Fr* pows = base;
    Fr* hbuf = base + levels + 2;
    Fr* sbuf = hbuf + batch * total;
    Fr* stop = sbuf + batch * total;
    const unsigned by = (unsigned)batch;
    const unsigned tb = (unsigned)((batch + 63) / 64);
    (++fb::g_launches, k_level_pows<<<1, 1, 0, st>>>(d_x, pows, levels));
    std::vector<const Fr*> arr(levels + 1);
    std::vector<Fr*> sarr(levels + 1, nullptr);
    arr[0] = d_a;
    {
        Fr* hp = hbuf;
        Fr* sp = sbuf;
        for (int l = 1; l <= levels; l++) {
            (++fb::g_launches, k_horner_up<<<dim3((unsigned)((sz[l] + 127) / 128), by), 128, 0, st>>>(arr[l - 1], sz[l - 1], pows + (l - 1), hp, sz[l]));
            arr[l] = hp;
            sarr[l] = sp;
            hp += batch * sz[l];
            sp += batch * sz[l];
        }
    }
    if (levels == 0) {
        // short input: one thread per row does it all
        if (d_q) {
            (++fb::g_launches, k_horner_down<<<dim3(1, by), 1, 0, st>>>(d_a, n, pows, nullptr, 1, d_q, 1, d_y));
        } else {
            (++fb::g_launches, k_horner_top<<<tb, 64, 0, st>>>(d_a, n, pows, stop, batch));
            (++fb::g_launches, k_gather_first<<<tb, 64, 0, st>>>(stop, n, d_y, batch));
        }
        FB_CUDA(cudaGetLastError());
        return;
    }
    (++fb::g_launches, k_horner_top<<<tb, 64, 0, st>>>(arr[levels], sz[levels], pows + levels, sarr[levels], batch));
    if (!d_q) {
        (++fb::g_launches, k_gather_first<<<tb, 64, 0, st>>>(sarr[levels], sz[levels], d_y, batch));
        FB_CUDA(cudaGetLastError());
        return;
    }
    for (int l = levels - 1; l >= 1; l--)
        (++fb::g_launches, k_horner_down<<<dim3((unsigned)((sz[l + 1] + 127) / 128), by), 128, 0, st>>>(arr[l], sz[l], pows + l, sarr[l + 1], sz[l + 1],
                                                                            sarr[l], 0, nullptr));
    (++fb::g_launches, k_horner_down<<<dim3((unsigned)((sz[1] + 127) / 128), by), 128, 0, st>>>(d_a, n, pows, sarr[1], sz[1], d_q, 1, d_y));
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb

// ---------------------------------------------------------------------------------------------------------
// fb_dev_evaluate_noalloc - FsPoly::eval for callers that cannot allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as horner_quotient's evaluation-only path (k_level_pows,
// k_horner_up per level, k_horner_top, k_gather_first); the only difference is where the scratch lives:
// a file-scope __device__ array instead of a grow-on-demand cudaMalloc'ed buffer.
//
//   * kNoAllocMaxLen = 2^20 coefficients.  Scratch = level powers + the collapsed levels
//     (2^15 + 2^10 + 2^5 values) + the top level's suffix values, 33 920 Fr = 1.04 MiB, held by every GPU
//     that loads the library until the process exits.
//   * Every call shares the array and it keeps the last call's contents.  Nothing has to start at zero:
//     every word a call reads was written earlier in the same call (level l's sz[l] values by k_horner_up,
//     the top level's by k_horner_top), so stale data from a longer, earlier call is never read.
//   * Calls that overlap on the GPU (two streams, two host threads) RACE on the array: serialise them.
//   * The array's name is a host-side stand-in in host code, not a device address: the address is looked up
//     once per device, with that device current, by cudaGetSymbolAddress, and that pointer goes to the kernels.
//     The look-up happens on the first call for a device: make one call before starting a graph capture.
namespace fb {
namespace {

constexpr uint64_t kNoAllocMaxLen = 1ull << 20;
constexpr int kNoAllocMaxLevels = 4;  // 2^20 -> 2^15 -> 2^10 -> 2^5: three up-sweeps
constexpr size_t kNoAllocScratchFr = (kNoAllocMaxLevels + 2) + ((1u << 15) + (1u << 10) + (1u << 5)) + kChunk + 58;  // 33 920
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchFr * sizeof(Fr)];

constexpr int kNoAllocMaxDevices = 64;
Fr* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;

// device address of g_noalloc_scratch on `device` (which is current); nullptr + *err on failure
Fr* noalloc_scratch_for(int device, cudaError_t* err) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        *err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (*err != cudaSuccess) return nullptr;
        g_noalloc_ptr[device] = static_cast<Fr*>(p);
    }
    return g_noalloc_ptr[device];
}

}  // namespace
}  // namespace fb

extern "C" int fb_dev_evaluate_noalloc(int device, const void* d_poly_fr, uint64_t len, const void* d_x_fr, void* d_out_y_fr,
                                       void* cuda_stream) {
    using namespace fb;
    if (device < 0 || device >= kNoAllocMaxDevices) return ctx_fail(nullptr, FB_ERR_ARG, "fb_dev_evaluate_noalloc: device ordinal out of range");
    if (!d_x_fr || !d_out_y_fr || (len && !d_poly_fr)) return ctx_fail(nullptr, FB_ERR_ARG, "fb_dev_evaluate_noalloc: null pointer");
    if (len > kNoAllocMaxLen)
        return ctx_fail(nullptr, FB_ERR_ARG, "fb_dev_evaluate_noalloc: more than 2^20 coefficients (the static scratch is sized for 2^20); use fb_evaluate");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return ctx_fail(nullptr, FB_ERR_CUDA, "fb_dev_evaluate_noalloc: cudaSetDevice failed");
    }
    Fr* base = noalloc_scratch_for(device, &e);
    if (!base) {
        cudaGetLastError();
        return ctx_fail(nullptr, FB_ERR_CUDA, "fb_dev_evaluate_noalloc: cudaGetSymbolAddress failed");
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const Fr* d_a = static_cast<const Fr*>(d_poly_fr);
    const Fr* d_x = static_cast<const Fr*>(d_x_fr);
    Fr* d_y = static_cast<Fr*>(d_out_y_fr);
    if (len == 0) {  // the empty polynomial evaluates to 0
        e = cudaMemsetAsync(d_y, 0, sizeof(Fr), st);
    } else {
        size_t sz[kNoAllocMaxLevels + 1];
        int levels = 0;
        sz[0] = (size_t)len;
        while (sz[levels] > (size_t)kChunk) {
            sz[levels + 1] = (sz[levels] + kChunk - 1) >> kChunkLog;
            levels++;
        }
        // layout: pows[levels + 2] | h_1 .. h_levels | s_top (<= kChunk values)
        Fr* pows = base;
        Fr* hp = base + (kNoAllocMaxLevels + 2);
        (++fb::g_launches, k_level_pows<<<1, 1, 0, st>>>(d_x, pows, levels));
        const Fr* cur = d_a;
        for (int l = 1; l <= levels; l++) {
            (++fb::g_launches, k_horner_up<<<dim3((unsigned)((sz[l] + 127) / 128), 1), 128, 0, st>>>(cur, sz[l - 1], pows + (l - 1), hp, sz[l]));
            cur = hp;
            hp += sz[l];
        }
        Fr* stop = hp;  // suffix values of the top level (sz[levels] <= kChunk of them)
        (++fb::g_launches, k_horner_top<<<1, 64, 0, st>>>(cur, sz[levels], pows + levels, stop, 1));
        (++fb::g_launches, k_gather_first<<<1, 64, 0, st>>>(stop, sz[levels], d_y, 1));
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return ctx_fail(nullptr, FB_ERR_CUDA, "fb_dev_evaluate_noalloc: kernel launch failed");
    }
    return FB_OK;
}
