// g1ops.cu - batched G1 utilities around the MSM: wire codecs on the GPU (the
// setup-point loader), XYZZ -> affine, fixed-base multiplication by the generator
// (SRS generation) and plain point sums (master_commit / pi_0).
//
// Reference counterparts: FsG1::from_bytes / deserialize inside
// kzg::io_utils::batch_reader (src/engine/piano.rs:762-819), FsG1::to_bytes /
// serialize (piano.rs:649-706, src/rpc.rs:241, 268-269, 311, 330), G1::mul in
// generate_trusted_setup (piano.rs:916-1007) and the all-ones
// g1_linear_combination of master_commit / master_open (piano.rs:164-175,
// 258-265).
#include "common.cuh"

namespace fb {

namespace {

__global__ void __launch_bounds__(128) k_decompress(const uint8_t* __restrict__ in, G1Affine* __restrict__ out, size_t n, int* err) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    __align__(16) uint8_t buf[48];
    const uint4* src = reinterpret_cast<const uint4*>(in + 48 * i);
    uint4* b4 = reinterpret_cast<uint4*>(buf);
    b4[0] = src[0];
    b4[1] = src[1];
    b4[2] = src[2];
    G1Affine p;
    int rc = g1_decompress(buf, &p);
    if (rc) {
        atomicCAS(err, 0, rc);
        p = G1Affine::inf();
    }
    out[i] = p;
}

__global__ void __launch_bounds__(128) k_deserialize(const uint8_t* __restrict__ in, G1Affine* __restrict__ out, size_t n, int* err) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint8_t buf[96];
    for (int k = 0; k < 96; k++) buf[k] = in[96 * i + k];
    G1Affine p;
    int rc = g1_deserialize(buf, &p);
    if (rc) {
        atomicCAS(err, 0, rc);
        p = G1Affine::inf();
    }
    out[i] = p;
}

__global__ void __launch_bounds__(128) k_compress(const G1Affine* __restrict__ in, uint8_t* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint8_t buf[48];
    g1_compress(in[i], buf);
    for (int k = 0; k < 48; k++) out[48 * i + k] = buf[k];
}

__global__ void __launch_bounds__(128) k_serialize(const G1Affine* __restrict__ in, uint8_t* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint8_t buf[96];
    g1_serialize(in[i], buf);
    for (int k = 0; k < 96; k++) out[96 * i + k] = buf[k];
}

__global__ void __launch_bounds__(128) k_to_affine(const G1Xyzz* __restrict__ in, G1Affine* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // few results, one inversion each: the latency of the dependent chain is what
    // counts here, so the division-step inversion (fe_inv_safegcd) replaces Fermat's ladder
    const G1Xyzz p = in[i];
    if (p.is_inf()) {
        out[i] = G1Affine::inf();
        return;
    }
    out[i] = g1_to_affine_with_inv(p, fe_inv_safegcd<FpParams>(p.zzz));
}

// generator comb: tab[w*255 + d-1] = d * 2^(8w) * G  (w < 32, 1 <= d <= 255)
__global__ void k_gen_table_xyzz(G1Xyzz* tab) {
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= 32) return;
    G1Xyzz base = G1Xyzz::from_affine(g1_generator());
    for (int i = 0; i < 8 * w; i++) base = g1_dbl(base);
    G1Xyzz acc = base;
    tab[w * 255] = acc;
    for (int d = 2; d <= 255; d++) {
        g1_add(acc, base);
        tab[w * 255 + d - 1] = acc;
    }
}

__global__ void __launch_bounds__(128) k_fixed_base_mul(const Fr* __restrict__ scalars, size_t n, const G1Affine* __restrict__ tab,
                                                        G1Affine* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr k = fe_from_mont<FrParams>(scalars[i]);
    G1Xyzz acc = G1Xyzz::inf();
    for (int w = 0; w < 32; w++) {
        uint32_t d = (k.v[w >> 2] >> (8 * (w & 3))) & 0xffu;
        if (d) g1_add_mixed(acc, tab[w * 255 + d - 1]);
    }
    out[i] = g1_to_affine(acc);
}

__global__ void __launch_bounds__(128) k_sum_affine(const G1Affine* __restrict__ in, size_t n, G1Xyzz* __restrict__ out, size_t nout) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nout) return;
    size_t lo = t * 8, hi = lo + 8 < n ? lo + 8 : n;
    G1Xyzz acc = G1Xyzz::inf();
    for (size_t k = lo; k < hi; k++) g1_add_mixed(acc, in[k]);
    out[t] = acc;
}

__global__ void __launch_bounds__(128) k_sum_xyzz(const G1Xyzz* __restrict__ in, size_t n, G1Xyzz* __restrict__ out, size_t nout) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nout) return;
    size_t lo = t * 4, hi = lo + 4 < n ? lo + 4 : n;
    G1Xyzz acc = in[lo];
    for (size_t k = lo + 1; k < hi; k++) g1_add(acc, in[k]);
    out[t] = acc;
}

// parts: nparts x (96 | 32 | 96) bytes; y is every part's (they agree) and is copied from part 0.  out: 48 | 32 | 48.
// One warp: lane t takes kind t & 1 (0 the commitments' partials, 1 the proofs') of parts t >> 1, t >> 1 + 16, ...;
// the lanes of one kind then add up in a shared-memory tree - log2(nparts) dependent additions instead of nparts.
__global__ void __launch_bounds__(32) k_combine_parts(const uint8_t* __restrict__ parts, size_t nparts, uint8_t* __restrict__ out,
                                                      int* err) {
    __shared__ G1Xyzz sm[32];
    const int t = threadIdx.x, kind = t & 1;
    G1Xyzz acc = G1Xyzz::inf();
    for (size_t p = t >> 1; p < nparts; p += 16) {
        uint8_t buf[96];
        const uint8_t* src = parts + p * 224 + (kind ? 128 : 0);
        for (int k = 0; k < 96; k++) buf[k] = src[k];
        G1Affine q;
        int rc = g1_deserialize(buf, &q);
        if (rc) {
            atomicCAS(err, 0, rc);
            q = G1Affine::inf();
        }
        g1_add_mixed(acc, q);
    }
    sm[t] = acc;
    for (int s = 16; s >= 2; s >>= 1) {
        __syncwarp();
        if (t < s) {
            G1Xyzz x = sm[t];
            g1_add(x, sm[t + s]);  // t and t + s are of one kind: s is even
            sm[t] = x;
        }
    }
    __syncwarp();
    if (t == 2)
        for (int k = 0; k < 32; k++) out[48 + k] = parts[96 + k];
    if (t >= 2) return;
    acc = sm[t];
    G1Affine r = acc.is_inf() ? G1Affine::inf() : g1_to_affine_with_inv(acc, fe_inv_safegcd<FpParams>(acc.zzz));
    uint8_t o[48];
    g1_compress(r, o);
    uint8_t* dst = out + (t ? 80 : 0);
    for (int k = 0; k < 48; k++) dst[k] = o[k];
}

// generator comb tables, one per device, built once and published only after the build has finished on the
// device: contexts on several GPUs (or several contexts on one GPU) may run setup concurrently
constexpr int kMaxDevices = 64;
G1Affine* g_gen_table[kMaxDevices] = {};
std::mutex g_gen_table_mu;

}  // namespace

void g1_decompress_batch(const uint8_t* d_in48, G1Affine* d_out, size_t n, int* d_err, cudaStream_t st) {
    if (!n) return;
    (++fb::g_launches, k_decompress<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_in48, d_out, n, d_err));
    FB_CUDA(cudaGetLastError());
}
void g1_deserialize_batch(const uint8_t* d_in96, G1Affine* d_out, size_t n, int* d_err, cudaStream_t st) {
    if (!n) return;
    (++fb::g_launches, k_deserialize<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_in96, d_out, n, d_err));
    FB_CUDA(cudaGetLastError());
}
void g1_compress_batch(const G1Affine* d_in, uint8_t* d_out48, size_t n, cudaStream_t st) {
    if (!n) return;
    (++fb::g_launches, k_compress<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_in, d_out48, n));
    FB_CUDA(cudaGetLastError());
}
void g1_serialize_batch(const G1Affine* d_in, uint8_t* d_out96, size_t n, cudaStream_t st) {
    if (!n) return;
    (++fb::g_launches, k_serialize<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_in, d_out96, n));
    FB_CUDA(cudaGetLastError());
}
void g1_xyzz_to_affine_batch(const G1Xyzz* d_in, G1Affine* d_out, size_t n, cudaStream_t st) {
    if (!n) return;
    (++fb::g_launches, k_to_affine<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_in, d_out, n));
    FB_CUDA(cudaGetLastError());
}

void g1_combine_parts(const uint8_t* d_parts224, size_t nparts, uint8_t* d_out128, int* d_err, cudaStream_t st) {
    (++fb::g_launches, k_combine_parts<<<1, 32, 0, st>>>(d_parts224, nparts, d_out128, d_err));
    FB_CUDA(cudaGetLastError());
}

void g1_fixed_base_mul(const Fr* d_scalars, size_t n, G1Affine* d_out, DevBuf& scratch, cudaStream_t st) {
    if (!n) return;
    int dev;
    FB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) throw CudaError{"device ordinal out of range"};
    const G1Affine* tab;
    {
        std::lock_guard<std::mutex> lk(g_gen_table_mu);
        if (!g_gen_table[dev]) {
            G1Xyzz* tmp = scratch.as<G1Xyzz>(32 * 255);
            G1Affine* t;
            FB_CUDA(cudaMalloc(&t, 32 * 255 * sizeof(G1Affine)));
            (++fb::g_launches, k_gen_table_xyzz<<<1, 32, 0, st>>>(tmp));
            (++fb::g_launches, k_to_affine<<<(32 * 255 + 127) / 128, 128, 0, st>>>(tmp, t, 32 * 255));
            FB_CUDA(cudaStreamSynchronize(st));  // other streams of this device read it from now on
            g_gen_table[dev] = t;
        }
        tab = g_gen_table[dev];
    }
    (++fb::g_launches, k_fixed_base_mul<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_scalars, n, tab, d_out));
    FB_CUDA(cudaGetLastError());
}

void g1_sum_affine(const G1Affine* d_in, size_t n, G1Xyzz* d_out, DevBuf& scratch, cudaStream_t st) {
    if (n == 0) {
        FB_CUDA(cudaMemsetAsync(d_out, 0, sizeof(G1Xyzz), st));
        return;
    }
    size_t n1 = (n + 7) / 8;
    G1Xyzz* buf = scratch.as<G1Xyzz>(2 * n1 + 8);
    G1Xyzz* a = buf;
    G1Xyzz* b = buf + n1 + 4;
    (++fb::g_launches, k_sum_affine<<<(unsigned)((n1 + 127) / 128), 128, 0, st>>>(d_in, n, n1 == 1 ? d_out : a, n1));
    size_t m = n1;
    while (m > 1) {
        size_t mo = (m + 3) / 4;
        (++fb::g_launches, k_sum_xyzz<<<(unsigned)((mo + 127) / 128), 128, 0, st>>>(a, m, mo == 1 ? d_out : b, mo));
        G1Xyzz* t = a;
        a = b;
        b = t;
        m = mo;
    }
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
