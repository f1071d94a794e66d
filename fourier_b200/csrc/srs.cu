// srs.cu - scalar side of the trusted-setup generator.
//
// Replaces the field arithmetic of generate_trusted_setup (reference
// src/engine/piano.rs:883-1037): tau = hash_to_bls_field(secret) (:890-891),
// powers tau^k (:916-949), R_i(tau_Y) and L_j(tau_X) (:956-988) and the products
// R_i * L_j (:995-1007).  The reference gets each Lagrange value by an inverse
// FFT of a unit vector plus Horner (O(T^2 log T) in total); here the closed form
//     L_j(tau) = w^j (tau^T - 1) / (T (tau - w^j))
// gives the same field element in O(1) per entry.  The group side (g^scalar) is
// g1_fixed_base_mul in g1ops.cu.
#include "common.cuh"

namespace fb {

namespace {

__constant__ uint32_t kRoot32F[8] = {0x5f0e466au, 0xb9b58d8cu, 0x1819d7ecu, 0x5b1b4c80u,
                                     0x52a31e64u, 0x0af53ae3u, 0x19e9b27bu, 0x5bf3addau};

__device__ __forceinline__ Fr pow_u64(Fr base, uint64_t e) {
    Fr r = Fr::one();
    while (e) {
        if (e & 1) r = fe_mul<FrParams>(r, base);
        e >>= 1;
        if (e) base = fe_sqr<FrParams>(base);
    }
    return r;
}

// 32 bytes big-endian -> integer mod r -> Montgomery
__global__ void k_secret_to_fr(const uint8_t* in, Fr* out) {
    if (threadIdx.x || blockIdx.x) return;
    Fr x;
    for (int i = 0; i < 8; i++) {
        const uint8_t* b = in + 4 * (7 - i);
        x.v[i] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    Fr m = Fr::modulus();
    for (int rep = 0; rep < 3; rep++) {
        if (fe_geq<FrParams>(x.v, m.v)) {
            x.v[0] = sub_cc(x.v[0], m.v[0]);
            for (int i = 1; i < 7; i++) x.v[i] = subc_cc(x.v[i], m.v[i]);
            x.v[7] = subc(x.v[7], m.v[7]);
        }
    }
    *out = fe_to_mont<FrParams>(x);
}

// hdr[0] = tau^N - 1, hdr[1] = 1/N, hdr[2] = w_N
__global__ void k_lagrange_hdr(const Fr* tau, int log_n, Fr* hdr) {
    if (threadIdx.x || blockIdx.x) return;
    Fr t = *tau;
    for (int i = 0; i < log_n; i++) t = fe_sqr<FrParams>(t);
    hdr[0] = fe_sub<FrParams>(t, Fr::one());
    Fr two = fe_dbl<FrParams>(Fr::one());
    Fr n = Fr::one();
    for (int i = 0; i < log_n; i++) n = fe_mul<FrParams>(n, two);
    hdr[1] = fe_inv<FrParams>(n);
    Fr w;
    for (int i = 0; i < 8; i++) w.v[i] = kRoot32F[i];
    for (int i = log_n; i < 32; i++) w = fe_sqr<FrParams>(w);
    hdr[2] = w;
}

__global__ void __launch_bounds__(128) k_lagrange_at(const Fr* tau, const Fr* hdr, size_t n, Fr* out) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    Fr wj = pow_u64(hdr[2], j);
    Fr d = fe_sub<FrParams>(*tau, wj);
    Fr zn = hdr[0];
    if (zn.is_zero()) {  // tau lies in the domain: L_j(tau) is a Kronecker delta
        out[j] = d.is_zero() ? Fr::one() : Fr::zero();
        return;
    }
    Fr v = fe_mul<FrParams>(fe_mul<FrParams>(wj, zn), hdr[1]);
    out[j] = fe_mul<FrParams>(v, fe_inv<FrParams>(d));
}

__global__ void __launch_bounds__(128) k_powers(const Fr* tau, size_t n, Fr* out) {
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    out[k] = pow_u64(*tau, k);
}

__global__ void __launch_bounds__(256) k_outer(const Fr* __restrict__ r, size_t nrows, const Fr* __restrict__ l, size_t T,
                                               Fr* __restrict__ out) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows * T) return;
    size_t i = t / T, j = t - i * T;
    out[t] = fe_mul<FrParams>(r[i], l[j]);
}

}  // namespace

void srs_secret_to_fr(const uint8_t* d_secret32, Fr* d_out, cudaStream_t st) {
    (++fb::g_launches, k_secret_to_fr<<<1, 1, 0, st>>>(d_secret32, d_out));
    FB_CUDA(cudaGetLastError());
}

void srs_lagrange_at(const Fr* d_tau, int log_n, Fr* d_out, Fr* d_hdr3, cudaStream_t st) {
    size_t n = (size_t)1 << log_n;
    (++fb::g_launches, k_lagrange_hdr<<<1, 1, 0, st>>>(d_tau, log_n, d_hdr3));
    (++fb::g_launches, k_lagrange_at<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_tau, d_hdr3, n, d_out));
    FB_CUDA(cudaGetLastError());
}

void srs_powers(const Fr* d_tau, size_t n, Fr* d_out, cudaStream_t st) {
    (++fb::g_launches, k_powers<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_tau, n, d_out));
    FB_CUDA(cudaGetLastError());
}

void srs_outer(const Fr* d_r, size_t nrows, const Fr* d_l, size_t T, Fr* d_out, cudaStream_t st) {
    size_t n = nrows * T;
    (++fb::g_launches, k_outer<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_r, nrows, d_l, T, d_out));
    FB_CUDA(cudaGetLastError());
}

}  // namespace fb
