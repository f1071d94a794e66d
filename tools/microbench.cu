// microbench.cu - integer-pipe issue rates and field/curve throughput on the
// device, plus a self-check of the carry-chain multiplier against the portable
// one.  Prints one JSON object per line.  Build: make tools/microbench
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "../fourier_b200/csrc/g1.cuh"

using namespace fb;

#define CK(x)                                                                     \
    do {                                                                          \
        cudaError_t e = (x);                                                      \
        if (e != cudaSuccess) {                                                   \
            printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); \
            exit(1);                                                              \
        }                                                                         \
    } while (0)

constexpr int ITER = 4096;

// ILP independent chains of IMAD (32-bit)
template <int ILP>
__global__ void k_imad(uint32_t* out, uint32_t a, uint32_t b) {
    uint32_t x[ILP];
    for (int i = 0; i < ILP; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) x[i] = x[i] * a + b;
    }
    uint32_t s = 0;
    for (int i = 0; i < ILP; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// ILP independent chains of IMAD.WIDE.U32 (64-bit accumulate)
template <int ILP>
__global__ void k_imad_wide(uint64_t* out, uint32_t a, uint32_t b) {
    uint64_t x[ILP];
    for (int i = 0; i < ILP; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            uint32_t lo = (uint32_t)x[i] ^ b;
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[i]) : "r"(lo), "r"(a));
        }
    }
    uint64_t s = 0;
    for (int i = 0; i < ILP; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// carry chains: ILP independent chains, each a run of 8 lo/hi MAD pairs
template <int ILP>
__global__ void k_imad_chain(uint32_t* out, uint32_t a, uint32_t b) {
    uint32_t x[ILP][16];
    for (int i = 0; i < ILP; i++)
        for (int j = 0; j < 16; j++) x[i][j] = threadIdx.x + i + j;
    for (int it = 0; it < ITER / 8; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            x[i][0] = mad_lo_cc(a, b + i, x[i][0]);
            x[i][1] = madc_hi_cc(a, b + i, x[i][1]);
#pragma unroll
            for (int j = 2; j < 16; j += 2) {
                x[i][j] = madc_lo_cc(a + j, b + i, x[i][j]);
                x[i][j + 1] = madc_hi_cc(a + j, b + i, x[i][j + 1]);
            }
        }
    }
    uint32_t s = 0;
    for (int i = 0; i < ILP; i++)
        for (int j = 0; j < 16; j++) s += x[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class P, int VARIANT, int ILP>
__global__ void k_mul_loop(Fe<P>* data, int iters) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    Fe<P> x[ILP], y = data[t];
    for (int i = 0; i < ILP; i++) {
        x[i] = y;
        x[i].v[0] ^= i;
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) x[i] = VARIANT == 0 ? fe_mul<P>(x[i], y) : fe_mul_portable<P>(x[i], y);
    }
    Fe<P> s = x[0];
    for (int i = 1; i < ILP; i++) s = fe_add<P>(s, x[i]);
    data[t] = s;
}

__global__ void k_madd_loop(G1Affine* pts, G1Xyzz* out, int iters, int npts) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    G1Xyzz acc = G1Xyzz::inf();
    for (int it = 0; it < iters; it++) g1_add_mixed(acc, pts[(t + it * 37) % npts]);
    out[t] = acc;
}

// latency of ONE binary-GCD inversion on ONE lane per warp (how the batched-affine rounds and the
// evaluation-form opening use it): every warp inverts its own element, 20 warps per SM
template <class P, int ALG = 0>  // ALG 0: fe_inv_gcd (shift-and-subtract, ALU pipe), 1: fe_inv_safegcd (division-step batches, wide MADs)
__global__ void k_inv_lane0(Fe<P>* data) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if ((threadIdx.x & 31) == 0) data[warp] = ALG ? fe_inv_safegcd<P>(data[warp]) : fe_inv_gcd<P>(data[warp]);
}

template <class P>
__global__ void k_check(const Fe<P>* a, const Fe<P>* b, int n, int* bad) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    Fe<P> x = fe_mul<P>(a[t], b[t]), y = fe_mul_portable<P>(a[t], b[t]);
    if (x != y) atomicAdd(bad, 1);
    // algebra: (a+b)^2 == a^2 + 2ab + b^2, a*inv... skipped (slow); sub/add inverse
    Fe<P> s = fe_add<P>(a[t], b[t]);
    Fe<P> l = fe_sqr<P>(s);
    Fe<P> r = fe_add<P>(fe_add<P>(fe_sqr<P>(a[t]), fe_sqr<P>(b[t])), fe_dbl<P>(x));
    if (l != r) atomicAdd(bad, 1);
    if (fe_sub<P>(s, b[t]) != a[t]) atomicAdd(bad, 1);
}

template <class F>
float time_ms(F f, int reps = 3) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0));
        f();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

template <class P>
void fill_random(Fe<P>* h, int n, uint64_t seed) {
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < P::N; j++) {
            seed = seed * 6364136223846793005ULL + 1442695040888963407ULL;
            h[i].v[j] = (uint32_t)(seed >> 32);
        }
        h[i].v[P::N - 1] &= (P::N == 8 ? 0x3fffffffu : 0x0fffffffu);  // < modulus
    }
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"max_clock_mhz\": %d}\n", prop.name, sms, clk_khz / 1000);
    const int threads = 256, blocks = sms * 8;
    void* buf;
    CK(cudaMalloc(&buf, (size_t)threads * blocks * 8));
    double total_threads = (double)threads * blocks;
    {
        float ms;
        double ops;
#define ONE(name, call, ilp, per)                                                                                  \
    ms = time_ms([&] { call; });                                                                                   \
    ops = total_threads * ITER * (double)(ilp) * (per);                                                            \
    printf("{\"bench\": \"%s\", \"ilp\": %d, \"ms\": %.4f, \"gops\": %.1f, \"ops_per_clk_per_sm_at_max\": %.2f}\n", name, \
           ilp, ms, ops / ms / 1e6, ops / (ms * 1e-3) / sms / (clk_khz * 1e3));
        ONE("imad32", (k_imad<1><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 1, 1)
        ONE("imad32", (k_imad<4><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 4, 1)
        ONE("imad32", (k_imad<8><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 8, 1)
        ONE("imad_wide", (k_imad_wide<1><<<blocks, threads>>>((uint64_t*)buf, 0x9e3779b1u, 12345u)), 1, 1)
        ONE("imad_wide", (k_imad_wide<4><<<blocks, threads>>>((uint64_t*)buf, 0x9e3779b1u, 12345u)), 4, 1)
        ONE("imad_wide", (k_imad_wide<8><<<blocks, threads>>>((uint64_t*)buf, 0x9e3779b1u, 12345u)), 8, 1)
        // chain kernels: per outer iteration (ITER/8 of them) each chain issues 8 wide MADs => ITER wide MADs per chain
        ONE("imad_wide_carry_chain", (k_imad_chain<1><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 1, 1)
        ONE("imad_wide_carry_chain", (k_imad_chain<2><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 2, 1)
        ONE("imad_wide_carry_chain", (k_imad_chain<4><<<blocks, threads>>>((uint32_t*)buf, 0x9e3779b1u, 12345u)), 4, 1)
    }
    // ---- field multiply correctness + throughput
    const int N = 1 << 20;
    {
        Fp *ha = (Fp*)malloc(sizeof(Fp) * N), *hb = (Fp*)malloc(sizeof(Fp) * N), *da, *db;
        fill_random<FpParams>(ha, N, 1);
        fill_random<FpParams>(hb, N, 2);
        for (int i = 0; i < 12; i++) { ha[0].v[i] = FpParams::mod(i); hb[0].v[i] = FpParams::mod(i); }
        ha[0].v[0] -= 1; hb[0].v[0] -= 1;  // (p-1)*(p-1)
        CK(cudaMalloc(&da, sizeof(Fp) * N));
        CK(cudaMalloc(&db, sizeof(Fp) * N));
        CK(cudaMemcpy(da, ha, sizeof(Fp) * N, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(db, hb, sizeof(Fp) * N, cudaMemcpyHostToDevice));
        int* dbad;
        CK(cudaMalloc(&dbad, 4));
        CK(cudaMemset(dbad, 0, 4));
        k_check<FpParams><<<N / 256, 256>>>(da, db, N, dbad);
        int bad;
        CK(cudaMemcpy(&bad, dbad, 4, cudaMemcpyDeviceToHost));
        printf("{\"check\": \"fp_mul chain vs portable + identities\", \"n\": %d, \"mismatches\": %d}\n", N, bad);
        CK(cudaMemset(dbad, 0, 4));
        {
            Fr *hra = (Fr*)malloc(sizeof(Fr) * N), *hrb = (Fr*)malloc(sizeof(Fr) * N);
            fill_random<FrParams>(hra, N, 3);
            fill_random<FrParams>(hrb, N, 4);
            CK(cudaMemcpy(da, hra, sizeof(Fr) * N, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(db, hrb, sizeof(Fr) * N, cudaMemcpyHostToDevice));
            free(hra);
            free(hrb);
        }
        k_check<FrParams><<<N / 256, 256>>>((Fr*)da, (Fr*)db, N, dbad);
        CK(cudaMemcpy(&bad, dbad, 4, cudaMemcpyDeviceToHost));
        printf("{\"check\": \"fr_mul chain vs portable + identities\", \"n\": %d, \"mismatches\": %d}\n", N, bad);
        CK(cudaMemcpy(da, ha, sizeof(Fp) * N, cudaMemcpyHostToDevice));
        const int iters = 256;
        int nthreads = sms * 2048;  // full occupancy worth of threads
        if (nthreads > N) nthreads = N;
#define MULB(name, P, variant, ilp, tpb)                                                                           \
    {                                                                                                              \
        float ms = time_ms([&] { k_mul_loop<P, variant, ilp><<<nthreads / tpb, tpb>>>((Fe<P>*)da, iters); });      \
        double muls = (double)nthreads * iters * ilp;                                                              \
        printf("{\"bench\": \"%s\", \"variant\": \"%s\", \"ilp\": %d, \"ms\": %.3f, \"gmul_per_s\": %.2f}\n", name, \
               variant == 0 ? "carry_chain" : "portable_u64", ilp, ms, muls / ms / 1e6);                           \
    }
        MULB("fp_mul", FpParams, 0, 1, 128)
        MULB("fp_mul", FpParams, 0, 2, 128)
        MULB("fp_mul", FpParams, 1, 1, 128)
        MULB("fr_mul", FrParams, 0, 1, 128)
        MULB("fr_mul", FrParams, 0, 2, 128)
        MULB("fr_mul", FrParams, 1, 1, 128)
        // rate of dependent fe_mul<Fp> chains against warps per SM sub-partition (one CTA per SM of 4*k warps): where the
        // multiplier saturates, and whether two independent chains per thread (ilp 2) reach it with fewer warps
#define OCC(ilp, k)                                                                                                 \
    {                                                                                                              \
        float ms = time_ms([&] { k_mul_loop<FpParams, 0, ilp><<<sms, 128 * k>>>((Fp*)da, iters); });               \
        double muls = (double)sms * 128 * k * iters * ilp;                                                         \
        printf("{\"bench\": \"fp_mul_occupancy\", \"warps_per_smsp\": %d, \"ilp\": %d, \"ms\": %.3f, \"gmul_per_s\": %.2f}\n", k, ilp, \
               ms, muls / ms / 1e6);                                                                               \
    }
        OCC(1, 1) OCC(1, 2) OCC(1, 3) OCC(1, 4) OCC(1, 6) OCC(1, 8)
        OCC(2, 1) OCC(2, 2) OCC(2, 3) OCC(2, 4)
        OCC(4, 1) OCC(4, 2)
        // inversion latency: 148 x 5 CTAs x 4 warps, one lane each
        {
            const int nw = sms * 20;
            Fp* dinv;
            CK(cudaMalloc(&dinv, sizeof(Fp) * nw));
            CK(cudaMemcpy(dinv, da, sizeof(Fp) * nw, cudaMemcpyDeviceToDevice));
            float ms = time_ms([&] { k_inv_lane0<FpParams><<<sms * 5, 128>>>(dinv); });
            printf("{\"bench\": \"fe_inv_gcd<Fp> one lane per warp, 20 warps/SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FrParams><<<sms * 5, 128>>>((Fr*)dinv); });
            printf("{\"bench\": \"fe_inv_gcd<Fr> one lane per warp, 20 warps/SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FpParams, 1><<<sms * 5, 128>>>(dinv); });
            printf("{\"bench\": \"fe_inv_safegcd<Fp> one lane per warp, 20 warps/SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FrParams, 1><<<sms * 5, 128>>>((Fr*)dinv); });
            printf("{\"bench\": \"fe_inv_safegcd<Fr> one lane per warp, 20 warps/SM\", \"ms\": %.4f}\n", ms);
            // one warp per SM: the latency of a lone inversion (how the end-of-call steps see it)
            ms = time_ms([&] { k_inv_lane0<FpParams, 0><<<sms, 32>>>(dinv); });
            printf("{\"bench\": \"fe_inv_gcd<Fp> one lane, one warp per SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FpParams, 1><<<sms, 32>>>(dinv); });
            printf("{\"bench\": \"fe_inv_safegcd<Fp> one lane, one warp per SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FrParams, 0><<<sms, 32>>>((Fr*)dinv); });
            printf("{\"bench\": \"fe_inv_gcd<Fr> one lane, one warp per SM\", \"ms\": %.4f}\n", ms);
            ms = time_ms([&] { k_inv_lane0<FrParams, 1><<<sms, 32>>>((Fr*)dinv); });
            printf("{\"bench\": \"fe_inv_safegcd<Fr> one lane, one warp per SM\", \"ms\": %.4f}\n", ms);
            cudaFree(dinv);
        }
        // mixed additions
        G1Affine* pts;
        G1Xyzz* out;
        const int npts = 4096;
        CK(cudaMalloc(&pts, sizeof(G1Affine) * npts));
        CK(cudaMalloc(&out, sizeof(G1Xyzz) * nthreads));
        // any field values work for timing (the formulas do not branch on curve membership)
        CK(cudaMemcpy(pts, da, sizeof(G1Affine) * npts, cudaMemcpyDeviceToDevice));
        for (int tpb : {64, 128}) {
            int nt = sms * 1024;
            float ms = time_ms([&] { k_madd_loop<<<nt / tpb, tpb>>>(pts, out, 64, npts); });
            printf("{\"bench\": \"g1_add_mixed\", \"threads\": %d, \"tpb\": %d, \"ms\": %.3f, \"madd_per_s\": %.3e}\n", nt, tpb, ms,
                   (double)nt * 64 / (ms * 1e-3));
        }
    }
    return 0;
}
