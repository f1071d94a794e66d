// batchinv.cuh - one field inversion per CTA shared by all its threads (Montgomery's trick).
//
// The batched-affine MSM rounds (msm.cu) and the evaluation-form opening (evalform.cu) need 1/t for a
// non-zero t held by every thread.  Lanes combine by shuffled prefix/suffix products, the NW warp totals
// meet in shared memory, and ONE thread runs the binary-GCD inversion (fe_inv_gcd: data-dependent
// branches, so it runs on one lane whatever the batching).  Measured on B200 (tools/microbench,
// profiles/microbench_r1.jsonl): with one inverting lane in each of 20 resident warps per SM the
// schedulers time-slice the 20 instruction streams and an Fp inversion takes 176 us; with one per CTA
// the stream count drops by NW and the inversion runs near its single-warp latency.
#pragma once
#include "field.cuh"

namespace fb {

template <class P>
__device__ __forceinline__ Fe<P> shfl_fe(const Fe<P>& v, int src) {
    Fe<P> r;
#pragma unroll
    for (int i = 0; i < P::N; i++) r.v[i] = __shfl_sync(0xffffffffu, v.v[i], src);
    return r;
}

// v[0..NW) -> their inverses, on one thread: product tree up, one inversion, tree down (3(NW-1) multiplies,
// dependent depth 2 log2 NW)
template <class P, int NW>
__device__ __forceinline__ void tree_inverse(Fe<P>* v) {
    if constexpr (NW == 1) {
        v[0] = fe_inv_gcd<P>(v[0]);
    } else if constexpr (NW % 2 != 0 || (NW & (NW - 1)) != 0) {
        // any count: running products up, one inversion, peel back (3(NW-1) dependent multiplies)
        Fe<P> pre[NW];
        pre[0] = v[0];
#pragma unroll
        for (int i = 1; i < NW; i++) pre[i] = fe_mul<P>(pre[i - 1], v[i]);
        Fe<P> inv = fe_inv_gcd<P>(pre[NW - 1]);
#pragma unroll
        for (int i = NW - 1; i >= 1; i--) {
            const Fe<P> vi = fe_mul<P>(inv, pre[i - 1]);
            inv = fe_mul<P>(inv, v[i]);
            v[i] = vi;
        }
        v[0] = inv;
    } else {
        Fe<P> prod[NW / 2];
#pragma unroll
        for (int i = 0; i < NW / 2; i++) prod[i] = fe_mul<P>(v[2 * i], v[2 * i + 1]);
        tree_inverse<P, NW / 2>(prod);
#pragma unroll
        for (int i = 0; i < NW / 2; i++) {
            const Fe<P> a = fe_mul<P>(prod[i], v[2 * i + 1]);
            const Fe<P> b = fe_mul<P>(prod[i], v[2 * i]);
            v[2 * i] = a;
            v[2 * i + 1] = b;
        }
    }
}

// Every thread of a CTA of NW full warps calls this with a non-zero t and gets 1/t.  Contains two
// __syncthreads(): no thread of the CTA may have exited before the call.
template <class P, int NW>
__device__ __forceinline__ Fe<P> block_batch_inverse(const Fe<P>& t) {
    __shared__ Fe<P> s_tot[NW];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Fe<P> pre = t, suf = t;  // inclusive prefix / suffix products over the lanes
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const Fe<P> up = shfl_fe<P>(pre, lane - d < 0 ? lane : lane - d);
        const Fe<P> dn = shfl_fe<P>(suf, lane + d > 31 ? lane : lane + d);
        if (lane >= d) pre = fe_mul<P>(pre, up);
        if (lane + d <= 31) suf = fe_mul<P>(suf, dn);
    }
    if (lane == 31) s_tot[warp] = pre;
    __syncthreads();
    if (threadIdx.x == 0) {
        Fe<P> v[NW];
#pragma unroll
        for (int i = 0; i < NW; i++) v[i] = s_tot[i];
        tree_inverse<P, NW>(v);
#pragma unroll
        for (int i = 0; i < NW; i++) s_tot[i] = v[i];
    }
    __syncthreads();
    Fe<P> r = s_tot[warp];                                              // 1 / (warp total)
    const Fe<P> ex_pre = shfl_fe<P>(pre, lane == 0 ? 0 : lane - 1);    // product of the lanes below
    const Fe<P> ex_suf = shfl_fe<P>(suf, lane == 31 ? 31 : lane + 1);  // product of the lanes above
    if (lane > 0) r = fe_mul<P>(r, ex_pre);
    if (lane < 31) r = fe_mul<P>(r, ex_suf);
    return r;
}

}  // namespace fb
