// vgs_device.cuh -- device helpers shared by the kernel files (arithmetic that must match Go's int32 /
// float32 semantics, tile loads, publication of per-pod candidates).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vgs_layout.h"

namespace vgs {

// ------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------
__device__ __forceinline__ int32_t wrap_sub(int32_t a, int32_t b) {
    return (int32_t)((uint32_t)a - (uint32_t)b);
}
__device__ __forceinline__ int32_t wrap_add(int32_t a, int32_t b) {
    return (int32_t)((uint32_t)a + (uint32_t)b);
}
// Go: Totalmem * MemPercentagereq / 100 in int32 (score.go:119): wrapping multiply,
// then division truncating toward zero.
__device__ __forceinline__ int32_t pct_mem(int32_t totalmem, int32_t pct) {
    return (int32_t)((uint32_t)totalmem * (uint32_t)pct) / 100;
}
// score.go:180  float32(total)/float32(free) + float32(len(node.Devices)-sums)
__device__ __forceinline__ float container_score(int32_t total, int32_t free_, int32_t rest) {
    return __fadd_rn(__fdiv_rn(__int2float_rn(total), __int2float_rn(free_)),
                     __int2float_rn(rest));
}
__device__ __forceinline__ bool type_ok(uint32_t lo, uint32_t hi, uint32_t meta) {
    uint32_t t = meta & M_TYPE_MASK;
    uint32_t w = (t & 32u) ? hi : lo;
    return (w >> (t & 31u)) & 1u;
}

__device__ __forceinline__ DevReq load_req(const DevReq *__restrict__ p) {
    const int4 *q = reinterpret_cast<const int4 *>(p);
    int4 a = __ldg(q), b = __ldg(q + 1);
    DevReq r;
    r.nums = a.x;
    r.memfix = a.y;
    r.pct = a.z;
    r.coresreq = a.w;
    r.blockmask = (uint32_t)b.x;
    r.flags = (uint32_t)b.y;
    r.mask_lo = (uint32_t)b.z;
    r.mask_hi = (uint32_t)b.w;
    return r;
}

template <int DM>
__device__ __forceinline__ void load_word(const int4 *__restrict__ tile, int word, int lane,
                                          int32_t (&dst)[DM]) {
#pragma unroll
    for (int q = 0; q < DM / 4; q++) {
        int4 x = __ldg(&tile[(word * (DM / 4) + q) * kWarp + lane]);
        dst[4 * q + 0] = x.x;
        dst[4 * q + 1] = x.y;
        dst[4 * q + 2] = x.z;
        dst[4 * q + 3] = x.w;
    }
}

// ------------------------------------------------------------------------------------
// warp-level publication of one pod's candidates: lane k keeps the result of the k-th
// pod of a group of 32 so that the flush is one coalesced atomic per 32 pods.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ void publish(bool fit, float score, int k, int lane, uint32_t row0,
                                        unsigned long long &mykey, int32_t &mycnt) {
    unsigned bal = __ballot_sync(0xffffffffu, fit);
    if (bal) {
        uint32_t sb = fit ? order_bits(__float_as_uint(score)) : 0u;
        uint32_t m = __reduce_max_sync(0xffffffffu, sb);
        unsigned who = __ballot_sync(0xffffffffu, sb == m);
        uint32_t hl = 31u - (uint32_t)__clz((int)who); // ties -> highest row
        if (lane == k) {
            mykey = ((unsigned long long)m << 32) | (unsigned long long)(row0 + hl);
            mycnt = __popc(bal);
        }
    }
}

// Small batches (one pod chunk of <= 32 pods, e.g. the single pod of a Filter call): every warp
// of every item would hit the same few keys[] words with atomics -- 31250 same-address pairs at
// 1M nodes.  Instead lane k keeps pod k's candidate across all the CTA's items and the CTA
// publishes once.
template <int WARPS>
__device__ __forceinline__ void flush_small(unsigned long long acckey, int32_t acccnt, int lane, int warp,
                                            int n_pods, const int32_t *__restrict__ orig_or_null,
                                            const GenPod *__restrict__ pods_or_null,
                                            unsigned long long *__restrict__ keys, int32_t *__restrict__ counts) {
    __shared__ unsigned long long s_k[WARPS][kWarp];
    __shared__ int32_t s_c[WARPS][kWarp];
    s_k[warp][lane] = acckey;
    s_c[warp][lane] = acccnt;
    __syncthreads();
    if (warp == 0 && lane < n_pods) {
        unsigned long long kk = 0ull;
        int32_t cc = 0;
        const int nw = (int)(blockDim.x >> 5); // <= WARPS: small snapshots run fewer warps per CTA
        for (int w = 0; w < nw; w++) {
            kk = s_k[w][lane] > kk ? s_k[w][lane] : kk;
            cc += s_c[w][lane];
        }
        if (kk != 0ull) {
            const int op = orig_or_null ? orig_or_null[lane] : pods_or_null[lane].orig;
            atomicMax(&keys[op], kk);
            atomicAdd(&counts[op], cc);
        }
    }
}

} // namespace vgs
