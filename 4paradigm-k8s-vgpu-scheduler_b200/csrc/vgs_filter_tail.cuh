// vgs_filter_tail.cuh -- the tail of a single-pod score kernel launched as the whole Filter call: the CTA that
// finishes last resolves the winner (one warp, vgs_resolve.cuh), optionally commits it (sequential drain), and
// stores result, picks and the status word into the device arrays AND the pinned host mirrors.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"
#include "vgs_prepare.cuh"
#include "vgs_resolve.cuh"

namespace vgs {

// every thread of the CTA calls it after the CTA's candidates have been published (keys[0] / counts[0])
template <int DM>
__device__ __forceinline__ void filter_tail(const FilterOne &fo, unsigned long long *__restrict__ keys,
                                            int32_t *__restrict__ counts, int32_t n_nodes, int32_t row_base) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        // last CTA out resolves the winner (thread 0 published this CTA's candidate just above)
        __shared__ bool s_last;
        if (threadIdx.x == 0) {
            __threadfence();
            s_last = atomicAdd(fo.done, 1u) == gridDim.x - 1u;
        }
        __syncthreads();
        if (!s_last || warp != 0) return;
        // warp 0 of the last CTA: lane 0 reads the candidate, the warp resolves it (lane = device slot)
        unsigned long long key = 0ull;
        vgs_result res;
        res.fit_count = 0;
        if (lane == 0) {
            *fo.done = 0u; // zero at rest: the next call counts again
            __threadfence();
            key = atomicMax(&keys[0], 0ull);
            res.fit_count = atomicAdd(&counts[0], 0);
            keys[0] = 0ull; // the candidate block goes back to its resting state
            counts[0] = 0;
        }
        key = __shfl_sync(0xffffffffu, key, 0);
        const int32_t pick_lo = fo.pick_begin[0], pick_hi = fo.pick_begin[fo.pod.n_reqs];
        auto finish = [&](bool zero_picks) { // lane 0 only
            if (zero_picks) {
                vgs_pick z;
                z.slot = z.usedmem = z.usedcores = 0;
                for (int32_t i = pick_lo; i < pick_hi; i++) fo.picks[i] = z, fo.host_picks[i] = z;
            }
            fo.results[0] = res;
            fo.host_results[0] = res;
            __threadfence();
            fo.host_status[0] = *reinterpret_cast<volatile uint32_t *>(fo.status);
        };
        const uint32_t kbits = unorder_bits((uint32_t)(key >> 32));
        const int32_t row = (int32_t)(uint32_t)key - row_base;
        if (key == 0ull || row < 0 || row >= n_nodes) { // nobody fits (a foreign row cannot happen on one shard)
            if (lane == 0) {
                res.best_row = key == 0ull ? -1 : (int32_t)(uint32_t)key;
                res.score = key == 0ull ? 0.0f : __uint_as_float(kbits);
                res.n_picks = key == 0ull ? 0 : pick_hi - pick_lo;
                finish(true);
            }
            return;
        }
        bool alive;
        float score;
        int32_t nd, used, usedmem, usedcores;
        resolve_row_warp(
            fo.raw, row, fo.dmax, fo.pod.n_reqs, [&](int r) { return fo.reqs[r]; },
            [&](int r, int t, const vgs_pick &pk) {
                fo.picks[fo.pick_begin[r] + t] = pk;
                fo.host_picks[fo.pick_begin[r] + t] = pk;
            },
            alive, score, nd, used, usedmem, usedcores);
        if (!alive || __float_as_uint(score) != kbits) {
            if (lane == 0) atomicOr(fo.status, S_MISMATCH);
        }
        if (fo.commit_walk && alive) { // sequential drain: the placement becomes usage (addPod, scheduler.go:280-297,394)
            if (lane < nd) {
                const size_t at = (size_t)row * fo.dmax + lane;
                const_cast<int32_t *>(fo.raw.p[VGS_PLANE_USED])[at] = used;
                const_cast<int32_t *>(fo.raw.p[VGS_PLANE_USEDMEM])[at] = usedmem;
                const_cast<int32_t *>(fo.raw.p[VGS_PLANE_USEDCORES])[at] = usedcores;
            }
            __syncwarp();
            // ... and the touched row of the walk copy is prepared again
            if (lane == 0) prepare_node<DM>(fo.raw, row, n_nodes, fo.dmax, fo.commit_walk, fo.status);
        }
        __syncwarp();
        if (lane == 0) {
            res.best_row = (int32_t)(uint32_t)key;
            res.score = __uint_as_float(kbits);
            res.n_picks = pick_hi - pick_lo;
            finish(false);
        }
    }
}

} // namespace vgs
