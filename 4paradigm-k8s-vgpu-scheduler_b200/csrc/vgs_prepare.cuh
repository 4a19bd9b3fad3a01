// vgs_prepare.cuh -- prepare_node: one node of the raw planes -> its slot of the walk-ordered tile copy (K0's body;
// also run by the resolve paths when a drain commits a pod and the touched row has to be prepared again).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vgpuscore.h"
#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"

namespace vgs {

// one node: raw planes -> its slot of the walk-ordered tile (also called by K3 when a drain commits a pod)
template <int DM>
// stage: nullptr = store straight into the tile (single rows: deltas, drain commits); else a shared-memory
// block of kWalkWords * DM * 32 int32 owned by the calling WARP, which must be working on one whole tile
// (lane = node): the words are collected there and leave as 512-byte warp stores
__device__ __forceinline__ void prepare_node(const RawPlanes &raw, int row, int32_t n_nodes, int32_t dmax,
                                             int4 *__restrict__ walk, uint32_t *__restrict__ status,
                                             int32_t *__restrict__ stage = nullptr) {
    const bool valid = row < n_nodes;
    int32_t nd = valid ? raw.ndev[row] : 0;
    uint32_t bad = 0;
    if (nd < 0 || nd > dmax) {
        bad |= S_NDEV;
        nd = 0;
    }
    int32_t used[DM], count[DM], usedmem[DM], totalmem[DM], usedcores[DM], totalcore[DM],
        numa[DM], type[DM];
    // the slots are loaded for j < dmax, not j < ndev (ndev's own load would sit in front of all of them),
    // as two 128-bit loads per plane where the rows are 32 bytes, and masked afterwards
    int32_t *const dst[VGS_NUM_PLANES] = {used, count, usedmem, totalmem, usedcores, totalcore, numa, type};
    unsigned long long all_ptrs = 0ull; // caller-owned planes (vgs_score_batch_noalloc) may be 4-byte aligned only
#pragma unroll
    for (int f = 0; f < VGS_NUM_PLANES; f++) all_ptrs |= reinterpret_cast<unsigned long long>(raw.p[f]);
    if (DM == 8 && dmax == 8 && valid && (all_ptrs & 15ull) == 0ull) {
#pragma unroll
        for (int f = 0; f < VGS_NUM_PLANES; f++) {
            const int4 *q = reinterpret_cast<const int4 *>(raw.p[f] + (size_t)row * 8);
            const int4 lo = q[0], hi = q[1];
            dst[f][0] = lo.x, dst[f][1] = lo.y, dst[f][2] = lo.z, dst[f][3] = lo.w;
            dst[f][4] = hi.x, dst[f][5] = hi.y, dst[f][6] = hi.z, dst[f][7] = hi.w;
        }
    } else {
#pragma unroll
        for (int f = 0; f < VGS_NUM_PLANES; f++)
#pragma unroll
            for (int j = 0; j < DM; j++) dst[f][j] = (valid && j < dmax) ? raw.p[f][(size_t)row * dmax + j] : 0;
    }
#pragma unroll
    for (int j = 0; j < DM; j++) {
        bool on = j < nd;
        if (!on) used[j] = count[j] = usedmem[j] = totalmem[j] = usedcores[j] = totalcore[j] = numa[j] = type[j] = 0;
        if (on) {
            if (used[j] < 0 || used[j] >= VGS_MAX_SHARE || count[j] < 0 || count[j] >= VGS_MAX_SHARE)
                bad |= S_RANGE;
            if (type[j] < 0 || type[j] >= VGS_MAX_TYPES) {
                bad |= S_TYPE;
                type[j] = 0;
            }
            if (totalmem[j] < 0 || totalmem[j] > kSliceMaxTotalmem) bad |= S_BIGMEM;
        }
    }
    if (bad) atomicOr(status, bad);
    {
        int mt = 0;
#pragma unroll
        for (int j = 0; j < DM; j++) mt = max(mt, type[j]);
        if (mt > 0) atomicMax(status + 1, (uint32_t)mt); // status[1] = largest type id in use
    }

    const int tile = row / kWarp, lane = row % kWarp;
    int32_t *wbase = reinterpret_cast<int32_t *>(walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp);
    auto put = [&](int word, int w, int32_t v) {
        if (stage) stage[(word * DM + w) * kWarp + lane] = v; // bank = lane: conflict-free whatever w is
        else wbase[((word * (DM / 4) + (w >> 2)) * kWarp + lane) * 4 + (w & 3)] = v;
    };
#pragma unroll
    for (int j = 0; j < DM; j++) {
        if (j < nd) {
            // ascending position under score.go:45-50 with insertion-sort stability:
            // (Numa, Count-Used, registration slot)
            int rank = 0, nrank = 0;
            int32_t fj = wrap_sub(count[j], used[j]);
#pragma unroll
            for (int i = 0; i < DM; i++) {
                if (i < nd && i != j) {
                    int32_t fi = wrap_sub(count[i], used[i]);
                    bool less = numa[i] < numa[j] ||
                                (numa[i] == numa[j] && (fi < fj || (fi == fj && i < j)));
                    rank += less ? 1 : 0;
                    nrank += numa[i] < numa[j] ? 1 : 0;
                }
            }
            int w = nd - 1 - rank; // the reverse walk (score.go:92) visits the last first
            uint32_t meta = (uint32_t)type[j] | ((uint32_t)nrank << M_NUMA_SHIFT) |
                            ((uint32_t)j << M_SLOT_SHIFT) | ((uint32_t)nd << M_NDEV_SHIFT) | M_ALWAYS;
            if (count[j] <= used[j]) meta |= M_NOFREE;
            if (totalcore[j] == 100) meta |= M_IS100;
            if (totalcore[j] != 0) meta |= M_CORENZ;
            if (totalcore[j] == 100 && used[j] > 0) meta |= M_EXCL;
            if (totalcore[j] != 0 && usedcores[j] == totalcore[j]) meta |= M_FULL;
            if (raw.idle_types >> type[j] & 1ull) meta |= used[j] > 0 ? (M_T370 | M_BUSY370) : M_T370;
            float s1 = fj > 0 ? container_score(count[j], fj, nd - 1) : 0.0f;
            put(W_TOTALMEM, w, totalmem[j]);
            put(W_FREEMEM, w, wrap_sub(totalmem[j], usedmem[j]));
            put(W_FREECORES, w, wrap_sub(totalcore[j], usedcores[j]));
            put(W_COUNT, w, count[j]);
            put(W_FREESLOTS, w, fj);
            put(W_META, w, (int32_t)meta);
            put(W_SCORE1, w, __float_as_int(s1));
            put(W_TOTALCORE, w, totalcore[j]);
        } else {
            uint32_t meta = M_PAD | M_NOFREE | M_ALWAYS | ((uint32_t)j << M_SLOT_SHIFT) |
                            ((uint32_t)nd << M_NDEV_SHIFT);
            put(W_TOTALMEM, j, 0);
            put(W_FREEMEM, j, 0);
            put(W_FREECORES, j, 0);
            put(W_COUNT, j, 0);
            put(W_FREESLOTS, j, 0);
            put(W_META, j, (int32_t)meta);
            put(W_SCORE1, j, 0);
            put(W_TOTALCORE, j, 0);
        }
    }
    if (stage) {
        __syncwarp();
        int4 *tile4 = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
#pragma unroll
        for (int word = 0; word < kWalkWords; word++)
#pragma unroll
            for (int q = 0; q < DM / 4; q++) {
                const int32_t *sp = stage + ((word * DM + 4 * q) * kWarp + lane);
                tile4[(word * (DM / 4) + q) * kWarp + lane] = make_int4(sp[0], sp[kWarp], sp[2 * kWarp], sp[3 * kWarp]);
            }
        __syncwarp();
    }
}

// The same words for ONE row by a whole warp (lane = device slot): the drain's commits, where the row comes back
// into play at once and a single thread's run through prepare_node would sit on the critical path.  mirror: a
// second tile (shared memory, the row at column mirror_lane) that receives the same words; may be null.
// the core: lane j holds the columns of device slot j (nd already clamped to [0, dmax]); walk or mirror may be null
template <int DM>
__device__ __forceinline__ void prepare_lanes(int row, int32_t nd, uint32_t bad, int32_t used, int32_t count,
                                              int32_t usedmem, int32_t totalmem, int32_t usedcores, int32_t totalcore,
                                              int32_t numa, int32_t type, uint64_t idle_types, int4 *__restrict__ walk,
                                              uint32_t *__restrict__ status, int4 *mirror, int mirror_lane) {
    constexpr unsigned kAll = 0xffffffffu;
    const int j = (int)(threadIdx.x & 31u);
    const bool on = j < nd;
    if (on) {
        if (used < 0 || used >= VGS_MAX_SHARE || count < 0 || count >= VGS_MAX_SHARE) bad |= S_RANGE;
        if (type < 0 || type >= VGS_MAX_TYPES) {
            bad |= S_TYPE;
            type = 0;
        }
        if (totalmem < 0 || totalmem > kSliceMaxTotalmem) bad |= S_BIGMEM;
    }
    bad = __reduce_or_sync(kAll, bad);
    const int32_t mt = __reduce_max_sync(kAll, on ? type : 0);
    if (j == 0) {
        if (bad) atomicOr(status, bad);
        if (mt > 0) atomicMax(status + 1, (uint32_t)mt);
    }
    const int32_t fj = wrap_sub(count, used);
    int rank = 0, nrank = 0;
#pragma unroll
    for (int i = 0; i < DM; i++) {
        const int32_t fi = __shfl_sync(kAll, fj, i), ni = __shfl_sync(kAll, numa, i);
        if (i < nd && i != j) {
            const bool less = ni < numa || (ni == numa && (fi < fj || (fi == fj && i < j)));
            rank += less ? 1 : 0;
            nrank += ni < numa ? 1 : 0;
        }
    }
    if (j >= DM) return;
    const int w = on ? nd - 1 - rank : j;
    uint32_t meta;
    int32_t words[kWalkWords];
#pragma unroll
    for (int k = 0; k < kWalkWords; k++) words[k] = 0;
    if (on) {
        meta = (uint32_t)type | ((uint32_t)nrank << M_NUMA_SHIFT) | ((uint32_t)j << M_SLOT_SHIFT) |
               ((uint32_t)nd << M_NDEV_SHIFT) | M_ALWAYS;
        if (count <= used) meta |= M_NOFREE;
        if (totalcore == 100) meta |= M_IS100;
        if (totalcore != 0) meta |= M_CORENZ;
        if (totalcore == 100 && used > 0) meta |= M_EXCL;
        if (totalcore != 0 && usedcores == totalcore) meta |= M_FULL;
        if (idle_types >> type & 1ull) meta |= used > 0 ? (M_T370 | M_BUSY370) : M_T370;
        const float s1 = fj > 0 ? container_score(count, fj, nd - 1) : 0.0f;
        words[W_TOTALMEM] = totalmem;
        words[W_FREEMEM] = wrap_sub(totalmem, usedmem);
        words[W_FREECORES] = wrap_sub(totalcore, usedcores);
        words[W_COUNT] = count;
        words[W_FREESLOTS] = fj;
        words[W_SCORE1] = __float_as_int(s1);
        words[W_TOTALCORE] = totalcore;
    } else {
        meta = M_PAD | M_NOFREE | M_ALWAYS | ((uint32_t)j << M_SLOT_SHIFT) | ((uint32_t)nd << M_NDEV_SHIFT);
    }
    words[W_META] = (int32_t)meta;
    const int tile = row / kWarp, lane = row % kWarp;
    int32_t *wbase = walk ? reinterpret_cast<int32_t *>(walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp) : nullptr;
    int32_t *mbase = reinterpret_cast<int32_t *>(mirror);
#pragma unroll
    for (int k = 0; k < kWalkWords; k++) {
        if (wbase) wbase[((k * (DM / 4) + (w >> 2)) * kWarp + lane) * 4 + (w & 3)] = words[k];
        if (mbase) mbase[((k * (DM / 4) + (w >> 2)) * kWarp + mirror_lane) * 4 + (w & 3)] = words[k];
    }
}

template <int DM>
__device__ __forceinline__ void prepare_node_warp(const RawPlanes &raw, int row, int32_t dmax, int4 *__restrict__ walk,
                                                  uint32_t *__restrict__ status, int4 *mirror, int mirror_lane) {
    const int j = (int)(threadIdx.x & 31u);
    int32_t nd = raw.ndev[row];
    uint32_t bad = 0;
    if (nd < 0 || nd > dmax) {
        bad |= S_NDEV;
        nd = 0;
    }
    const bool on = j < nd;
    const size_t at = (size_t)row * dmax + (j < dmax ? j : 0);
    const int32_t used = on ? raw.p[VGS_PLANE_USED][at] : 0, count = on ? raw.p[VGS_PLANE_COUNT][at] : 0;
    const int32_t usedmem = on ? raw.p[VGS_PLANE_USEDMEM][at] : 0, totalmem = on ? raw.p[VGS_PLANE_TOTALMEM][at] : 0;
    const int32_t usedcores = on ? raw.p[VGS_PLANE_USEDCORES][at] : 0, totalcore = on ? raw.p[VGS_PLANE_TOTALCORE][at] : 0;
    const int32_t numa = on ? raw.p[VGS_PLANE_NUMA][at] : 0, type = on ? raw.p[VGS_PLANE_TYPEID][at] : 0;
    prepare_lanes<DM>(row, nd, bad, used, count, usedmem, totalmem, usedcores, totalcore, numa, type, raw.idle_types, walk,
                      status, mirror, mirror_lane);
}

} // namespace vgs
