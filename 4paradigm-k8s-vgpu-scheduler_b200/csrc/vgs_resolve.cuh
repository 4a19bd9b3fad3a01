// vgs_resolve.cuh -- one (pod, node) pair re-run from the RAW planes in the plain way: stable insertion sort by
// (Numa, Count-Used) before every request (score.go:45-50,165), reverse walk with the six predicates and numa-bind
// (score.go:86-152), usage charged to the taken devices (score.go:168-174), picks emitted in take order.  It is
// the resolve kernel's body (K3: the winner node of every pod, self-check against the score kernels' key) and the
// tail of the single-pod Filter kernel (vgs_general.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vgpuscore.h"
#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"

namespace vgs {

struct ResolveOut {
    bool alive;    // every request of the pod found its devices on this node
    float score;   // NodeScore.score (score.go:180,203)
    int32_t nd;    // len(node.Devices), clamped to [0, dmax]
    // the node's usage as the pod leaves it (what addPod would fold in, scheduler.go:280-297)
    int32_t used[VGS_MAX_DEVICES], usedmem[VGS_MAX_DEVICES], usedcores[VGS_MAX_DEVICES];
};

// req_at(r) -> DevReq of the pod's r-th request; emit(r, t, pick) receives the t-th device of request r
template <typename ReqAt, typename Emit>
__device__ __forceinline__ void resolve_row(const RawPlanes &raw, int32_t row, int32_t dmax, int32_t n_reqs,
                                            ReqAt req_at, Emit emit, ResolveOut &o) {
    int32_t nd = raw.ndev[row];
    const size_t base = (size_t)row * dmax;
    int32_t count[VGS_MAX_DEVICES], totalmem[VGS_MAX_DEVICES], totalcore[VGS_MAX_DEVICES];
    int32_t numa[VGS_MAX_DEVICES], type[VGS_MAX_DEVICES], ord[VGS_MAX_DEVICES];
#pragma unroll
    for (int d = 0; d < VGS_MAX_DEVICES; d++) { // all loads issued at once: this thread is latency-bound
        const bool on = d < dmax; // (not d < ndev: that would put ndev's own load in front of all of these)
        o.used[d] = on ? raw.p[VGS_PLANE_USED][base + d] : 0;
        o.usedmem[d] = on ? raw.p[VGS_PLANE_USEDMEM][base + d] : 0;
        o.usedcores[d] = on ? raw.p[VGS_PLANE_USEDCORES][base + d] : 0;
        count[d] = on ? raw.p[VGS_PLANE_COUNT][base + d] : 0;
        totalmem[d] = on ? raw.p[VGS_PLANE_TOTALMEM][base + d] : 0;
        totalcore[d] = on ? raw.p[VGS_PLANE_TOTALCORE][base + d] : 0;
        numa[d] = on ? raw.p[VGS_PLANE_NUMA][base + d] : 0;
        type[d] = on ? raw.p[VGS_PLANE_TYPEID][base + d] : 0;
        ord[d] = d;
    }
    nd = min(max(nd, 0), dmax);
    o.nd = nd;

    float score = 0.0f;
    bool alive = true, in_ctr = false;
    int32_t total = 0, free_ = 0, sums = 0;
    for (int r = 0; r < n_reqs && alive; r++) {
        const DevReq rq = req_at(r);
        if (rq.flags & R_NEWCTR) {
            if (in_ctr) score = __fadd_rn(score, container_score(total, free_, nd - sums));
            total = free_ = sums = 0;
            in_ctr = true;
        }
        sums += rq.nums;
        if (rq.nums > nd) { // score.go:162
            alive = false;
            break;
        }
        // stable insertion sort, ascending (Numa, Count-Used)   score.go:45-50,165
        for (int i = 1; i < nd; i++) {
            for (int j = i; j > 0; j--) {
                int a = ord[j], b = ord[j - 1];
                bool less = numa[a] == numa[b]
                                ? wrap_sub(count[a], o.used[a]) < wrap_sub(count[b], o.used[b])
                                : numa[a] < numa[b];
                if (!less) break;
                ord[j] = b;
                ord[j - 1] = a;
            }
        }
        int32_t left = rq.nums, prevnuma = -1, nt = 0;
        int32_t tdev[VGS_MAX_DEVICES], tmem[VGS_MAX_DEVICES];
        bool ok = false;
        for (int i = nd - 1; i >= 0; i--) { // score.go:92
            const int d = ord[i];
            uint32_t t = (uint32_t)type[d];
            if (!(((t & 32u) ? rq.mask_hi : rq.mask_lo) >> (t & 31u) & 1u)) continue;
            if ((rq.flags & R_IDLE) && (raw.idle_types >> t & 1ull) && o.used[d] > 0)
                continue; // cambricon/device.go:97-99, part of the type check
            if ((rq.flags & R_NUMA) && prevnuma != numa[d]) { // :99-104
                left = rq.nums;
                prevnuma = numa[d];
                nt = 0;
            }
            if (count[d] <= o.used[d]) continue; // :107
            if (rq.coresreq > 100) break;        // :110
            int32_t memreq = (rq.flags & R_USEPCT) ? pct_mem(totalmem[d], rq.pct) : rq.memfix;
            if (wrap_sub(totalmem[d], o.usedmem[d]) < memreq) continue;
            if (wrap_sub(totalcore[d], o.usedcores[d]) < rq.coresreq) continue;
            if (totalcore[d] == 100 && rq.coresreq == 100 && o.used[d] > 0) continue;
            if (totalcore[d] != 0 && o.usedcores[d] == totalcore[d] && rq.coresreq == 0) continue;
            if (left > 0) {
                left--;
                tdev[nt] = d;
                tmem[nt] = memreq;
                nt++;
            }
            if (left == 0) {
                ok = true;
                break;
            }
        }
        if (!ok) {
            alive = false;
            break;
        }
        for (int t = 0; t < nt; t++) { // :168-174
            const int d = tdev[t];
            total = wrap_add(total, count[d]);
            free_ = wrap_add(free_, wrap_sub(count[d], o.used[d]));
            o.used[d] = wrap_add(o.used[d], 1);
            o.usedcores[d] = wrap_add(o.usedcores[d], rq.coresreq);
            o.usedmem[d] = wrap_add(o.usedmem[d], tmem[t]);
            vgs_pick pk;
            pk.slot = d;
            pk.usedmem = tmem[t];
            pk.usedcores = rq.coresreq;
            emit(r, t, pk);
        }
    }
    if (alive && in_ctr) score = __fadd_rn(score, container_score(total, free_, nd - sums));
    o.alive = alive;
    o.score = score;
}

// The same restatement carried by one WARP: lane d holds device slot d of the node.  The stable insertion sort
// becomes "new position = number of devices that are smaller under (Numa, Count-Used, previous position)" (a stable
// sort is the lexicographic sort with the old position as last key), the reverse walk becomes bit algebra in
// position space: m = OR over lanes of (fit << position); a fitting device is the t-th pick when t fitting devices
// stand at higher positions, t = popc(m >> (position + 1)), and it is taken when t < Nums.  numa-bind: the walk
// drops its picks at every change of Numa among the devices that pass the type check (score.go:94-104), and equal
// Numa values are contiguous after the sort, so the group with the highest Numa holding >= Nums fitting devices
// supplies the picks.  Used by the single-pod Filter kernel (latency: ~1 us instead of ~9 for the one-thread
// version above, whose local-memory arrays make every step a dependent L1 round trip).
// All 32 lanes must call it.  emit(r, t, pick) is called by the lane that owns the pick.  On return lane d < nd holds
// the device's usage as the pod leaves it in (used, usedmem, usedcores).
// the columns of device slot d = lane that do not change with usage (for a preparation without another load)
struct RowStatic {
    int32_t count, totalmem, totalcore, numa, type;
};
template <typename ReqAt, typename Emit>
__device__ __forceinline__ void resolve_row_warp(const RawPlanes &raw, int32_t row, int32_t dmax, int32_t n_reqs,
                                                 ReqAt req_at, Emit emit, bool &alive_out, float &score_out,
                                                 int32_t &nd_out, int32_t &used, int32_t &usedmem, int32_t &usedcores,
                                                 RowStatic *fixed = nullptr) {
    constexpr unsigned kAll = 0xffffffffu;
    const int d = (int)(threadIdx.x & 31u);
    const bool on = d < dmax;
    const size_t at = (size_t)row * dmax + (on ? d : 0);
    int32_t nd = raw.ndev[row];
    used = on ? raw.p[VGS_PLANE_USED][at] : 0;
    usedmem = on ? raw.p[VGS_PLANE_USEDMEM][at] : 0;
    usedcores = on ? raw.p[VGS_PLANE_USEDCORES][at] : 0;
    const int32_t count = on ? raw.p[VGS_PLANE_COUNT][at] : 0, totalmem = on ? raw.p[VGS_PLANE_TOTALMEM][at] : 0;
    const int32_t totalcore = on ? raw.p[VGS_PLANE_TOTALCORE][at] : 0, numa = on ? raw.p[VGS_PLANE_NUMA][at] : 0;
    const uint32_t type = on ? (uint32_t)raw.p[VGS_PLANE_TYPEID][at] : 0u;
    if (fixed) fixed->count = count, fixed->totalmem = totalmem, fixed->totalcore = totalcore, fixed->numa = numa, fixed->type = (int32_t)type;
    nd = min(max(nd, 0), dmax);
    nd_out = nd;
    const bool valid = d < nd;
    const unsigned validmask = nd >= 32 ? kAll : ((1u << nd) - 1u);
    int pos = d; // position in node.Devices (registration order at first, the previous sort's order afterwards)

    float score = 0.0f;
    bool alive = true, in_ctr = false;
    uint32_t total = 0u, free_ = 0u;
    int32_t sums = 0;
    for (int r = 0; r < n_reqs && alive; r++) { // every branch below is warp-uniform
        const DevReq rq = req_at(r);
        if (rq.flags & R_NEWCTR) {
            if (in_ctr) score = __fadd_rn(score, container_score((int32_t)total, (int32_t)free_, nd - sums));
            total = free_ = 0u;
            sums = 0;
            in_ctr = true;
        }
        sums += rq.nums;
        if (rq.nums > nd || rq.coresreq > 100) { // score.go:162; :110-113 (false whichever device is reached first)
            alive = false;
            break;
        }
        // sort.Sort(node.Devices): stable, ascending (Numa, Count-Used)   score.go:45-50,165
        const int32_t fr = wrap_sub(count, used);
        int npos = 0;
        for (int i = 0; i < nd; i++) {
            const int32_t ni = __shfl_sync(kAll, numa, i), fi = __shfl_sync(kAll, fr, i);
            const int pi = __shfl_sync(kAll, pos, i);
            const bool less = ni < numa || (ni == numa && (fi < fr || (fi == fr && pi < pos)));
            npos += less ? 1 : 0;
        }
        pos = valid ? npos : 31;
        // the predicates of fitInCertainDevice, per device   score.go:94-134
        const bool typeok = valid && ((((type & 32u) ? rq.mask_hi : rq.mask_lo) >> (type & 31u)) & 1u) &&
                            !((rq.flags & R_IDLE) && (raw.idle_types >> type & 1ull) && used > 0);
        const int32_t memreq = (rq.flags & R_USEPCT) ? pct_mem(totalmem, rq.pct) : rq.memfix;
        const bool fit = typeok && count > used && wrap_sub(totalmem, usedmem) >= memreq &&
                         wrap_sub(totalcore, usedcores) >= rq.coresreq &&
                         !(totalcore == 100 && rq.coresreq == 100 && used > 0) &&
                         !(totalcore != 0 && usedcores == totalcore && rq.coresreq == 0);
        bool ingroup = fit;
        if (rq.flags & R_NUMA) {
            const unsigned fitlanes = __ballot_sync(kAll, fit);
            const unsigned same = __match_any_sync(kAll, numa) & validmask;
            const bool cand = fit && __popc(fitlanes & same) >= rq.nums;
            const unsigned any = __ballot_sync(kAll, cand);
            const int32_t gsel = __reduce_max_sync(kAll, cand ? numa : INT32_MIN);
            ingroup = any != 0u && fit && numa == gsel;
        }
        const unsigned m = __reduce_or_sync(kAll, ingroup ? (1u << pos) : 0u); // fitting devices, in position space
        if (__popc(m) < rq.nums) {
            alive = false;
            break;
        }
        const int t = __popc(m >> 1 >> pos); // fitting devices the reverse walk meets before this one
        const bool taken = ingroup && t < rq.nums;
        if (taken) { // score.go:135-145
            vgs_pick pk;
            pk.slot = d;
            pk.usedmem = memreq;
            pk.usedcores = rq.coresreq;
            emit(r, t, pk);
        }
        total += __reduce_add_sync(kAll, taken ? (uint32_t)count : 0u);           // :169
        free_ += __reduce_add_sync(kAll, taken ? (uint32_t)wrap_sub(count, used) : 0u); // :170, before Used++
        if (taken) {                                                             // :171-173
            used = wrap_add(used, 1);
            usedcores = wrap_add(usedcores, rq.coresreq);
            usedmem = wrap_add(usedmem, memreq);
        }
    }
    if (alive && in_ctr) score = __fadd_rn(score, container_score((int32_t)total, (int32_t)free_, nd - sums));
    alive_out = alive;
    score_out = score;
}

} // namespace vgs
