// vgs_walk_eval.cuh -- one (pod, node) pair of the walk kernel (vgs_general.cu explains the ordering keys): the
// device function shared by k1_score_walk (thread = node of a tile of the walk copy, loop over pods) and by the
// batch-scale drain (vgs_drain.cu: the same evaluation on the rows a chunk of pods has touched, from copies of their
// tile words in shared memory).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"

namespace vgs {

// key of a device that fits, else 0: Count > Used (level field above the bias; pads sit on the bias),
// free memory >= request, free cores >= threshold -- three chained compares and a select
__device__ __forceinline__ uint32_t cand_of(uint32_t key, int32_t fm, int32_t mem, int32_t fc, int32_t thr) {
    uint32_t r;
    asm("{\n"
        ".reg .pred p;\n"
        ".reg .u32 t;\n"
        "and.b32 t, %1, %6;\n"
        "setp.gt.u32 p, t, %7;\n"
        "setp.ge.and.s32 p, %2, %3, p;\n"
        "setp.ge.and.s32 p, %4, %5, p;\n"
        "selp.u32 %0, %1, 0, p;\n"
        "}"
        : "=r"(r)
        : "r"(key), "r"(fm), "r"(mem), "r"(fc), "r"(thr), "n"((1u << kNumaShift) - 1u),
          "n"((kFsBias << kStampBits) | kStampMask));
    return r;
}

// if (cand == mm): key = (key | stamp mask) + nrestamp; fm += dmem; fc += dcores; total += count; cand = 0.
// `one` is a register holding 1 that the compiler cannot see through: a * 1 + b stays a multiply-add.
__device__ __forceinline__ void take_if(uint32_t &cand, uint32_t mm, uint32_t &key, uint32_t nrestamp, int32_t &fm,
                                        uint32_t dmem, int32_t &fc, uint32_t dcores, uint32_t &total, uint32_t count,
                                        uint32_t one) {
    asm("{\n"
        ".reg .pred p;\n"
        ".reg .u32 t;\n"
        "setp.eq.u32 p, %0, %5;\n"
        "or.b32 t, %1, %6;\n"
        "@p mad.lo.u32 %1, %7, %8, t;\n"
        "@p mad.lo.u32 %2, %7, %9, %2;\n"
        "@p mad.lo.u32 %3, %7, %10, %3;\n"
        "@p mad.lo.u32 %4, %7, %11, %4;\n"
        "@p mov.u32 %0, 0;\n"
        "}"
        : "+r"(cand), "+r"(key), "+r"(fm), "+r"(fc), "+r"(total)
        : "r"(mm), "n"(kStampMask), "r"(one), "r"(nrestamp), "r"(dmem), "r"(dcores), "r"(count));
}
// numa-bind (score.go:99-104): the walk drops its picks at every change of Numa, so the highest numa
// group holding >= Nums fitting devices supplies them all; every other candidate is struck out
template <int DM>
__device__ __forceinline__ void numa_filter(uint32_t (&cand)[DM], int32_t nums) {
    uint32_t c2[DM];
#pragma unroll
    for (int j = 0; j < DM; j++) c2[j] = cand[j];
    uint32_t gsel = 0xffffffffu; // none
    bool done = false;
    for (int round = 0; round < DM; round++) {
        uint32_t m = 0u;
#pragma unroll
        for (int j = 0; j < DM; j++) m = max(m, c2[j]);
        const uint32_t g = m >> kNumaShift;
        int cnt = 0;
#pragma unroll
        for (int j = 0; j < DM; j++) cnt += (c2[j] != 0u && (c2[j] >> kNumaShift) == g) ? 1 : 0;
        if (!done && m != 0u && cnt >= nums) {
            gsel = g;
            done = true;
        }
        if (m == 0u) done = true;
#pragma unroll
        for (int j = 0; j < DM; j++) c2[j] = (c2[j] >> kNumaShift) == g ? 0u : c2[j];
        if (__all_sync(0xffffffffu, done)) break;
    }
#pragma unroll
    for (int j = 0; j < DM; j++) cand[j] = (cand[j] >> kNumaShift) == gsel ? cand[j] : 0u;
}

// NC: the words come through the read-only path (the score kernels: nothing writes the walk copy while they run);
// else plain loads -- a generic pointer (shared or global memory) that the running kernel may have written
template <int DM, bool NC>
__device__ __forceinline__ void tile_word(const int4 *tile, int word, int lane, int32_t (&dst)[DM]) {
    if constexpr (NC) {
        load_word<DM>(tile, word, lane, dst);
    } else {
#pragma unroll
        for (int q = 0; q < DM / 4; q++) {
            const int4 x = tile[(word * (DM / 4) + q) * kWarp + lane];
            dst[4 * q + 0] = x.x, dst[4 * q + 1] = x.y, dst[4 * q + 2] = x.z, dst[4 * q + 3] = x.w;
        }
    }
}

// what a thread keeps of its node across pods: Count, the static meta bits, the initial keys
template <int DM>
struct WalkNode {
    const int4 *tp; // the node's tile
    int lane;       // its column there
    int32_t count[DM], meta0[DM];
    uint32_t key0[DM];
    int32_t ndev;
};

template <int DM, bool NC>
__device__ __forceinline__ void walk_node_load(const int4 *tp, int lane, WalkNode<DM> &nd) {
    int32_t fs0[DM];
    nd.tp = tp, nd.lane = lane;
    tile_word<DM, NC>(tp, W_COUNT, lane, nd.count);
    tile_word<DM, NC>(tp, W_META, lane, nd.meta0);
    tile_word<DM, NC>(tp, W_FREESLOTS, lane, fs0);
    nd.ndev = (nd.meta0[0] >> M_NDEV_SHIFT) & 15;
#pragma unroll
    for (int j = 0; j < DM; j++) // device j of the tile is the j-th the first walk visits: stamp DM-1-j
        nd.key0[j] = ((((uint32_t)nd.meta0[j] >> M_NUMA_SHIFT) & 15u) << kNumaShift) |
                     (((uint32_t)fs0[j] + kFsBias) << kStampBits) | (uint32_t)(DM - 1 - j);
}

// One pod on one node: does every request find its devices (alive), and NodeScore.score.  All 32 lanes of a warp
// call it with the SAME pod (trip counts and the numa rounds are warp-uniform); req_at(r) = the pod's r-th request.
// SHARE: consecutive pods with the same first request (G_SAME_FIRST / G_SAVE_FIRST, set by the host) share
// everything up to and including that request through `save` ([3 * DM/4 + 1][128] int4, column = threadIdx.x);
// may_restore = something is parked there for this node.  `one` = a register holding 1 the compiler cannot see
// through (take_if).
template <int DM, bool NC, bool SHARE, typename ReqAt>
__device__ __forceinline__ void walk_eval(const WalkNode<DM> &nd, const GenPod &pod, ReqAt req_at, uint32_t one,
                                          int4 (*save)[128], bool may_restore, bool &alive_out, float &score_out) {
    // the pod's private, mutable view of the node
    int32_t fm[DM], fc[DM];
    uint32_t key[DM], mt[DM];
    const bool uniform_type = pod.flags & G_UNIFORM_TYPE;
    const bool track_used = !(pod.flags & G_NO_USED_TRACK);
    bool alive = true, in_ctr = false;
    float score = 0.0f;
    uint32_t total = 0u, free_ = 0u; // wrapping, like Go's int32
    int32_t sums = 0;
    uint32_t stamp = (uint32_t)DM; // next arrival stamp
    int r_first = 0;
#pragma unroll
    for (int j = 0; j < DM; j++) mt[j] = (uint32_t)nd.meta0[j];
    // (item == first pod of this warp's run over the tile: nothing parked yet)
    if (SHARE && (pod.flags & G_SAME_FIRST) && may_restore) {
#pragma unroll
        for (int q = 0; q < DM / 4; q++) {
            const int4 a = save[q][threadIdx.x], b = save[DM / 4 + q][threadIdx.x],
                       c = save[2 * (DM / 4) + q][threadIdx.x];
            fm[4 * q] = a.x, fm[4 * q + 1] = a.y, fm[4 * q + 2] = a.z, fm[4 * q + 3] = a.w;
            fc[4 * q] = b.x, fc[4 * q + 1] = b.y, fc[4 * q + 2] = b.z, fc[4 * q + 3] = b.w;
            key[4 * q] = (uint32_t)c.x, key[4 * q + 1] = (uint32_t)c.y, key[4 * q + 2] = (uint32_t)c.z,
                    key[4 * q + 3] = (uint32_t)c.w;
        }
        const int4 sc = save[3 * (DM / 4)][threadIdx.x];
        total = (uint32_t)sc.x, free_ = (uint32_t)sc.y, sums = sc.z;
        stamp = (uint32_t)sc.w & 0xffffu;
        alive = (sc.w >> 16) & 1;
        in_ctr = true;
        r_first = 1;
    } else {
        // free memory and cores re-read from the (L1-resident) tile, keys from the registers above
        tile_word<DM, NC>(nd.tp, W_FREEMEM, nd.lane, fm);
        tile_word<DM, NC>(nd.tp, W_FREECORES, nd.lane, fc);
#pragma unroll
        for (int j = 0; j < DM; j++) key[j] = nd.key0[j];
        if (uniform_type) {
            // one type mask for the whole pod: a device that fails it never fits, whatever the
            // request -- it gets a free memory no request is small enough for
            const DevReq r0 = req_at(0);
#pragma unroll
            for (int j = 0; j < DM; j++)
                fm[j] = type_ok(r0.mask_lo, r0.mask_hi, mt[j]) ? fm[j] : INT32_MIN;
        }
    }

    for (int r = r_first; r < pod.n_reqs; r++) {
        const DevReq rq = req_at(r);
        if (rq.flags & R_NEWCTR) { // score.go:180,203: one container done
            // (a lane that is out has taken nothing: free = 0 would send the whole warp through the
            // division's slow path)
            if (in_ctr) score = __fadd_rn(score, container_score((int32_t)total, alive ? (int32_t)free_ : 1, nd.ndev - sums));
            total = 0u;
            free_ = 0u;
            sums = 0;
            in_ctr = true;
        }
        sums += rq.nums;
        // (Nums > len(node.Devices), score.go:162: the rounds below run out of candidates; the trip
        // count is capped so that an absurd Nums costs nothing)
        const int32_t rounds = min(rq.nums, DM + 1);
        // ---- which devices fit (score.go:94-134), all against the state before this request
        uint32_t cand[DM];
        // the bits of the blockmask that are not arithmetic here
        const uint32_t bm = rq.blockmask & (M_EXCL | M_BUSY370 | M_ALWAYS);
        auto body = [&](auto PCT) {
            constexpr bool kPct = decltype(PCT)::value;
            int32_t mem[kPct ? DM : 1];
            if constexpr (kPct) {
                int32_t totalmem[DM];
                tile_word<DM, NC>(nd.tp, W_TOTALMEM, nd.lane, totalmem);
#pragma unroll
                for (int j = 0; j < DM; j++) mem[j] = pct_mem(totalmem[j], rq.pct);
            } else {
                mem[0] = rq.memfix;
            }
            if (rq.blockmask & M_FULL) { // Coresreq == 0 (score.go:132): free cores >= 0, and a card
                // whose cores are all in use (Totalcore != 0 && Usedcores == Totalcore) loses
#pragma unroll
                for (int j = 0; j < DM; j++)
                    cand[j] = cand_of(key[j], fm[j], mem[kPct ? j : 0], fc[j], (int32_t)((mt[j] >> 23) & 1u));
            } else {
#pragma unroll
                for (int j = 0; j < DM; j++) cand[j] = cand_of(key[j], fm[j], mem[kPct ? j : 0], fc[j], rq.coresreq);
            }
            if (bm) { // exclusive request, idle-share request, Coresreq > 100: rare
#pragma unroll
                for (int j = 0; j < DM; j++) cand[j] = (mt[j] & bm) ? 0u : cand[j];
            }
            if (!uniform_type) {
#pragma unroll
                for (int j = 0; j < DM; j++) cand[j] = type_ok(rq.mask_lo, rq.mask_hi, mt[j]) ? cand[j] : 0u;
            }
            if (rq.flags & R_NUMA) numa_filter<DM>(cand, rq.nums);
            // ---- the first Nums candidates in walk order = Nums rounds of "largest key"; each take
            // (score.go:168-174) is applied at once -- fit was decided above, so later rounds do not see
            // it.  One compare per device, the updates predicated on it and placed on the multiply-add
            // unit (take_if), beside the compare / select unit that the fit tests keep busy.
            const uint32_t ncores = 0u - (uint32_t)rq.coresreq;
            for (int i = 0; i < rounds; i++) { // warp-uniform trip count
                uint32_t m = 0u;
#pragma unroll
                for (int j = 0; j < DM; j++) m = max(m, cand[j]);
                alive = alive && m != 0u;
                // the taken device's Count - Used is the level field of its key (score.go:170)
                free_ += ((m >> kStampBits) & ((1u << kFsBits) - 1u)) - kFsBias;
                const uint32_t mm = m != 0u ? m : 0xffffffffu; // keys never reach this value
                // one level down and stamped as the newest arrival there; devices taken by one request
                // keep their relative order: the first taken gets the request's highest stamp
                const uint32_t nrestamp = (stamp + (uint32_t)(rounds - 1 - i)) - (0x80u + kStampMask);
                if (track_used) { // Used > 0 from here on: exclusive cards, busy 370s
#pragma unroll
                    for (int j = 0; j < DM; j++)
                        mt[j] |= cand[j] == mm ? (((mt[j] >> 3) & M_EXCL) | ((mt[j] >> 1) & M_BUSY370)) : 0u;
                }
#pragma unroll
                for (int j = 0; j < DM; j++)
                    take_if(cand[j], mm, key[j], nrestamp, fm[j], 0u - (uint32_t)mem[kPct ? j : 0], fc[j], ncores,
                            total, (uint32_t)nd.count[j], one);
            }
        };
        if (rq.flags & R_USEPCT) body(std::true_type{}); // warp-uniform
        else body(std::false_type{});
        stamp += (uint32_t)rounds;
        if (SHARE && r == 0 && (pod.flags & G_SAVE_FIRST)) { // the next pod starts from here
#pragma unroll
            for (int q = 0; q < DM / 4; q++) {
                save[q][threadIdx.x] = make_int4(fm[4 * q], fm[4 * q + 1], fm[4 * q + 2], fm[4 * q + 3]);
                save[DM / 4 + q][threadIdx.x] = make_int4(fc[4 * q], fc[4 * q + 1], fc[4 * q + 2], fc[4 * q + 3]);
                save[2 * (DM / 4) + q][threadIdx.x] = make_int4((int32_t)key[4 * q], (int32_t)key[4 * q + 1],
                                                               (int32_t)key[4 * q + 2], (int32_t)key[4 * q + 3]);
            }
            save[3 * (DM / 4)][threadIdx.x] =
                make_int4((int32_t)total, (int32_t)free_, sums, (int32_t)(stamp | (alive ? 0x10000u : 0u)));
        }
        if (!__any_sync(0xffffffffu, alive)) break;
    }
    if (in_ctr) score = __fadd_rn(score, container_score((int32_t)total, alive ? (int32_t)free_ : 1, nd.ndev - sums));
    alive_out = alive;
    score_out = score;
}

} // namespace vgs
