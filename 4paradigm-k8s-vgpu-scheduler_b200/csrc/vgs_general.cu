// vgs_general.cu -- k1_score_walk: the score kernel for pods with several requests (containers x
// vendors), numa-bind, or one request for three and more devices.  Thread = node, loop over pods.
//
// What it restates: calcScore / fitInDevices / fitInCertainDevice with the state carried from one
// request of a pod to the next (pkg/scheduler/score.go:154-214): sort.Sort before every request
// (:165, stable by (Numa, Count-Used)), reverse walk taking the first Nums fitting devices (:92-149),
// then Used++, Usedmem += , Usedcores += on the taken devices (:168-174).
//
// How: nothing is ever sorted.  Every device carries one 32-bit ordering key
//
//        [ numa rank : 4 ][ Count - Used + bias : 21 ][ arrival stamp : 7 ]
//
// and the reverse walk is "descending key".  The stamp restates the stability of the insertion
// sort: among devices with equal (Numa, Count-Used) the walk order is the order they had before
// this sort.  A device only ever changes level by being taken (free slots - 1); it then lands
// behind everything with more free slots and -- because it had a larger key than every device
// already on its new level, and a stable sort keeps that -- ahead of all of them.  So "later
// arrival first" orders a level, initial residents in K0's walk order: stamp = a per-pod counter
// that grows from request to request (K0 walk position for the initial residents).  Devices taken
// by one request keep their relative order: within a request the stamps fall in take order.
// The first Nums fitting devices are then Nums rounds of "largest key among the candidates", and
// a take rewrites one key: (key | stamp mask) - (0x80 + 0x7f - stamp) lowers the level and stamps it.
//
// Usage flags are not maintained: Count <= Used is the key's level field, "cores full" is
// Totalcore - Usedcores == 0, and the two rules that read Used > 0 (exclusive card,
// score.go:128; Cambricon idle share, cambricon/device.go:97) set their bit at the take, only
// for pods that carry such a request.
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/vgpuscore.h"
#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"
#include "vgs_filter_tail.cuh"

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

// a * b + c on the integer multiply-add unit, kept as one instruction
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
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
__device__ __forceinline__ uint32_t eq01(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("{\n.reg .pred p;\nsetp.eq.u32 p, %1, %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t ltu01(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("{\n.reg .pred p;\nsetp.lt.u32 p, %1, %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(r) : "r"(a), "r"(b));
    return r;
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

// FILTER: the single-pod Filter call in ONE launch.  The pod and its requests arrive as launch parameters (no
// staging copy in front of the kernel), and the CTA that finishes last resolves the winner (the resolve kernel's
// body, vgs_resolve.cuh) and stores result, picks and the status word straight into the pinned host mirrors:
// the call is one launch + one stream synchronisation.
template <int DM, bool VERIFY, bool FILTER>
__global__ void __launch_bounds__(128)
    k1_score_walk(const int4 *__restrict__ walk, int32_t n_tiles, int32_t n_nodes, int32_t row_base,
                  const DevReq *__restrict__ reqs_g, const GenPod *__restrict__ pods_g, int32_t n_pods, ScoreLaunch L,
                  unsigned long long *__restrict__ keys, int32_t *__restrict__ counts, uint32_t *__restrict__ vfit,
                  float *__restrict__ vscore, int32_t vwords, const __grid_constant__ FilterOne fo) {
    const DevReq *reqs = FILTER ? fo.reqs : reqs_g;
    const GenPod *pods = FILTER ? &fo.pod : pods_g;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_items = L.n_chunks * L.n_tile_groups;
    const bool small = n_pods <= kWarp; // see flush_small
    // Consecutive pods with the same first request (the host sorts the class by it, G_SAME_FIRST / G_SAVE_FIRST)
    // share everything up to and including that request: the node state it leaves is parked here, per
    // thread, and the followers start from it at their second request.
    __shared__ int4 s_save[3 * (DM / 4) + 1][128];
    uint32_t one = 1u;
    asm volatile("" : "+r"(one)); // opaque 1 (see take_if)
    unsigned long long acckey = 0ull;
    int32_t acccnt = 0;

    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int chunk = item / L.n_tile_groups, tg = item - chunk * L.n_tile_groups;
        const int pod0 = chunk * L.chunk_len;
        const int npod = min(L.chunk_len, n_pods - pod0);
        const int tile = tg * L.warps_per_cta + warp;
        if (tile >= n_tiles) continue;

        const int4 *tp = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
        // per node, for the whole item: Count, the static meta bits, the initial keys
        int32_t count[DM], meta0[DM], fs0[DM];
        uint32_t key0[DM];
        load_word<DM>(tp, W_COUNT, lane, count);
        load_word<DM>(tp, W_META, lane, meta0);
        load_word<DM>(tp, W_FREESLOTS, lane, fs0);
        const int32_t ndev = (meta0[0] >> M_NDEV_SHIFT) & 15;
#pragma unroll
        for (int j = 0; j < DM; j++) // device j of the tile is the j-th the first walk visits: stamp DM-1-j
            key0[j] = ((((uint32_t)meta0[j] >> M_NUMA_SHIFT) & 15u) << kNumaShift) |
                      (((uint32_t)fs0[j] + kFsBias) << kStampBits) | (uint32_t)(DM - 1 - j);
        const uint32_t row0 = (uint32_t)(row_base + tile * kWarp);
        const int row = tile * kWarp + lane;

        for (int pp = 0; pp < npod; pp += kWarp) {
            unsigned long long mykey = 0ull;
            int32_t mycnt = 0;
            const int kend = min(kWarp, npod - pp);
            for (int k = 0; k < kend; k++) {
                const GenPod pod = pods[pod0 + pp + k];
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
                for (int j = 0; j < DM; j++) mt[j] = (uint32_t)meta0[j];
                // (item == first pod of this warp's run over the tile: nothing parked yet)
                if ((pod.flags & G_SAME_FIRST) && (pp + k) > 0) {
#pragma unroll
                    for (int q = 0; q < DM / 4; q++) {
                        const int4 a = s_save[q][threadIdx.x], b = s_save[DM / 4 + q][threadIdx.x],
                                   c = s_save[2 * (DM / 4) + q][threadIdx.x];
                        fm[4 * q] = a.x, fm[4 * q + 1] = a.y, fm[4 * q + 2] = a.z, fm[4 * q + 3] = a.w;
                        fc[4 * q] = b.x, fc[4 * q + 1] = b.y, fc[4 * q + 2] = b.z, fc[4 * q + 3] = b.w;
                        key[4 * q] = (uint32_t)c.x, key[4 * q + 1] = (uint32_t)c.y, key[4 * q + 2] = (uint32_t)c.z,
                                key[4 * q + 3] = (uint32_t)c.w;
                    }
                    const int4 sc = s_save[3 * (DM / 4)][threadIdx.x];
                    total = (uint32_t)sc.x, free_ = (uint32_t)sc.y, sums = sc.z;
                    stamp = (uint32_t)sc.w & 0xffffu;
                    alive = (sc.w >> 16) & 1;
                    in_ctr = true;
                    r_first = 1;
                } else {
                    // free memory and cores re-read from the (L1-resident) tile, keys from the registers above
                    load_word<DM>(tp, W_FREEMEM, lane, fm);
                    load_word<DM>(tp, W_FREECORES, lane, fc);
#pragma unroll
                    for (int j = 0; j < DM; j++) key[j] = key0[j];
                    if (uniform_type) {
                        // one type mask for the whole pod: a device that fails it never fits, whatever the
                        // request -- it gets a free memory no request is small enough for
                        const DevReq r0 = FILTER ? reqs[pod.req_begin] : load_req(&reqs[pod.req_begin]);
#pragma unroll
                        for (int j = 0; j < DM; j++)
                            fm[j] = type_ok(r0.mask_lo, r0.mask_hi, mt[j]) ? fm[j] : INT32_MIN;
                    }
                }

                for (int r = r_first; r < pod.n_reqs; r++) {
                    const DevReq rq = FILTER ? reqs[pod.req_begin + r] : load_req(&reqs[pod.req_begin + r]);
                    if (rq.flags & R_NEWCTR) { // score.go:180,203: one container done
                        // (a lane that is out has taken nothing: free = 0 would send the whole warp through the
                        // division's slow path)
                        if (in_ctr) score = __fadd_rn(score, container_score((int32_t)total, alive ? (int32_t)free_ : 1, ndev - sums));
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
                            load_word<DM>(tp, W_TOTALMEM, lane, totalmem);
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
                                        total, (uint32_t)count[j], one);
                        }
                    };
                    if (rq.flags & R_USEPCT) body(std::true_type{}); // warp-uniform
                    else body(std::false_type{});
                    stamp += (uint32_t)rounds;
                    if (r == 0 && (pod.flags & G_SAVE_FIRST)) { // the next pod starts from here
#pragma unroll
                        for (int q = 0; q < DM / 4; q++) {
                            s_save[q][threadIdx.x] = make_int4(fm[4 * q], fm[4 * q + 1], fm[4 * q + 2], fm[4 * q + 3]);
                            s_save[DM / 4 + q][threadIdx.x] = make_int4(fc[4 * q], fc[4 * q + 1], fc[4 * q + 2], fc[4 * q + 3]);
                            s_save[2 * (DM / 4) + q][threadIdx.x] = make_int4((int32_t)key[4 * q], (int32_t)key[4 * q + 1],
                                                                           (int32_t)key[4 * q + 2], (int32_t)key[4 * q + 3]);
                        }
                        s_save[3 * (DM / 4)][threadIdx.x] =
                            make_int4((int32_t)total, (int32_t)free_, sums, (int32_t)(stamp | (alive ? 0x10000u : 0u)));
                    }
                    if (!__any_sync(0xffffffffu, alive)) break;
                }
                if (in_ctr) score = __fadd_rn(score, container_score((int32_t)total, alive ? (int32_t)free_ : 1, ndev - sums));
                const bool fitn = alive;
                if (VERIFY) {
                    unsigned b = __ballot_sync(0xffffffffu, fitn);
                    if (lane == 0) vfit[(size_t)pod.orig * vwords + tile] = b;
                    if (row < n_nodes) vscore[(size_t)pod.orig * n_nodes + row] = fitn ? score : 0.0f;
                }
                publish(fitn, score, k, lane, row0, mykey, mycnt);
            }
            if (small) {
                acckey = mykey > acckey ? mykey : acckey;
                acccnt += mycnt;
            } else if (lane < kend && mykey != 0ull) {
                int op = pods[pod0 + pp + lane].orig;
                atomicMax(&keys[op], mykey);
                atomicAdd(&counts[op], mycnt);
            }
        }
    }
    if (small) flush_small<4>(acckey, acccnt, lane, warp, n_pods, nullptr, pods, keys, counts);
    if constexpr (FILTER) filter_tail<DM>(fo, keys, counts, n_nodes, row_base);
}

template <int DM>
static cudaError_t launch_walk_t(int dev, const ScoreArgs &a, const DevReq *reqs, const GenPod *pods, int32_t n_pods,
                                 cudaStream_t st) {
    ScoreLaunch L;
    cudaError_t e;
    static const FilterOne none{};
    if (a.vfit) {
        auto k = k1_score_walk<DM, true, false>;
        if ((e = plan_score_launch(reinterpret_cast<const void *>(k), dev, a.n_tiles, n_pods, 4, 0, &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, 0, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, pods, n_pods, L, a.keys, a.counts,
                                     a.vfit, a.vscore, a.vwords, none);
    } else {
        auto k = k1_score_walk<DM, false, false>;
        if ((e = plan_score_launch(reinterpret_cast<const void *>(k), dev, a.n_tiles, n_pods, 4, 0, &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, 0, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, pods, n_pods, L, a.keys, a.counts,
                                     nullptr, nullptr, 0, none);
    }
    return cudaGetLastError();
}

template <int DM>
static cudaError_t launch_filter_t(int dev, const ScoreArgs &a, const FilterOne &fo, cudaStream_t st) {
    ScoreLaunch L;
    cudaError_t e;
    auto k = k1_score_walk<DM, false, true>;
    if ((e = plan_score_launch(reinterpret_cast<const void *>(k), dev, a.n_tiles, 1, 4, 0, &L)) != cudaSuccess) return e;
    k<<<L.grid, L.block, 0, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, nullptr, nullptr, 1, L, a.keys, a.counts,
                                 nullptr, nullptr, 0, fo);
    return cudaGetLastError();
}

cudaError_t launch_filter_one(int dm, int dev, const ScoreArgs &a, const FilterOne &fo, cudaStream_t st) {
    if (a.n_tiles <= 0) return cudaErrorInvalidValue; // the caller takes the phased path for an empty snapshot
    const DevReq &q = fo.reqs[0];
    if (fo.pod.n_reqs == 1 && !(fo.pod.flags & G_LEGACY) && (!(q.flags & R_NUMA) || q.nums == 1))
        return launch_filter_single(dm, q.nums == 1, dev, a, fo, st);
    return dm == 8 ? launch_filter_t<8>(dev, a, fo, st) : launch_filter_t<12>(dev, a, fo, st);
}

cudaError_t launch_score_walk(int dm, int dev, const ScoreArgs &a, const DevReq *reqs, const GenPod *pods,
                              int32_t n_pods, cudaStream_t st) {
    if (n_pods <= 0 || a.n_tiles <= 0) return cudaSuccess;
    return dm == 8 ? launch_walk_t<8>(dev, a, reqs, pods, n_pods, st) : launch_walk_t<12>(dev, a, reqs, pods, n_pods, st);
}

} // namespace vgs
