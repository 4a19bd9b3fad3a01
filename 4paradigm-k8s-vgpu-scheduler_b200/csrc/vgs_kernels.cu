// vgs_kernels.cu -- sm_100a kernels of libvgpuscore.so
//
//   K0  k0_prepare        raw planes -> walk-ordered warp tiles (+ validation)
//   K1t k1t_fill/prefix   bit tables of a pod-sliced class (one bit per pod)
//   K1p k1p_collect/sort/classify   per device: table rows it selects, candidate scores, score class
//   K1  k1_score_sliced   classes of >= 512 single-request pods asking for 1 or 2 devices:
//                         32 pods per register, the CTA walks one node at a time, SliceDev tiles
//                         staged with cp.async.bulk + mbarrier; scout launch + main launch
//       k1_score_nodesliced  one pod word of such a class, lane = node: the words beyond a multiple
//                         of 32 and classes of few words
//       k1_score_single   per (pod, node) pair: thread = node (state in registers), loop over a
//                         pod chunk staged in shared memory
//       k1_score_general  any number of containers / requests, numa-bind
//   K2  k2_merge          max over shards of the per-pod candidate keys
//   K3  k3_resolve        re-run the winner node of every pod, emit the picks (drain: also commit)
//   K4  k4_apply_delta    scatter usage deltas into the raw planes
//       k_copy_planes     pinned host planes -> device planes over PCIe (zero-copy upload)
//
// Reference semantics restated (never copied): pkg/scheduler/score.go:45-50 (device
// order), :86-152 (fitInCertainDevice), :154-181 (fitInDevices), :183-214 (calcScore),
// pkg/scheduler/scheduler.go:386-387 (best node).
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <vector>

#include "../../include/vgpuscore.h"
#include "vgs_filter_tail.cuh"
#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"
#include "vgs_prepare.cuh"
#include "vgs_resolve.cuh"

namespace vgs {

// ------------------------------------------------------------------------------------
// K0: prepare the walk-ordered copy.  One thread per node.
// ------------------------------------------------------------------------------------
template <int DM>
__global__ void __launch_bounds__(128)
    k0_prepare(RawPlanes raw, int32_t n_nodes, int32_t dmax, int4 *__restrict__ walk,
               int32_t n_tiles, uint32_t *__restrict__ status, const int32_t *__restrict__ rowlist,
               int32_t n_list) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int row;
    if (rowlist) {
        if (idx >= n_list) return;
        row = rowlist[idx];
        if (row < 0) return;
    } else {
        row = idx;
    }
    if (row >= n_tiles * kWarp) return;
    if (rowlist) { // scattered rows: straight stores
        prepare_node<DM>(raw, row, n_nodes, dmax, walk, status);
        return;
    }
    // whole snapshot: warp = tile, the tile's words are collected in shared memory and stored as 512-byte rows
    extern __shared__ __align__(16) int32_t k0_stage[];
    prepare_node<DM>(raw, row, n_nodes, dmax, walk, status, k0_stage + (threadIdx.x >> 5) * (kWalkWords * DM * kWarp));
}

// ------------------------------------------------------------------------------------
// K1 (single-request pods).  CTA item = (group of warps_per_cta node tiles, pod chunk).
// NUMS1: Nums == 1 -> the first fitting device in walk order decides, its score was
// prepared by K0.  Otherwise: first Nums fitting devices, total/free accumulated.
// ------------------------------------------------------------------------------------
// FILTER: the launch is a whole Filter call for one pod -- the request travels in the launch parameters and the
// CTA that finishes last resolves the winner (vgs_filter_tail.cuh).
template <int DM, bool NUMS1, bool VERIFY, bool FILTER>
__global__ void __launch_bounds__(256)
    k1_score_single(const int4 *__restrict__ walk, int32_t n_tiles, int32_t n_nodes,
                    int32_t row_base, const DevReq *__restrict__ reqs,
                    const int32_t *__restrict__ orig, int32_t n_pods, ScoreLaunch L,
                    unsigned long long *__restrict__ keys, int32_t *__restrict__ counts,
                    uint32_t *__restrict__ vfit, float *__restrict__ vscore, int32_t vwords,
                    const __grid_constant__ FilterOne fo) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DevReq *s_req = reinterpret_cast<DevReq *>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_items = L.n_chunks * L.n_tile_groups;
    const bool small = n_pods <= kWarp; // one chunk, one group of pods: see flush_small
    const bool one_chunk = L.n_chunks == 1;
    bool need_totalmem = true;
    unsigned long long acckey = 0ull;
    int32_t acccnt = 0;

    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int chunk = item / L.n_tile_groups, tg = item - chunk * L.n_tile_groups;
        const int pod0 = chunk * L.chunk_len;
        const int npod = min(L.chunk_len, n_pods - pod0);
        if constexpr (FILTER) {
            if (item == (int)blockIdx.x) {
                if (threadIdx.x == 0) s_req[0] = fo.reqs[0];
                need_totalmem = (fo.reqs[0].flags & R_USEPCT) != 0;
                __syncthreads();
            }
        } else if (!one_chunk || item == (int)blockIdx.x) {
            // stage the chunk's requests; with a single chunk once per CTA, and the warps then
            // stream their node tiles without meeting at a barrier again
            __syncthreads(); // previous item's readers are done
            const int4 *src = reinterpret_cast<const int4 *>(reqs + pod0);
            int4 *dst = reinterpret_cast<int4 *>(s_req);
            int pct = 0;
            for (int i = threadIdx.x; i < npod * 2; i += blockDim.x) {
                const int4 x = __ldg(&src[i]);
                dst[i] = x;
                if (i & 1) pct |= (uint32_t)x.y & R_USEPCT; // second half of a DevReq: .y = flags
            }
            need_totalmem = __syncthreads_or(pct) != 0;
        }
        const int tile = tg * L.warps_per_cta + warp;
        if (tile >= n_tiles) continue;

        const int4 *tp = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
        int32_t totalmem[DM], freemem[DM], freecores[DM], meta[DM];
        int32_t count[DM], freeslots[DM], score1[DM];
        if (need_totalmem) { // Totalmem only enters through memory percentages (score.go:117-120)
            load_word<DM>(tp, W_TOTALMEM, lane, totalmem);
        } else {
#pragma unroll
            for (int j = 0; j < DM; j++) totalmem[j] = 0;
        }
        load_word<DM>(tp, W_FREEMEM, lane, freemem);
        load_word<DM>(tp, W_FREECORES, lane, freecores);
        load_word<DM>(tp, W_META, lane, meta);
        if (NUMS1) {
            load_word<DM>(tp, W_SCORE1, lane, score1);
        } else {
            load_word<DM>(tp, W_COUNT, lane, count);
            load_word<DM>(tp, W_FREESLOTS, lane, freeslots);
        }
        const int32_t ndev = (meta[0] >> M_NDEV_SHIFT) & 15;
        const uint32_t row0 = (uint32_t)(row_base + tile * kWarp);
        const int row = tile * kWarp + lane;

        for (int pp = 0; pp < npod; pp += kWarp) {
            unsigned long long mykey = 0ull;
            int32_t mycnt = 0;
            const int kend = min(kWarp, npod - pp);
            for (int k = 0; k < kend; k++) {
                const DevReq rq = s_req[pp + k];
                bool fit;
                float score;
                if (NUMS1) {
                    int32_t sb = 0;
                    if (rq.flags & R_USEPCT) {
#pragma unroll
                        for (int j = DM - 1; j >= 0; j--) {
                            bool f = !(meta[j] & rq.blockmask) &&
                                     type_ok(rq.mask_lo, rq.mask_hi, meta[j]) &&
                                     freemem[j] >= pct_mem(totalmem[j], rq.pct) &&
                                     freecores[j] >= rq.coresreq;
                            sb = f ? score1[j] : sb;
                        }
                    } else {
#pragma unroll
                        for (int j = DM - 1; j >= 0; j--) {
                            bool f = !(meta[j] & rq.blockmask) &&
                                     type_ok(rq.mask_lo, rq.mask_hi, meta[j]) &&
                                     freemem[j] >= rq.memfix && freecores[j] >= rq.coresreq;
                            sb = f ? score1[j] : sb;
                        }
                    }
                    fit = sb != 0; // a prepared score is >= 1.0f, never 0
                    score = __int_as_float(sb);
                } else {
                    int32_t cnt = 0, total = 0, free_ = 0;
                    const bool usepct = rq.flags & R_USEPCT;
#pragma unroll
                    for (int j = 0; j < DM; j++) {
                        int32_t mem = usepct ? pct_mem(totalmem[j], rq.pct) : rq.memfix;
                        bool f = !(meta[j] & rq.blockmask) &&
                                 type_ok(rq.mask_lo, rq.mask_hi, meta[j]) && freemem[j] >= mem &&
                                 freecores[j] >= rq.coresreq;
                        bool take = f && cnt < rq.nums;
                        cnt += f ? 1 : 0;
                        total += take ? count[j] : 0;
                        free_ += take ? freeslots[j] : 0;
                    }
                    fit = cnt >= rq.nums;
                    score = fit ? container_score(total, free_, ndev - rq.nums) : 0.0f;
                }
                if (VERIFY) {
                    int op = orig[pod0 + pp + k];
                    unsigned b = __ballot_sync(0xffffffffu, fit);
                    if (lane == 0) vfit[(size_t)op * vwords + tile] = b;
                    if (row < n_nodes) vscore[(size_t)op * n_nodes + row] = fit ? score : 0.0f;
                }
                publish(fit, score, k, lane, row0, mykey, mycnt);
            }
            if (small) {
                acckey = mykey > acckey ? mykey : acckey;
                acccnt += mycnt;
            } else if (lane < kend && mykey != 0ull) {
                int op = orig[pod0 + pp + lane];
                atomicMax(&keys[op], mykey);
                atomicAdd(&counts[op], mycnt);
            }
        }
    }
    if constexpr (FILTER) {
        flush_small<8>(acckey, acccnt, lane, warp, 1, nullptr, &fo.pod, keys, counts);
        filter_tail<DM>(fo, keys, counts, n_nodes, row_base);
    } else {
        if (small) flush_small<8>(acckey, acccnt, lane, warp, n_pods, orig, nullptr, keys, counts);
    }
}

// ------------------------------------------------------------------------------------
// K1 (general pods): any number of containers / requests, numa-bind, state carried
// from one request to the next exactly as calcScore/fitInDevices mutate node.Devices.
//
// Devices stay in their K0 walk order in registers.  The order the reference would see
// after each sort.Sort is tracked as pos[j] (ascending position); the reverse walk then
// visits devices by descending key (numa rank, free slots, pos).  "First Nums fitting
// devices of the walk" = fitting devices with fewer than Nums fitting devices ahead.
// ------------------------------------------------------------------------------------
template <int DM, bool VERIFY>
__global__ void __launch_bounds__(128)
    k1_score_general(const int4 *__restrict__ walk, int32_t n_tiles, int32_t n_nodes,
                     int32_t row_base, const DevReq *__restrict__ reqs,
                     const GenPod *__restrict__ pods, int32_t n_pods, ScoreLaunch L,
                     unsigned long long *__restrict__ keys, int32_t *__restrict__ counts,
                     uint32_t *__restrict__ vfit, float *__restrict__ vscore, int32_t vwords) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_items = L.n_chunks * L.n_tile_groups;
    const bool small = n_pods <= kWarp; // see flush_small
    unsigned long long acckey = 0ull;
    int32_t acccnt = 0;

    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int chunk = item / L.n_tile_groups, tg = item - chunk * L.n_tile_groups;
        const int pod0 = chunk * L.chunk_len;
        const int npod = min(L.chunk_len, n_pods - pod0);
        const int tile = tg * L.warps_per_cta + warp;
        if (tile >= n_tiles) continue;

        const int4 *tp = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
        // immutable per node: stays in registers for the whole item
        int32_t totalmem[DM], count[DM];
        load_word<DM>(tp, W_TOTALMEM, lane, totalmem);
        load_word<DM>(tp, W_COUNT, lane, count);
        const int32_t ndev = (__ldg(reinterpret_cast<const int32_t *>(&tp[(W_META * (DM / 4)) * kWarp + lane])) >>
                              M_NDEV_SHIFT) & 15;
        const uint32_t row0 = (uint32_t)(row_base + tile * kWarp);
        const int row = tile * kWarp + lane;

        for (int pp = 0; pp < npod; pp += kWarp) {
            unsigned long long mykey = 0ull;
            int32_t mycnt = 0;
            const int kend = min(kWarp, npod - pp);
            for (int k = 0; k < kend; k++) {
                const GenPod pod = pods[pod0 + pp + k];
                // the pod's private, mutable copy of the node: re-read from the (L1-resident) walk
                // tile instead of being kept twice in registers
                int32_t fm[DM], fc[DM], fs[DM], pos[DM], mt_[DM];
                uint32_t mt[DM];
                load_word<DM>(tp, W_FREEMEM, lane, fm);
                load_word<DM>(tp, W_FREECORES, lane, fc);
                load_word<DM>(tp, W_FREESLOTS, lane, fs);
                load_word<DM>(tp, W_META, lane, mt_);
#pragma unroll
                for (int j = 0; j < DM; j++) {
                    mt[j] = (uint32_t)mt_[j];
                    pos[j] = DM - 1 - j;
                }
                bool alive = true, in_ctr = false;
                float score = 0.0f;
                int32_t total = 0, free_ = 0, sums = 0;
                // One request.  The order of the walk is kept as bit masks: gt[j] = devices visited before
                // j (descending key).  FIRST: the devices still stand in K0's walk order, device i < j is
                // simply ahead of j and no keys are compared.  NUMA: numa-bind grouping (rare).
                // "First Nums fitting devices" = fitting devices with fewer than Nums fitting devices ahead
                // = popc(gt & fitmask) < Nums; the position after this request's sort = 7 - popc(gt).
                auto request = [&](auto FIRST, auto NUMA, const DevReq &rq) {
                    if (rq.flags & R_NEWCTR) {
                        if (in_ctr) score = __fadd_rn(score, container_score(total, free_, ndev - sums));
                        total = 0;
                        free_ = 0;
                        sums = 0;
                        in_ctr = true;
                    }
                    sums += rq.nums;
                    const bool usepct = rq.flags & R_USEPCT;
                    int32_t mem[DM];
                    uint32_t gt[DM], fitmask = 0u;
                    if (usepct) { // warp-uniform: the multiply + divide only where the request asks for it
#pragma unroll
                        for (int j = 0; j < DM; j++) mem[j] = pct_mem(totalmem[j], rq.pct);
                    } else {
#pragma unroll
                        for (int j = 0; j < DM; j++) mem[j] = rq.memfix;
                    }
#pragma unroll
                    for (int j = 0; j < DM; j++) {
                        const bool f = !(mt[j] & rq.blockmask) && type_ok(rq.mask_lo, rq.mask_hi, mt[j]) &&
                                       fm[j] >= mem[j] && fc[j] >= rq.coresreq;
                        fitmask |= f ? (1u << j) : 0u;
                    }
                    if constexpr (decltype(FIRST)::value) {
#pragma unroll
                        for (int j = 0; j < DM; j++) gt[j] = (1u << j) - 1u;
                    } else {
                        int32_t key[DM];
#pragma unroll
                        for (int j = 0; j < DM; j++) {
                            key[j] = (int32_t)(((mt[j] >> M_NUMA_SHIFT) & 15u) << (kKeyFreeBits + kKeyPosBits)) |
                                     ((fs[j] + kKeyFreeBias) << kKeyPosBits) | pos[j];
                            gt[j] = 0u;
                        }
#pragma unroll
                        for (int i = 0; i < DM; i++) {
#pragma unroll
                            for (int j = i + 1; j < DM; j++) { // keys are distinct (pos is a permutation)
                                const bool igtj = key[i] > key[j];
                                gt[j] |= igtj ? (1u << i) : 0u;
                                gt[i] |= igtj ? 0u : (1u << j);
                            }
                        }
                    }
                    bool ok;
                    uint32_t takemask; // fitting devices the walk takes
                    if constexpr (decltype(NUMA)::value) {
                        // the walk restarts its picks at every numa change (score.go:99-104): the first numa
                        // group (descending) holding >= Nums fitting devices wins
                        uint32_t grp[DM]; // devices of my numa group (me included)
                        int gstar = -1;
#pragma unroll
                        for (int j = 0; j < DM; j++) grp[j] = 1u << j;
#pragma unroll
                        for (int i = 0; i < DM; i++) {
#pragma unroll
                            for (int j = i + 1; j < DM; j++) {
                                const bool eq = ((mt[i] ^ mt[j]) & (15u << M_NUMA_SHIFT)) == 0u;
                                grp[i] |= eq ? (1u << j) : 0u;
                                grp[j] |= eq ? (1u << i) : 0u;
                            }
                        }
#pragma unroll
                        for (int j = 0; j < DM; j++)
                            if ((fitmask >> j & 1u) && __popc(fitmask & grp[j]) >= rq.nums)
                                gstar = max(gstar, (int)((mt[j] >> M_NUMA_SHIFT) & 15u));
                        ok = gstar >= 0;
                        takemask = 0u;
#pragma unroll
                        for (int j = 0; j < DM; j++) {
                            const bool t = ok && (fitmask >> j & 1u) && (int)((mt[j] >> M_NUMA_SHIFT) & 15u) == gstar &&
                                           __popc(gt[j] & fitmask & grp[j]) < rq.nums;
                            takemask |= t ? (1u << j) : 0u;
                        }
                    } else if constexpr (decltype(FIRST)::value) {
                        // K0's order: device i < j is ahead of j, so the walk takes the lowest Nums set bits
                        ok = __popc(fitmask) >= rq.nums;
                        takemask = 0u;
                        uint32_t m = ok ? fitmask : 0u;
                        for (int n = 0; n < rq.nums && n < DM; n++) { // warp-uniform trip count
                            takemask |= m & (0u - m);
                            m &= m - 1u;
                        }
                    } else {
                        ok = __popc(fitmask) >= rq.nums;
                        takemask = 0u;
#pragma unroll
                        for (int j = 0; j < DM; j++) {
                            const bool t = ok && (fitmask >> j & 1u) && __popc(gt[j] & fitmask) < rq.nums;
                            takemask |= t ? (1u << j) : 0u;
                        }
                    }
                    alive = alive && ok;
#pragma unroll
                    for (int j = 0; j < DM; j++) {
                        if (takemask >> j & 1u) { // score.go:168-174
                            total = wrap_add(total, count[j]);
                            free_ = wrap_add(free_, fs[j]);
                            fs[j] -= 1;
                            fm[j] = wrap_sub(fm[j], mem[j]);
                            fc[j] = wrap_sub(fc[j], rq.coresreq);
                            uint32_t m = mt[j] & ~(M_NOFREE | M_EXCL | M_FULL);
                            if (fs[j] <= 0) m |= M_NOFREE;
                            if (m & M_IS100) m |= M_EXCL;
                            if ((m & M_CORENZ) && fc[j] == 0) m |= M_FULL;
                            m |= (m >> 1) & M_BUSY370; // Used > 0 from here on
                            mt[j] = m;
                        }
                        // ascending position under the keys this request sorted by = devices with a smaller key
                        if constexpr (!decltype(FIRST)::value) pos[j] = DM - 1 - __popc(gt[j]);
                    }
                };
                using T_ = std::true_type;
                using F_ = std::false_type;
                for (int r = 0; r < pod.n_reqs; r++) {
                    const DevReq rq = load_req(&reqs[pod.req_begin + r]);
                    const bool numa = rq.flags & R_NUMA;
                    if (numa) { // rare: one variant, keys compared even for the first request
                        request(F_{}, T_{}, rq);
                    } else if (r == 0) {
                        request(T_{}, F_{}, rq); // positions stay as they are
                    } else {
                        request(F_{}, F_{}, rq);
                    }
                    if (!__any_sync(0xffffffffu, alive)) break;
                }
                if (alive && in_ctr) score = __fadd_rn(score, container_score(total, free_, ndev - sums));
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
}


// ------------------------------------------------------------------------------------
// K1 (pod-sliced): single-request pods with Nums == 1 or Nums == 2, 32 pods per register.
//
// The whole CTA works on ONE node at a time (every branch on node data is uniform); thread
// t owns pod word w = 32 pods of the class.  A device's fit word is four table rows ANDed:
// the type row, the memory row (fixed-MB pods | percentage pods), the cores row -- rows are
// chosen by how many of the class's distinct thresholds the device's free capacity admits
// (binary search per device, once per tile) -- and the flag word (exclusive / full-core /
// never).  The reverse walk becomes word algebra: first = fit & ~seen, second = fit & seen
// & ~seen2.  A candidate score is a scalar of the node (device or device pair), so the
// per-pod best is touched only when that scalar beats the thread's current minimum: after
// a few hundred nodes almost nothing does.  Node tiles (8 KB, 32 nodes) are staged into
// shared memory with cp.async.bulk + mbarrier, double buffered, and reused by every pod
// word of the CTA.  Nodes are visited in DESCENDING row order so that "strictly greater"
// keeps the highest row among equal scores inside a CTA; across CTAs the 64-bit key max does.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

constexpr int kTileInt4 = kWalkWords * 2 * kWarp; // DM = 8: 512 int4 = 8 KB per 32-node tile
constexpr int kDict = kSliceDict;
constexpr int kPairs = 28;  // device pairs (a < b) of an 8-device node
constexpr int kDevTileBytes = 32 * 8 * (int)sizeof(SliceDev); // 4 KB: the SliceDev records of 32 nodes
constexpr int kPclsTileBytes = 32 * 32;                       // 1 KB: pair classes of 32 nodes

__constant__ uint8_t kPairA[kPairs] = {0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 5, 5, 6};
__constant__ uint8_t kPairB[kPairs] = {1, 2, 3, 4, 5, 6, 7, 2, 3, 4, 5, 6, 7, 3, 4, 5, 6, 7, 4, 5, 6, 7, 5, 6, 7, 6, 7, 7};
__host__ __device__ constexpr int pair_index(int a, int b) { return a * 7 - a * (a - 1) / 2 + (b - a - 1); }

// number of sorted values v[0..n) that satisfy the monotone predicate (true prefix); the
// iteration count depends on n only, so a warp never diverges here
template <typename F>
__device__ __forceinline__ int count_prefix(const int32_t *__restrict__ v, int n, F pred) {
    int pos = 0;
    for (int step = n > 0 ? 1 << (31 - __clz(n)) : 0; step > 0; step >>= 1) {
        const int np = pos + step;
        if (np <= n && pred(__ldg(&v[np - 1]))) pos = np;
    }
    return pos;
}

// sum = a^b^c, carry = majority: one step of a carry-save adder on 32 independent bit lanes
__device__ __forceinline__ void csa(uint32_t &carry, uint32_t &sum, uint32_t a, uint32_t b, uint32_t c) {
    sum = a ^ b ^ c;
    carry = (a & b) | (a & c) | (b & c);
}

__device__ __forceinline__ void hset_insert(uint32_t *__restrict__ hset, uint32_t key, uint32_t *status, bool &full) {
    // The set only has to hold enough distinct scores to place the class boundaries well: a snapshot with more
    // distinct candidate scores than it can take (S_DICTFULL) simply stops feeding it -- a score that is not in
    // the set still gets a valid class (class_of), results stay exact.
    if (full) return; // (read once per thread at kernel start, set when an own insert gives up)
    uint32_t h = (key * 2654435761u) >> 20; // 12 bits
    for (int probe = 0; probe < 48; probe++) {
        const uint32_t cur = hset[h];
        if (cur == key) return;
        if (cur == 0u) {
            const uint32_t old = atomicCAS(&hset[h], 0u, key);
            if (old == 0u || old == key) return;
        }
        h = (h + 1) & (kSliceHash - 1);
    }
    atomicOr(status, S_DICTFULL); // a probe run of 48: the table is as good as full
    full = true;
}

// ------------------------------------------------------------------------------------
// K1t: the bit tables of a sliced class (layout: SliceTab), filled from the class's request
// records.  fill: thread = pod, one atomicOr per table the pod belongs to; prefix: thread =
// pod word, turns the per-value rows into cumulative ("threshold admitted") rows.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ int lower_bound_i32(const int32_t *__restrict__ v, int n, int32_t x) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(&v[mid]) < x) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
__global__ void __launch_bounds__(128) k1t_fill(SliceTab tb, const DevReq *__restrict__ req) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tb.n_pods) return;
    uint32_t *base = const_cast<uint32_t *>(tb.base);
    const DevReq d = load_req(&req[i]);
    const uint32_t bit = 1u << (i & 31);
    const size_t wd = (size_t)(i >> 5), Wp = (size_t)tb.Wp;
    const int r_mem = VGS_MAX_TYPES, r_pct = r_mem + tb.km + 1, r_core = r_pct + tb.kp + 1;
    for (uint32_t m = d.mask_lo; m; m &= m - 1) atomicOr(&base[(size_t)(__ffs((int)m) - 1) * Wp + wd], bit);
    for (uint32_t m = d.mask_hi; m; m &= m - 1) atomicOr(&base[(size_t)(32 + __ffs((int)m) - 1) * Wp + wd], bit);
    if (d.flags & R_USEPCT) atomicOr(&base[(size_t)(r_pct + lower_bound_i32(tb.pctv, tb.kp, d.pct) + 1) * Wp + wd], bit);
    else atomicOr(&base[(size_t)(r_mem + lower_bound_i32(tb.memv, tb.km, d.memfix) + 1) * Wp + wd], bit);
    // cores x flags: one row per flag combination f = excl | full << 1; Coresreq == 100 loses on
    // excl (score.go:128), == 0 on full (:132), > 100 always (:110)
    const int cut = lower_bound_i32(tb.corev, tb.kc, d.coresreq) + 1;
    for (int f = 0; f < 4; f++) {
        const bool lose = d.coresreq > 100 || ((f & 1) && d.coresreq == 100) || ((f & 2) && d.coresreq == 0);
        if (!lose) atomicOr(&base[(size_t)(r_core + cut * 4 + f) * Wp + wd], bit);
    }
}
// rows [r0, r0 + n) of column x, stride `step` rows apart, turned into their running OR -- eight loads in flight
// at a time (the naive loop is one dependent L2 round trip per row)
__device__ __forceinline__ void prefix_or_rows(uint32_t *__restrict__ col, size_t Wp, int r0, int n, int step) {
    uint32_t acc = 0u;
    for (int r = 0; r < n; r += 8) {
        uint32_t v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = r + u < n ? col[(size_t)(r0 + (r + u) * step) * Wp] : 0u;
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (r + u < n) col[(size_t)(r0 + (r + u) * step) * Wp] = acc |= v[u];
    }
}
__global__ void __launch_bounds__(128) k1t_prefix(SliceTab tb) {
    // thread = (pod word, chain): the memory chain, the percentage chain and the four cores x flags chains of a
    // column are independent
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) / 6, chain = (blockIdx.x * blockDim.x + threadIdx.x) % 6;
    if (x >= tb.W) return;
    uint32_t *col = const_cast<uint32_t *>(tb.base) + x;
    const size_t Wp = (size_t)tb.Wp;
    const int r_mem = VGS_MAX_TYPES, r_pct = r_mem + tb.km + 1, r_core = r_pct + tb.kp + 1;
    if (chain == 0) prefix_or_rows(col, Wp, r_mem, tb.km + 1, 1);
    else if (chain == 1) prefix_or_rows(col, Wp, r_pct, tb.kp + 1, 1);
    else prefix_or_rows(col, Wp, r_core + (chain - 2), tb.kc + 1, 4);
}
cudaError_t launch_slice_tables(const SliceTab &tb, const DevReq *dreq, cudaStream_t st) {
    if (tb.n_pods <= 0) return cudaSuccess;
    k1t_fill<<<(tb.n_pods + 127) / 128, 128, 0, st>>>(tb, dreq);
    k1t_prefix<<<(tb.W * 6 + 127) / 128, 128, 0, st>>>(tb);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------
// K1p: per batch and per sliced class, everything that depends on (device, class) but not on
// the pod: which table rows the device selects (binary searches over the class's distinct
// thresholds) and its candidate scores.  Thread = node; the searches of 800k devices run in
// parallel here instead of serially inside every score CTA.
// ------------------------------------------------------------------------------------
template <int NUMS>
__global__ void __launch_bounds__(128)
    k1p_collect(const int4 *__restrict__ walk, int32_t n_tiles, SliceTab tb, uint32_t *__restrict__ status,
                bool want_scores) { // want_scores: feed the distinct-score set (the pod-sliced kernel's classes)
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n_tiles * kWarp) return;
    const int tile = node >> 5, lane = node & 31;
    const int4 *tp = walk + (size_t)tile * kTileInt4;
    bool set_full = want_scores && (*reinterpret_cast<volatile uint32_t *>(status) & S_DICTFULL);
    int32_t totalmem[8], freemem[8], freecores[8], meta[8], wa[8], wb[8];
    load_word<8>(tp, W_TOTALMEM, lane, totalmem);
    load_word<8>(tp, W_FREEMEM, lane, freemem);
    load_word<8>(tp, W_FREECORES, lane, freecores);
    load_word<8>(tp, W_META, lane, meta);
    if (NUMS == 1) {
        load_word<8>(tp, W_SCORE1, lane, wa);
    } else {
        load_word<8>(tp, W_COUNT, lane, wa);
        load_word<8>(tp, W_FREESLOTS, lane, wb);
    }
    // rows as the score kernel indexes them (see SliceTab::row_shift)
    const int r_mem = VGS_MAX_TYPES - tb.row_shift, r_pct = r_mem + tb.km + 1, r_core = r_pct + tb.kp + 1;
    const uint32_t ndev = ((uint32_t)meta[0] >> M_NDEV_SHIFT) & 15u;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint32_t m = (uint32_t)meta[j];
        SliceDev d;
        if (m & (M_NOFREE | M_PAD)) { // Count <= Used (score.go:107) or not a device: never fits
            d.r_type = d.r_mem = d.r_pct = d.r_core = (uint16_t)r_mem; // memle row 0 is all zero
            d.a = 0u;
            d.b = NUMS == 1 ? (uint32_t)kDict : 0u;
        } else {
            const int32_t fm = freemem[j], fc = freecores[j], tm = totalmem[j];
            d.r_type = (uint16_t)(m & M_TYPE_MASK);
            d.r_mem = (uint16_t)(r_mem + count_prefix(tb.memv, tb.km, [&](int32_t v) { return v <= fm; }));
            d.r_pct = (uint16_t)(r_pct + count_prefix(tb.pctv, tb.kp, [&](int32_t v) { return pct_mem(tm, v) <= fm; }));
            // cores x flags table: row = cut * 4 + (excl | full << 1)
            d.r_core = (uint16_t)(r_core + 4 * count_prefix(tb.corev, tb.kc, [&](int32_t v) { return v <= fc; }) +
                                  ((m & M_EXCL) ? 1 : 0) + ((m & M_FULL) ? 2 : 0));
            if (NUMS == 1) {
                d.a = (uint32_t)wa[j];
                d.b = (uint32_t)kDict;
                if (want_scores) hset_insert(tb.hset, d.a, status, set_full);
            } else {
                d.a = (uint32_t)wa[j];
                d.b = (uint32_t)wb[j] | (ndev << 24);
            }
        }
        reinterpret_cast<SliceDev *>(tb.dinfo)[(size_t)node * 8 + j] = d;
    }
    if (NUMS == 2 && want_scores) {
#pragma unroll
        for (int a = 0; a < 7; a++)
#pragma unroll
            for (int b = a + 1; b < 8; b++)
                if (!(((uint32_t)meta[a] | (uint32_t)meta[b]) & (M_NOFREE | M_PAD)))
                    hset_insert(tb.hset,
                                __float_as_uint(container_score(wrap_add(wa[a], wa[b]), wrap_add(wb[a], wb[b]),
                                                                (int32_t)ndev - 2)),
                                status, set_full);
    }
}

// The Nums 1 class does not need the node as a whole: thread = DEVICE.  Eight times the threads of k1p_collect<1>,
// an eighth of its serial chain of searches and set probes each -- that kernel was latency bound at a third of the warp
// slots (thread = node leaves 21 warps per SM at 100k nodes) and heads the chain the cfg3 step waits for.
__global__ void __launch_bounds__(256)
    k1p_collect1_dev(const int4 *__restrict__ walk, int32_t n_tiles, SliceTab tb, uint32_t *__restrict__ status,
                     bool want_scores) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int node = t >> 3, j = t & 7;
    if (node >= n_tiles * kWarp) return;
    const int tile = node >> 5, lane = node & 31;
    const int32_t *tw = reinterpret_cast<const int32_t *>(walk + (size_t)tile * kTileInt4);
    auto word = [&](int w) { return __ldg(&tw[((w * 2 + (j >> 2)) * kWarp + lane) * 4 + (j & 3)]); };
    bool set_full = want_scores && (*reinterpret_cast<volatile uint32_t *>(status) & S_DICTFULL);
    const uint32_t m = (uint32_t)word(W_META);
    const int r_mem = VGS_MAX_TYPES - tb.row_shift, r_pct = r_mem + tb.km + 1, r_core = r_pct + tb.kp + 1;
    SliceDev d;
    if (m & (M_NOFREE | M_PAD)) { // Count <= Used (score.go:107) or not a device: never fits
        d.r_type = d.r_mem = d.r_pct = d.r_core = (uint16_t)r_mem; // memle row 0 is all zero
        d.a = 0u;
        d.b = (uint32_t)kDict;
    } else {
        const int32_t fm = word(W_FREEMEM), fc = word(W_FREECORES), tm = word(W_TOTALMEM);
        d.r_type = (uint16_t)(m & M_TYPE_MASK);
        d.r_mem = (uint16_t)(r_mem + count_prefix(tb.memv, tb.km, [&](int32_t v) { return v <= fm; }));
        d.r_pct = (uint16_t)(r_pct + count_prefix(tb.pctv, tb.kp, [&](int32_t v) { return pct_mem(tm, v) <= fm; }));
        d.r_core = (uint16_t)(r_core + 4 * count_prefix(tb.corev, tb.kc, [&](int32_t v) { return v <= fc; }) +
                              ((m & M_EXCL) ? 1 : 0) + ((m & M_FULL) ? 2 : 0));
        d.a = (uint32_t)word(W_SCORE1);
        d.b = (uint32_t)kDict;
        if (want_scores) hset_insert(tb.hset, d.a, status, set_full);
    }
    reinterpret_cast<SliceDev *>(tb.dinfo)[(size_t)node * 8 + j] = d;
}

// ... and its classify: the class of the device's own score
__device__ __forceinline__ uint32_t class_of(const uint32_t *__restrict__ gvals, int nv, uint32_t sb);
__global__ void __launch_bounds__(256) k1p_classify1_dev(int32_t n_tiles, SliceTab tb) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles * kWarp * 8) return;
    const int nv = (int)__ldg(&tb.gvals[kDict]);
    SliceDev *dv = reinterpret_cast<SliceDev *>(tb.dinfo) + t;
    const SliceDev d = *dv;
    if (d.r_type != d.r_mem || d.a) dv->b = d.a ? class_of(tb.gvals, nv, d.a) : (uint32_t)kDict;
}

// one CTA: compact the distinct scores out of the hash set, sort them, keep the kSliceDict
// largest (ascending) in gvals, their count in gvals[kSliceDict]
__global__ void __launch_bounds__(1024) k1p_sort(SliceTab tb) {
    __shared__ uint32_t v[kSliceHash];
    __shared__ int n_s;
    if (threadIdx.x == 0) n_s = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < kSliceHash; i += blockDim.x) {
        const uint32_t x = tb.hset[i];
        if (x) v[atomicAdd(&n_s, 1)] = x;
    }
    __syncthreads();
    const int n = n_s;
    int m = 2; // sort size: next power of two, zero padded (zeros sort first)
    while (m < n) m <<= 1;
    for (int i = n + threadIdx.x; i < m; i += blockDim.x) v[i] = 0u;
    __syncthreads();
    for (int k = 2; k <= m; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < m; i += blockDim.x) {
                const int x = i ^ j;
                if (x > i) {
                    const uint32_t a = v[i], b = v[x];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        v[i] = b;
                        v[x] = a;
                    }
                }
            }
            __syncthreads();
        }
    // <= kDict distinct scores: every one is a class of its own.  More: the class boundaries are the 16 largest
    // scores (where the per-pod bests end up, so resolution matters) and 16 evenly spaced ranks of the rest.
    const int keep = min(n, kDict);
    for (int i = threadIdx.x; i < kDict; i += blockDim.x) {
        uint32_t val = 0xffffffffu;
        if (n <= kDict) {
            if (i < keep) val = v[m - keep + i];
        } else if (i >= kDict / 2) {
            val = v[m - kDict + i];
        } else {
            const int rest = n - kDict / 2; // ranks 0 .. rest-1 of the ascending distinct scores
            val = v[m - n + (int)(((long long)(i + 1) * rest) / (kDict / 2 + 1))];
        }
        tb.gvals[i] = val;
    }
    if (threadIdx.x == 0) tb.gvals[kDict] = (uint32_t)keep;
}

// Detection class of a candidate score: the number of class boundaries below it.  below[k] holds the pods whose
// best is under boundary k, so a candidate of class k can only improve pods in below[k] -- exactly those when the
// score IS boundary k (few distinct scores: every score is one), a superset when it lies between two boundaries
// (the replay compares exactly).  Above the last boundary: kDict, the all-ones row.
__device__ __forceinline__ uint32_t class_of(const uint32_t *__restrict__ gvals, int nv, uint32_t sb) {
    int lo = 0, hi = nv;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(&gvals[mid]) < sb) lo = mid + 1;
        else hi = mid;
    }
    return lo < nv ? (uint32_t)lo : (uint32_t)kDict;
}

template <int NUMS>
__global__ void __launch_bounds__(128) k1p_classify(int32_t n_tiles, SliceTab tb) {
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n_tiles * kWarp) return;
    const int nv = (int)__ldg(&tb.gvals[kDict]);
    SliceDev *dv = reinterpret_cast<SliceDev *>(tb.dinfo) + (size_t)node * 8;
    if (NUMS == 1) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            if (dv[j].r_type != dv[j].r_mem || dv[j].a) dv[j].b = dv[j].a ? class_of(tb.gvals, nv, dv[j].a) : (uint32_t)kDict;
    } else {
        SliceDev d[8];
#pragma unroll
        for (int j = 0; j < 8; j++) d[j] = dv[j];
        uint32_t packed[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int a = 0; a < 7; a++)
#pragma unroll
            for (int b = a + 1; b < 8; b++) {
                uint32_t cl = (uint32_t)kDict;
                const bool off = (d[a].r_type == d[a].r_mem) || (d[b].r_type == d[b].r_mem);
                if (!off) {
                    const int32_t nd = (int32_t)(d[a].b >> 24);
                    cl = class_of(tb.gvals, nv,
                                  __float_as_uint(container_score(wrap_add((int32_t)d[a].a, (int32_t)d[b].a),
                                                                  wrap_add((int32_t)(d[a].b & 0xffffffu),
                                                                           (int32_t)(d[b].b & 0xffffffu)),
                                                                  nd - 2)));
                }
                const int q = pair_index(a, b);
                packed[q >> 2] |= cl << ((q & 3) * 8);
            }
        uint4 *dst = reinterpret_cast<uint4 *>(tb.pcls + (size_t)node * 32);
        dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
        dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    }
}

// shared-memory carve-up of k1_score_sliced (bytes); T = threads per CTA
struct SliceSmem {
    int tile, ptile, bar, vals, below, best, row, tab, total;
};
__host__ __device__ inline SliceSmem slice_smem(int T, int nums, int tab_rows) {
    SliceSmem m;
    int o = 0;
    m.tile = o, o += 2 * kDevTileBytes;
    m.ptile = o, o += nums == 2 ? 2 * kPclsTileBytes : 0;
    m.bar = o, o += 16;
    m.vals = o, o += (kDict + 4) * 4;
    m.below = o, o += (kDict + 1) * T * 4;
    m.best = o, o += 32 * T * 4;
    m.row = o, o += 32 * T * 2;
    m.tab = o, o += tab_rows * T * 4;
    m.total = o;
    return m;
}

// ------------------------------------------------------------------------------------
// K1 (pod-sliced): single-request pods with Nums == 1 or Nums == 2, 32 pods per register.
//
// The whole CTA works on ONE node at a time; thread t owns pod word w = 32 pods of the class.
// A device's fit word is four table rows ANDed (type, memory = fixed-MB | percentage pods,
// cores x flags) -- the rows were chosen by K1p.  The reverse walk becomes word algebra:
// first = fit & ~seen, second = fit & seen & ~seen2.  A candidate score is a scalar of the node
// (a device, or a device pair) and belongs to one of the kSliceDict score classes; below[k]
// holds my pods whose best is still under class k, so "can anybody improve here" is one AND.
// Nodes are visited in DESCENDING row order, so "strictly greater" keeps the highest row among
// equal scores inside a CTA; across CTAs the 64-bit key max does.  The 4 KB SliceDev tiles are
// staged with cp.async.bulk + mbarrier (double buffered) and shared by every pod word of the CTA.
// ------------------------------------------------------------------------------------
template <int NUMS, bool TABSMEM, bool VERIFY>
__global__ void __launch_bounds__(128)
    k1_score_sliced(int32_t n_tiles, int32_t n_nodes, int32_t row_base, SliceTab tb, int32_t n_types,
                    int32_t tiles_per_cta, int32_t phase, int32_t kScout, int32_t sub,
                    unsigned long long *__restrict__ keys,
                    int32_t *__restrict__ counts, uint32_t *__restrict__ vfit, float *__restrict__ vscore,
                    int32_t vwords) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int T = blockDim.x, tid = threadIdx.x;
    // table rows, global numbering: [type 0..63][mem 0..km][pct 0..kp][cores x flags: 4 * (kc + 1)]
    const int r_core = VGS_MAX_TYPES + tb.km + 1 + tb.kp + 1;
    const int rows_g = r_core + 4 * (tb.kc + 1);
    // the shared copy keeps only the type rows that can occur (ids < n_types)
    const int drop = TABSMEM ? VGS_MAX_TYPES - n_types : 0;
    const SliceSmem L = slice_smem(T, NUMS, TABSMEM ? rows_g - drop : 0);
    unsigned char *tilebuf = smem_raw + L.tile;   // [2][kDevTileBytes]
    unsigned char *ptilebuf = smem_raw + L.ptile; // [2][kPclsTileBytes]
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + L.bar);
    uint32_t *vals = reinterpret_cast<uint32_t *>(smem_raw + L.vals);
    // below[k][t]: my pods whose best is < vals[k] (may over-approximate, never under); row kDict = all ones
    uint32_t *below = reinterpret_cast<uint32_t *>(smem_raw + L.below);
    uint32_t *s_best = reinterpret_cast<uint32_t *>(smem_raw + L.best); // [32][T] raw score bits, 0 = none
    uint16_t *s_row = reinterpret_cast<uint16_t *>(smem_raw + L.row);   // [32][T] (virtual tile - t_lo) * 32 + lane: < 8192 by the counter cap
    const uint32_t *tabs = reinterpret_cast<const uint32_t *>(smem_raw + L.tab);

    const int w = blockIdx.y * T + tid;
    const bool active = w < tb.W;
    const int wc = min(w, tb.Wp - 1); // clamped column: inactive threads compute garbage, never publish
    const uint32_t stride = TABSMEM ? (uint32_t)T : (uint32_t)tb.Wp;
    const uint32_t col = TABSMEM ? (uint32_t)tid : (uint32_t)wc;
    // Two launches per class.  phase 0 ("scouts") walks every kScout-th tile with cold pruning
    // state and publishes its bests; phase 1 walks all the other tiles and starts from those
    // published bests, so that its detection words are almost always zero from the first node.
    // phase -1 = single launch over all tiles (small snapshots).  v = virtual tile index.
    auto tile_of = [&](int v) -> int {
        if (phase < 0) return v;
        if (phase == 0) return v * kScout;
        return (v / (kScout - 1)) * kScout + (v % (kScout - 1)) + 1;
    };
    const int n_virtual = phase < 0 ? n_tiles
                          : phase == 0 ? (n_tiles + kScout - 1) / kScout
                                       : n_tiles - (n_tiles + kScout - 1) / kScout;
    // scouts start cold (every node replays) and there are only ~100 scout tiles: `sub` CTAs share
    // each tile, one slice of its 8 four-node groups apiece, so that the launch fills the machine
    const int bx = (int)blockIdx.x / sub, lq_hi = ((int)blockIdx.x % sub + 1) * (8 / sub) - 1,
              lq_lo = ((int)blockIdx.x % sub) * (8 / sub);
    const int t_hi = n_virtual - bx * tiles_per_cta;
    const int t_lo = max(0, t_hi - tiles_per_cta);
    if (t_hi <= 0) return;
    const int n_my = t_hi - t_lo;
    const int node_hi = (tile_of(t_hi - 1) + 1) * kWarp;
    constexpr uint32_t kTx = kDevTileBytes + (NUMS == 2 ? kPclsTileBytes : 0);
    constexpr uint16_t kNotMine = 0xffffu; // s_row of a pod whose best was inherited, not found here

    auto fetch = [&](int tile, int buf) { // one thread: stage the records of 32 nodes
        mbar_expect_tx(&bar[buf], kTx);
        bulk_g2s(tilebuf + buf * kDevTileBytes, reinterpret_cast<const unsigned char *>(tb.dinfo) + (size_t)tile * kDevTileBytes,
                 kDevTileBytes, &bar[buf]);
        if (NUMS == 2)
            bulk_g2s(ptilebuf + buf * kPclsTileBytes, tb.pcls + (size_t)tile * kPclsTileBytes, kPclsTileBytes, &bar[buf]);
    };

    for (int k = tid; k <= kDict; k += T) vals[k] = __ldg(&tb.gvals[k]);
    for (int k = 0; k <= kDict; k++) below[k * T + tid] = phase == 1 ? 0u : 0xffffffffu;
    below[kDict * T + tid] = 0xffffffffu;
    for (int b = 0; b < 32; b++) {
        s_best[b * T + tid] = 0u;
        s_row[b * T + tid] = kNotMine;
    }
    if (TABSMEM) { // every row of my column with one asynchronous 4-byte copy each: one round trip, no registers
        uint32_t *tw = reinterpret_cast<uint32_t *>(smem_raw + L.tab);
        const int rows_s = rows_g - drop;
        for (int r = 0; r < rows_s; r++) {
            const uint32_t *src = &tb.base[(size_t)(r < n_types ? r : r + drop) * tb.Wp + wc];
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tw[r * T + tid]);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (TABSMEM) asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tid == 0) fetch(tile_of(t_hi - 1), 0);
    const int nv = (int)vals[kDict];
    if (phase == 1) {
        // inherit the scouts' bests.  A holder above all my rows can only be beaten by a strictly
        // larger score; one below some of my rows also by an equal score, so its score is
        // inherited one ulp lower (a local tie on a lower row is harmless: keys only grow).
        for (int b0 = 0; b0 < 32; b0 += 16) { // two rounds of 16 independent loads, not 32 dependent ones
            int op[16];
            unsigned long long kk[16];
#pragma unroll
            for (int u = 0; u < 16; u++) {
                const int i = w * 32 + b0 + u;
                op[u] = (active && i < tb.n_pods) ? __ldg(&tb.orig[i]) : -1;
            }
#pragma unroll
            for (int u = 0; u < 16; u++) kk[u] = op[u] >= 0 ? keys[op[u]] : 0ull;
#pragma unroll
            for (int u = 0; u < 16; u++) {
            const int b = b0 + u;
            const unsigned long long key = kk[u];
            uint32_t init = 0u;
            if (key) {
                const uint32_t gs = (uint32_t)(key >> 32) ^ 0x80000000u; // positive scores
                const int grow = (int)(uint32_t)key - row_base;
                init = grow >= node_hi ? gs : gs - 1u;
            }
            s_best[b * T + tid] = init;
            int c = 0; // classes with vals[k] <= init
            for (int step = 32; step > 0; step >>= 1)
                if (c + step <= nv && vals[c + step - 1] <= init) c += step;
            if (c < nv) below[c * T + tid] |= 1u << b; // the pod is under every class from c upward
            }
        }
        for (int k = 1; k < nv; k++) below[k * T + tid] |= below[(k - 1) * T + tid];
    }

    // K1p stored the rows already in this kernel's numbering: one multiply-add per table access
    const uint32_t stride_b = stride * 4u;
    const unsigned char *tab_col = reinterpret_cast<const unsigned char *>(TABSMEM ? tabs : tb.base) + (size_t)col * 4u;
    // the column address as an opaque 64-bit value: one IMAD.WIDE (row * stride + column) feeds the load
    unsigned long long tab_col_g = reinterpret_cast<unsigned long long>(tab_col);
    asm volatile("" : "+l"(tab_col_g));
    auto TAB = [&](uint32_t row) -> uint32_t {
        if (TABSMEM) return *reinterpret_cast<const uint32_t *>(tab_col + row * stride_b);
        return __ldg(reinterpret_cast<const uint32_t *>(tab_col_g + (unsigned long long)row * stride_b));
    };
    const unsigned char *below_tid = reinterpret_cast<const unsigned char *>(below + tid);
    const uint32_t tstride_b = (uint32_t)T * 4u;
    auto BELOW = [&](uint32_t cls) -> uint32_t { // below[cls * T + tid]
        return *reinterpret_cast<const uint32_t *>(below_tid + cls * tstride_b);
    };
    auto fit_of = [&](const uint4 &d) -> uint32_t {
        return TAB(d.x & 0xffffu) & (TAB(d.x >> 16) | TAB(d.y & 0xffffu)) & TAB(d.y >> 16);
    };
    // pods in `word` get candidate (sb, node); nodes only ever decrease, so equal scores keep
    // the earlier (higher) row.  A pod whose best reaches sb leaves every below[k] with vals[k] <= sb.
    auto raise = [&](uint32_t word, uint32_t sb, int vrow) { // vrow: the node's position among this CTA's virtual tiles
        uint32_t imp = 0u;
        for (uint32_t m = word; m; m &= m - 1) {
            const int b = __ffs((int)m) - 1;
            if (sb > s_best[b * T + tid]) {
                s_best[b * T + tid] = sb;
                s_row[b * T + tid] = (uint16_t)vrow;
                imp |= 1u << b;
            }
        }
        for (int k = 0; k < nv && vals[k] <= sb; k++) below[k * T + tid] &= ~imp;
    };
    auto mark = [&](uint32_t word, int node, uint32_t sb) { // verify mode: score of every fitting pair
        for (uint32_t mth = word; mth; mth &= mth - 1) {
            const int p = tb.orig[w * 32 + __ffs((int)mth) - 1];
            vscore[(size_t)p * n_nodes + node] = __uint_as_float(sb);
        }
    };

    uint32_t ones = 0u, twos = 0u, cnt[kSliceCountPlanes]; // fit counter: ones, twos, cnt[k] = bit k (k >= 2)
#pragma unroll
    for (int k = 0; k < kSliceCountPlanes; k++) cnt[k] = 0u;

    for (int it = 0; it < n_my; it++) {
        const int tile = tile_of(t_hi - 1 - it), buf = it & 1;
        const int vbase = (n_my - 1 - it) * kWarp; // 16-bit row of the tile's first node (s_row): independent of the scout stride
        if (tid == 0 && it + 1 < n_my) fetch(tile_of(t_hi - 2 - it), buf ^ 1); // released by the last sync
        mbar_wait(&bar[buf], (uint32_t)((it >> 1) & 1));
        const uint4 *dv = reinterpret_cast<const uint4 *>(tilebuf + buf * kDevTileBytes); // [32 nodes][8]
        const uint8_t *pc = ptilebuf + buf * kPclsTileBytes;                              // [32 nodes][32]

        // 32 nodes, descending, in groups of four.  Phase 1 is branch-free word algebra for all
        // four nodes (independent chains): fit words, the walk as masks, and a detection word
        // "could any pod of mine improve here?" against below[] as it stands (stale = larger =
        // safe).  Phase 2 replays, in order and against the live below[], only the nodes whose
        // detection word is non-zero.
        for (int lq = lq_hi; lq >= lq_lo; lq--) {
            uint32_t xs[4], hit[4];
            if (NUMS == 1) {
                uint32_t nw[4][8];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int l = lq * 4 + u;
                    uint32_t rem = 0xffffffffu, h = VERIFY ? 0xffffffffu : 0u;
#pragma unroll
                    for (int j = 0; j < 8; j++) { // walk order: the first fitting device is taken
                        const uint4 d = dv[l * 8 + j];
                        const uint32_t f = fit_of(d);
                        nw[u][j] = f & rem;
                        rem &= ~f;
                        h |= nw[u][j] & BELOW(d.w);
                    }
                    xs[u] = ~rem;
                    hit[u] = h;
                }
#pragma unroll
                for (int u = 3; u >= 0; u--) {
                    const int l = lq * 4 + u, node = tile * kWarp + l;
                    if (hit[u]) {
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                            const uint4 d = dv[l * 8 + j];
                            if (VERIFY && active) mark(nw[u][j], node, d.z);
                            const uint32_t h = nw[u][j] & BELOW(d.w);
                            if (h) raise(h, d.z, vbase + l);
                        }
                    }
                }
            } else {
                uint32_t first[4][8], second[4][8];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int l = lq * 4 + u;
                    uint32_t seen = 0u, seen2 = 0u, h = VERIFY ? 0xffffffffu : 0u;
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const uint32_t f = fit_of(dv[l * 8 + j]);
                        first[u][j] = f & ~seen;
                        second[u][j] = f & seen & ~seen2;
                        seen2 |= seen & f;
                        seen |= f;
                    }
                    const uint4 c0 = *reinterpret_cast<const uint4 *>(pc + l * 32);
                    const uint4 c1 = *reinterpret_cast<const uint4 *>(pc + l * 32 + 16);
                    const uint32_t cw[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                    for (int a = 0; a < 7; a++)
#pragma unroll
                        for (int b = a + 1; b < 8; b++) {
                            const int q = pair_index(a, b);
                            const uint32_t cl = (cw[q >> 2] >> ((q & 3) * 8)) & 0xffu;
                            h |= first[u][a] & second[u][b] & BELOW(cl);
                        }
                    xs[u] = seen2; // the node fits a pod iff two devices fit
                    hit[u] = h;
                }
#pragma unroll
                for (int u = 3; u >= 0; u--) {
                    const int l = lq * 4 + u, node = tile * kWarp + l;
                    if (hit[u]) {
#pragma unroll
                        for (int a = 0; a < 7; a++) {
                            if (first[u][a] == 0u) continue;
#pragma unroll
                            for (int b = a + 1; b < 8; b++) {
                                const uint32_t wab = first[u][a] & second[u][b]; // pods taking exactly a and b
                                if (wab == 0u) continue;
                                const uint4 da = dv[l * 8 + a], db = dv[l * 8 + b];
                                const uint32_t sb = __float_as_uint(container_score(
                                    wrap_add((int32_t)da.z, (int32_t)db.z),
                                    wrap_add((int32_t)(da.w & 0xffffffu), (int32_t)(db.w & 0xffffffu)),
                                    (int32_t)(da.w >> 24) - 2));
                                if (VERIFY && active) mark(wab, node, sb);
                                const uint32_t h = wab & BELOW((uint32_t)pc[l * 32 + pair_index(a, b)]);
                                if (h) raise(h, sb, vbase + l);
                            }
                        }
                    }
                }
            }
            if (VERIFY && active) {
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int node = tile * kWarp + lq * 4 + u;
                    for (uint32_t mth = xs[u]; mth; mth &= mth - 1) {
                        const int p = tb.orig[w * 32 + __ffs((int)mth) - 1];
                        atomicOr(&vfit[(size_t)p * vwords + (node >> 5)], 1u << (node & 31));
                    }
                }
            }
            // counter += xs[0] + xs[1] + xs[2] + xs[3] (bit-sliced, carry-save)
            uint32_t ca, sa, cb, cf;
            csa(ca, sa, ones, xs[0], xs[1]);
            csa(cb, ones, sa, xs[2], xs[3]);
            csa(cf, twos, twos, ca, cb);
#pragma unroll
            for (int k = 2; k < kSliceCountPlanes; k++) {
                const uint32_t t2 = cnt[k] & cf;
                cnt[k] ^= cf;
                cf = t2;
            }
        }
        __syncthreads(); // everyone is done with this tile buffer
    }

    if (!active) return;
    for (int b = 0; b < 32; b++) {
        const int i = w * 32 + b;
        if (i >= tb.n_pods) break;
        int32_t c = (int32_t)((ones >> b) & 1u) | ((int32_t)((twos >> b) & 1u) << 1);
#pragma unroll
        for (int k = 2; k < kSliceCountPlanes; k++) c |= (int32_t)((cnt[k] >> b) & 1u) << k;
        if (c == 0) continue; // none of my nodes fits this pod
        const int p = __ldg(&tb.orig[i]);
        atomicAdd(&counts[p], c);
        const uint16_t r = s_row[b * T + tid];
        if (r == kNotMine) continue; // nothing here beat the inherited best
        // scores of single-request pods are positive: ordered bits = bits | sign bit
        const unsigned long long key = ((unsigned long long)(s_best[b * T + tid] | 0x80000000u) << 32) |
                                       (unsigned long long)(uint32_t)(row_base + tile_of(t_lo + (int)(r >> 5)) * kWarp + (int)(r & 31u));
        atomicMax(&keys[p], key);
    }
}

// ------------------------------------------------------------------------------------
// K1 (node-sliced): ONE pod word (<= 32 pods) of a sliced class, lane = node.  The transposed use of
// the same word algebra: every lane builds the fit words of its own node's 8 devices from the
// word's table column (a few hundred bytes, kept in shared memory) and the K1p records, then the
// warp settles each pod of the word with one ballot + max-reduction over its 32 nodes.  About
// 1 warp instruction per (pod, node) -- 3x the pod-sliced kernel, 15x below the per-pair one -- which
// makes it the home of the words a class has beyond a multiple of 32 (a pod-sliced warp costs the
// same whether 1 or 32 of its words are in use) and of classes of a few words.
// ------------------------------------------------------------------------------------
template <int NUMS, bool VERIFY>
__global__ void __launch_bounds__(128)
    k1_score_nodesliced(int32_t n_virtual, int32_t tile_mul, int32_t n_nodes, int32_t row_base, SliceTab tb, int32_t word0,
                        unsigned long long *__restrict__ keys, int32_t *__restrict__ counts,
                        uint32_t *__restrict__ vfit, float *__restrict__ vscore, int32_t vwords) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *tcol = reinterpret_cast<uint32_t *>(smem_raw); // the word's column, rows in K1p's numbering
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int word = word0 + (int)blockIdx.y;
    const int rows_g = VGS_MAX_TYPES + tb.km + 1 + tb.kp + 1 + 4 * (tb.kc + 1);
    const int drop = tb.row_shift, kept_types = VGS_MAX_TYPES - drop;
    for (int r = threadIdx.x; r < rows_g - drop; r += blockDim.x)
        tcol[r] = __ldg(&tb.base[(size_t)(r < kept_types ? r : r + drop) * tb.Wp + word]);
    __syncthreads();
    const int npw = min(kWarp, tb.n_pods - word * kWarp); // pods of this word
    const int32_t *orig = tb.orig + word * kWarp;
    unsigned long long acckey = 0ull; // lane k: pod k of the word, over all this CTA's tiles
    int32_t acccnt = 0;

    // tiles v * tile_mul, v < n_virtual (tile_mul = 1: every tile; the scout stride: the scout tiles of a pod-sliced class)
    for (int v = (int)blockIdx.x * (int)(blockDim.x >> 5) + warp; v < n_virtual; v += (int)gridDim.x * (int)(blockDim.x >> 5)) {
        const int tile = v * tile_mul;
        const int row = tile * kWarp + lane;
        const uint4 *dv = tb.dinfo + (size_t)row * 8;
        uint4 d[8];
#pragma unroll
        for (int j = 0; j < 8; j++) d[j] = __ldg(&dv[j]);
        uint32_t f[8];
#pragma unroll
        for (int j = 0; j < 8; j++)
            f[j] = tcol[d[j].x & 0xffffu] & (tcol[d[j].x >> 16] | tcol[d[j].y & 0xffffu]) & tcol[d[j].y >> 16];
        uint32_t fitw; // pods this node fits
        uint32_t t1[8], t2[8];
        if (NUMS == 1) {
            uint32_t rem = 0xffffffffu;
#pragma unroll
            for (int j = 0; j < 8; j++) { // walk order: the first fitting device is taken
                t1[j] = f[j] & rem;
                rem &= ~f[j];
            }
            fitw = ~rem;
        } else {
            uint32_t seen = 0u, seen2 = 0u;
#pragma unroll
            for (int j = 0; j < 8; j++) { // the first two fitting devices are taken
                t1[j] = f[j] & ~seen;
                t2[j] = f[j] & seen & ~seen2;
                seen2 |= seen & f[j];
                seen |= f[j];
            }
            fitw = seen2;
#pragma unroll
            for (int j = 0; j < 8; j++) t1[j] = (t1[j] | t2[j]) & fitw; // taken by the pods that fit
        }
        const uint32_t row0 = (uint32_t)(row_base + tile * kWarp);
        unsigned long long mykey = 0ull;
        int32_t mycnt = 0;
        for (int b = 0; b < npw; b++) {
            const uint32_t bit = 1u << b;
            const bool fit = fitw & bit;
            float score;
            if (NUMS == 1) {
                uint32_t sb = 0u;
#pragma unroll
                for (int j = 0; j < 8; j++) sb = (t1[j] & bit) ? d[j].z : sb; // K0's score of that device alone
                score = __uint_as_float(sb);
            } else {
                int32_t total = 0, free_ = 0, nd = 2;
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (t1[j] & bit) { // score.go:168-170
                        total = wrap_add(total, (int32_t)d[j].z);
                        free_ = wrap_add(free_, (int32_t)(d[j].w & 0xffffffu));
                        nd = (int32_t)(d[j].w >> 24);
                    }
                score = fit ? container_score(total, free_, nd - 2) : 0.0f;
            }
            if (VERIFY) {
                const int op = orig[b];
                const unsigned bal = __ballot_sync(0xffffffffu, fit);
                if (lane == 0) vfit[(size_t)op * vwords + tile] = bal;
                if (row < n_nodes) vscore[(size_t)op * n_nodes + row] = fit ? score : 0.0f;
            }
            publish(fit, score, b, lane, row0, mykey, mycnt);
        }
        acckey = mykey > acckey ? mykey : acckey;
        acccnt += mycnt;
    }
    flush_small<4>(acckey, acccnt, lane, warp, npw, orig, nullptr, keys, counts);
}

// ------------------------------------------------------------------------------------
// K2: merge shard candidates (after the allgather)
// ------------------------------------------------------------------------------------
__global__ void k2_merge(const unsigned long long *__restrict__ gk, const int32_t *__restrict__ gc,
                         int32_t n_shards, int32_t stride, int32_t cstride, int32_t n_pods,
                         unsigned long long *__restrict__ keys, int32_t *__restrict__ counts) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    unsigned long long best = 0ull;
    int32_t c = 0;
    for (int g = 0; g < n_shards; g++) {
        unsigned long long k = gk[(size_t)g * stride + p];
        best = k > best ? k : best;
        if (gc) c += gc[(size_t)g * cstride + p];
    }
    keys[p] = best;
    if (gc) counts[p] = c;
}

// ------------------------------------------------------------------------------------
// K3: resolve.  One thread per pod re-runs the winner node from the RAW planes in the
// plain way (sort, reverse walk, mutate) and emits the picks in take order
// (score.go:138-144,175).  Its score must equal the key's score: self-check.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64)
    k3_resolve(RawPlanes raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
               const vgs_pod *__restrict__ pods, const DevReq *__restrict__ reqs,
               const int32_t *__restrict__ pick_begin, int32_t n_pods,
               unsigned long long *__restrict__ keys, int32_t *__restrict__ counts,
               vgs_result *__restrict__ results, vgs_pick *__restrict__ picks,
               uint32_t *__restrict__ status, int32_t *__restrict__ commit_rowlist,
               vgs_result *__restrict__ host_results, vgs_pick *__restrict__ host_picks, bool keep_candidates,
               uint32_t *__restrict__ host_status, int4 *__restrict__ commit_walk, int32_t commit_dm) {
    // host_results / host_picks != nullptr (small batches): the records are ALSO stored straight into
    // the caller-facing pinned host buffers (posted PCIe writes), which spares the two D2H copies
    // commit_rowlist != nullptr (sequential drain, one pod per launch): the winner's picks are
    // also folded into the raw planes (addPod, scheduler.go:280-297,394) and the touched row is
    // named for re-preparation
    // host_status != nullptr (single-CTA launches only): the last pod to finish also forwards the status
    // word to the host mirror, so that the download of such a batch is a stream synchronisation and nothing else
    __shared__ int s_left;
    if (threadIdx.x == 0) s_left = min((int)blockDim.x, n_pods - (int)(blockIdx.x * blockDim.x));
    __syncthreads();
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    auto leave = [&]() {
        if (host_status) {
            __threadfence(); // my status bits (if any) are out before I count myself done
            if (atomicSub(&s_left, 1) == 1) host_status[0] = *reinterpret_cast<volatile uint32_t *>(status);
        }
    };
    if (commit_rowlist) commit_rowlist[p] = -1;
    const unsigned long long key = keys[p];
    vgs_result res;
    res.fit_count = counts[p];
    // K3 is the last reader of the candidate block: leave it zero for the next score pass
    if (!keep_candidates) { // (kept for the shard exchange: k2x_merge clears them)
        keys[p] = 0ull;
        counts[p] = 0;
    }
    const vgs_pod pod = pods[p];
    const int32_t pick_lo = pick_begin[pod.req_begin], pick_hi = pick_begin[pod.req_begin + pod.n_reqs];
    // picks of a pod without a winner on this shard are all zero
    auto zero_picks = [&]() {
        vgs_pick z;
        z.slot = z.usedmem = z.usedcores = 0;
        for (int32_t i = pick_lo; i < pick_hi; i++) {
            picks[i] = z;
            if (host_picks) host_picks[i] = z;
        }
    };
    if (key == 0ull) {
        res.best_row = -1;
        res.score = 0.0f;
        res.n_picks = 0;
        results[p] = res;
        if (host_results) host_results[p] = res;
        zero_picks();
        leave();
        return;
    }
    const int32_t grow = (int32_t)(uint32_t)key;
    const uint32_t kbits = unorder_bits((uint32_t)(key >> 32));
    res.best_row = grow;
    res.score = __uint_as_float(kbits);
    res.n_picks = pick_hi - pick_lo;
    results[p] = res;
    if (host_results) host_results[p] = res;
    const int32_t row = grow - row_base;
    if (row < 0 || row >= n_nodes) { // another shard owns the winner
        zero_picks();
        leave();
        return;
    }

    const size_t base = (size_t)row * dmax;
    ResolveOut ro;
    resolve_row(
        raw, row, dmax, pod.n_reqs, [&](int r) { return reqs[pod.req_begin + r]; },
        [&](int r, int t, const vgs_pick &pk) {
            const int32_t pb = pick_begin[pod.req_begin + r];
            picks[pb + t] = pk;
            if (host_picks) host_picks[pb + t] = pk;
        },
        ro);
    const bool alive = ro.alive;
    const float score = ro.score;
    if (commit_rowlist && alive) {
        for (int d = 0; d < ro.nd; d++) { // the mutated copies ARE the committed state
            const_cast<int32_t *>(raw.p[VGS_PLANE_USED])[base + d] = ro.used[d];
            const_cast<int32_t *>(raw.p[VGS_PLANE_USEDMEM])[base + d] = ro.usedmem[d];
            const_cast<int32_t *>(raw.p[VGS_PLANE_USEDCORES])[base + d] = ro.usedcores[d];
        }
        commit_rowlist[p] = row;
        if (commit_walk) { // re-prepare the touched row right here: the next pod's score kernel reads it
            if (commit_dm == 8) prepare_node<8>(raw, row, n_nodes, dmax, commit_walk, status);
            else prepare_node<12>(raw, row, n_nodes, dmax, commit_walk, status);
        }
    }
    if (!alive || __float_as_uint(score) != kbits) atomicOr(status, S_MISMATCH);
    leave();
}

// K3 for whole batches: one WARP per pod (lane = device slot of the winner node, resolve_row_warp) instead of one
// thread per pod walking local-memory arrays: 4096 pods resolve in one wave of short warps.  Same outputs as
// k3_resolve without its commit / host-mirror extras (those are for single-CTA batches and the drain).
__global__ void __launch_bounds__(256)
    k3_resolve_warp(RawPlanes raw, int32_t n_nodes, int32_t dmax, int32_t row_base, const vgs_pod *__restrict__ pods,
                    const DevReq *__restrict__ reqs, const int32_t *__restrict__ pick_begin, int32_t n_pods,
                    unsigned long long *__restrict__ keys, int32_t *__restrict__ counts, vgs_result *__restrict__ results,
                    vgs_pick *__restrict__ picks, uint32_t *__restrict__ status, bool keep_candidates) {
    const int p = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = (int)(threadIdx.x & 31u);
    if (p >= n_pods) return; // whole warps
    unsigned long long key = 0ull;
    vgs_result res;
    res.fit_count = 0;
    if (lane == 0) {
        key = keys[p];
        res.fit_count = counts[p];
        if (!keep_candidates) { // last reader of the candidate block: leave it zero for the next score pass
            keys[p] = 0ull;
            counts[p] = 0;
        }
    }
    key = __shfl_sync(0xffffffffu, key, 0);
    const vgs_pod pod = pods[p];
    const int32_t pick_lo = pick_begin[pod.req_begin], pick_hi = pick_begin[pod.req_begin + pod.n_reqs];
    const uint32_t kbits = unorder_bits((uint32_t)(key >> 32));
    const int32_t grow = (int32_t)(uint32_t)key, row = grow - row_base;
    if (lane == 0) {
        res.best_row = key ? grow : -1;
        res.score = key ? __uint_as_float(kbits) : 0.0f;
        res.n_picks = key ? pick_hi - pick_lo : 0;
        results[p] = res;
    }
    if (key == 0ull || row < 0 || row >= n_nodes) { // nobody fits, or another shard owns the winner: zero picks
        vgs_pick z;
        z.slot = z.usedmem = z.usedcores = 0;
        for (int32_t i = pick_lo + lane; i < pick_hi; i += kWarp) picks[i] = z;
        return;
    }
    bool alive;
    float score;
    int32_t nd, used, usedmem, usedcores;
    resolve_row_warp(
        raw, row, dmax, pod.n_reqs, [&](int r) { return reqs[pod.req_begin + r]; },
        [&](int r, int t, const vgs_pick &pk) { picks[pick_begin[pod.req_begin + r] + t] = pk; }, alive, score, nd, used,
        usedmem, usedcores);
    if ((!alive || __float_as_uint(score) != kbits) && lane == 0) atomicOr(status, S_MISMATCH);
}

// ------------------------------------------------------------------------------------
// K4: usage deltas (what addPod/delPod change between two Filters, scheduler.go:280-297)
// ------------------------------------------------------------------------------------
__global__ void k4_apply_delta(RawPlanesMut raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
                               const vgs_delta *__restrict__ deltas, int32_t n,
                               int32_t *__restrict__ rowlist, uint32_t *__restrict__ status) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    vgs_delta d = deltas[i];
    int32_t row = d.row - row_base;
    if (row < 0 || row >= n_nodes) { // not this shard's row: ignore
        rowlist[i] = -1;
        return;
    }
    if (d.slot < 0 || d.slot >= dmax || d.slot >= raw.ndev[row]) {
        atomicOr(status, S_DELTA);
        rowlist[i] = -1;
        return;
    }
    size_t x = (size_t)row * dmax + d.slot;
    atomicAdd(&raw.p[VGS_PLANE_USED][x], d.d_used);
    atomicAdd(&raw.p[VGS_PLANE_USEDMEM][x], d.d_usedmem);
    atomicAdd(&raw.p[VGS_PLANE_USEDCORES][x], d.d_usedcores);
    rowlist[i] = row;
}

// ------------------------------------------------------------------------------------
// Host-pinned planes -> device planes, read straight over PCIe by the SMs (zero-copy).  On some
// boxes of the pool a 26 MB cudaMemcpyAsync from pinned memory ran at only ~15 GB/s (43-51 GB/s
// for 256 MB); many warps with 16-byte loads in flight reach ~50 GB/s at this size everywhere.
// ------------------------------------------------------------------------------------
struct PlaneCopy {
    const int4 *src[VGS_NUM_PLANES + 1];
    int4 *dst[VGS_NUM_PLANES + 1];
    int32_t n16[VGS_NUM_PLANES + 1]; // int4 elements per plane
    int32_t tail[VGS_NUM_PLANES + 1]; // leftover int32 elements
};
__global__ void __launch_bounds__(256) k_copy_planes(PlaneCopy pc) {
    const int f = blockIdx.y;
    const int4 *__restrict__ s = pc.src[f];
    int4 *__restrict__ d = pc.dst[f];
    const int n = pc.n16[f];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d[i] = s[i];
    if (blockIdx.x == 0 && threadIdx.x < pc.tail[f]) {
        const int32_t *s32 = reinterpret_cast<const int32_t *>(s + n);
        reinterpret_cast<int32_t *>(d + n)[threadIdx.x] = s32[threadIdx.x];
    }
}

cudaError_t launch_copy_planes(const int32_t *const src[VGS_NUM_PLANES], const int32_t *src_ndev,
                               int32_t *const dst[VGS_NUM_PLANES], int32_t *dst_ndev, size_t cells, size_t n_nodes,
                               cudaStream_t st) {
    PlaneCopy pc;
    for (int f = 0; f <= VGS_NUM_PLANES; f++) {
        const int32_t *s = f < VGS_NUM_PLANES ? src[f] : src_ndev;
        int32_t *d = f < VGS_NUM_PLANES ? dst[f] : dst_ndev;
        const size_t n = f < VGS_NUM_PLANES ? cells : n_nodes;
        if (((uintptr_t)s | (uintptr_t)d) & 15) return cudaErrorMisalignedAddress;
        pc.src[f] = reinterpret_cast<const int4 *>(s);
        pc.dst[f] = reinterpret_cast<int4 *>(d);
        pc.n16[f] = (int32_t)(n / 4);
        pc.tail[f] = (int32_t)(n % 4);
    }
    k_copy_planes<<<dim3(148 * 2, VGS_NUM_PLANES + 1), 256, 0, st>>>(pc);
    return cudaGetLastError();
}

// The same from planes in the narrow wire format (vgs_snapshot_upload_narrow): elements of 1, 2 or 4 bytes (signed),
// widened to the int32 device planes on the way.  The pass is bound by the bytes that cross PCIe -- 14 instead of
// 32 per device when the counters, core percentages, numa and type ids travel as bytes.  A thread takes 16 source
// bytes per turn (one load over PCIe, many in flight) and stores 1, 2 or 4 int4.
struct NarrowCopy {
    const int4 *src[VGS_NUM_PLANES + 1];
    int32_t *dst[VGS_NUM_PLANES + 1];
    int32_t n16[VGS_NUM_PLANES + 1];  // 16-byte source chunks per plane
    int32_t tail[VGS_NUM_PLANES + 1]; // leftover elements
    int32_t width[VGS_NUM_PLANES + 1];
};
__device__ __forceinline__ int4 widen_bytes(uint32_t w) {
    return make_int4((int32_t)(int8_t)(w & 0xffu), (int32_t)(int8_t)((w >> 8) & 0xffu), (int32_t)(int8_t)((w >> 16) & 0xffu),
                     (int32_t)(int8_t)(w >> 24));
}
__device__ __forceinline__ int4 widen_halves(uint32_t a, uint32_t b) {
    return make_int4((int32_t)(int16_t)(a & 0xffffu), (int32_t)(int16_t)(a >> 16), (int32_t)(int16_t)(b & 0xffffu),
                     (int32_t)(int16_t)(b >> 16));
}
__global__ void __launch_bounds__(256) k_copy_planes_narrow(NarrowCopy pc) {
    const int f = blockIdx.y;
    const int4 *__restrict__ s = pc.src[f];
    int4 *__restrict__ d = reinterpret_cast<int4 *>(pc.dst[f]);
    const int n = pc.n16[f], w = pc.width[f];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int4 v = s[i];
        if (w == 4) {
            d[i] = v;
        } else if (w == 2) {
            d[2 * (size_t)i] = widen_halves((uint32_t)v.x, (uint32_t)v.y);
            d[2 * (size_t)i + 1] = widen_halves((uint32_t)v.z, (uint32_t)v.w);
        } else {
            d[4 * (size_t)i] = widen_bytes((uint32_t)v.x);
            d[4 * (size_t)i + 1] = widen_bytes((uint32_t)v.y);
            d[4 * (size_t)i + 2] = widen_bytes((uint32_t)v.z);
            d[4 * (size_t)i + 3] = widen_bytes((uint32_t)v.w);
        }
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < pc.tail[f]) { // fewer than 16 elements
        const size_t e = (size_t)n * (16 / w) + threadIdx.x;
        const char *sb = reinterpret_cast<const char *>(s);
        int32_t x;
        if (w == 4) x = reinterpret_cast<const int32_t *>(sb)[e];
        else if (w == 2) x = reinterpret_cast<const int16_t *>(sb)[e];
        else x = reinterpret_cast<const int8_t *>(sb)[e];
        pc.dst[f][e] = x;
    }
}

cudaError_t launch_copy_planes_narrow(const void *const src[VGS_NUM_PLANES + 1], const uint8_t width[VGS_NUM_PLANES + 1],
                                      int32_t *const dst[VGS_NUM_PLANES], int32_t *dst_ndev, size_t cells, size_t n_nodes,
                                      cudaStream_t st) {
    NarrowCopy pc;
    for (int f = 0; f <= VGS_NUM_PLANES; f++) {
        int32_t *d = f < VGS_NUM_PLANES ? dst[f] : dst_ndev;
        const size_t n = f < VGS_NUM_PLANES ? cells : n_nodes;
        const int w = width[f];
        if (w != 1 && w != 2 && w != 4) return cudaErrorInvalidValue;
        if (((uintptr_t)src[f] | (uintptr_t)d) & 15) return cudaErrorMisalignedAddress;
        pc.src[f] = reinterpret_cast<const int4 *>(src[f]);
        pc.dst[f] = d;
        pc.n16[f] = src[f] ? (int32_t)(n / (16 / w)) : 0; // (a plane left out: the resident one stays)
        pc.tail[f] = src[f] ? (int32_t)(n % (16 / w)) : 0;
        pc.width[f] = w;
    }
    k_copy_planes_narrow<<<dim3(148 * 2, VGS_NUM_PLANES + 1), 256, 0, st>>>(pc);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------
static int g_sm_count[64];
static int sm_count(int dev) {
    if (dev < 0 || dev >= 64) return 148;
    if (g_sm_count[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        g_sm_count[dev] = n;
    }
    return g_sm_count[dev];
}

cudaError_t launch_prepare(int dm, const RawPlanes &raw, int32_t n_nodes, int32_t dmax, int4 *walk,
                           int32_t n_tiles, uint32_t *status, const int32_t *rowlist,
                           int32_t n_list, cudaStream_t st) {
    int n = rowlist ? n_list : n_tiles * kWarp;
    if (n <= 0) return cudaSuccess;
    int block = 128, grid = (n + block - 1) / block;
    const int smem = rowlist ? 0 : (block / kWarp) * kWalkWords * dm * kWarp * (int)sizeof(int32_t); // 32 / 48 KB
    if (dm == 8) {
        k0_prepare<8><<<grid, block, smem, st>>>(raw, n_nodes, dmax, walk, n_tiles, status, rowlist, n_list);
    } else {
        static std::once_flag once[64];
        int dev = 0;
        cudaGetDevice(&dev);
        std::call_once(once[dev & 63], [] {
            cudaFuncSetAttribute(k0_prepare<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * kWalkWords * 12 * kWarp * 4);
        });
        k0_prepare<12><<<grid, block, smem, st>>>(raw, n_nodes, dmax, walk, n_tiles, status, rowlist, n_list);
    }
    return cudaGetLastError();
}

// cudaOccupancyMaxActiveBlocksPerMultiprocessor costs microseconds and plan() asks several times
// per launch; the answer only depends on (kernel, device, block, dynamic smem)
template <typename K>
static cudaError_t cached_occupancy(int *occ, K kernel, int dev, int block, int smem) {
    struct Entry {
        const void *k;
        int dev, block, smem, occ;
    };
    static std::mutex mu;
    static std::vector<Entry> cache;
    const void *kp = reinterpret_cast<const void *>(kernel);
    {
        std::lock_guard<std::mutex> lk(mu);
        for (const Entry &e : cache)
            if (e.k == kp && e.dev == dev && e.block == block && e.smem == smem) {
                *occ = e.occ;
                return cudaSuccess;
            }
    }
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, kernel, block, smem);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    cache.push_back(Entry{kp, dev, block, smem, *occ});
    return cudaSuccess;
}

template <typename K>
static cudaError_t plan(K kernel, int dev, int32_t n_tiles, int32_t n_pods, int max_warps,
                        int smem_per_pod, ScoreLaunch *L) {
    // small snapshots: fewer warps per CTA so that every SM still gets work
    int sms = sm_count(dev);
    int wpc = max_warps;
    while (wpc > 1 && (n_tiles + wpc - 1) / wpc < 2 * sms) wpc >>= 1;
    L->warps_per_cta = wpc;
    L->block = wpc * kWarp;
    L->n_tile_groups = (n_tiles + wpc - 1) / wpc;
    int occ = 1;
    // pod chunk: enough items for >= 8 rounds per resident CTA, chunk a multiple of 32
    int chunk = 512;
    for (;;) {
        L->smem = chunk * smem_per_pod;
        cudaError_t e = cached_occupancy(&occ, kernel, dev, L->block, L->smem);
        if (e != cudaSuccess) return e;
        if (occ < 1) occ = 1;
        long long resident = (long long)sms * occ;
        long long items = (long long)L->n_tile_groups * ((n_pods + chunk - 1) / chunk);
        if (items >= 8 * resident || chunk <= 32) break;
        chunk >>= 1;
    }
    if (chunk > ((n_pods + 31) / 32) * 32) chunk = ((n_pods + 31) / 32) * 32;
    L->chunk_len = chunk;
    L->smem = chunk * smem_per_pod;
    L->n_chunks = (n_pods + chunk - 1) / chunk;
    long long items = (long long)L->n_tile_groups * L->n_chunks;
    long long resident = (long long)sms * occ;
    L->grid = (int)(items < resident ? items : resident);
    if (L->grid < 1) L->grid = 1;
    return cudaSuccess;
}

cudaError_t plan_score_launch(const void *kernel, int dev, int32_t n_tiles, int32_t n_pods, int max_warps,
                              int smem_per_pod, ScoreLaunch *L) {
    return plan(kernel, dev, n_tiles, n_pods, max_warps, smem_per_pod, L);
}

static const FilterOne kNoFilter{};

template <int DM, bool NUMS1>
static cudaError_t launch_single_t(int dev, const ScoreArgs &a, const DevReq *reqs, const int32_t *orig,
                                   int32_t n_pods, cudaStream_t st) {
    ScoreLaunch L;
    cudaError_t e;
    if (a.vfit) {
        auto k = k1_score_single<DM, NUMS1, true, false>;
        if ((e = plan(k, dev, a.n_tiles, n_pods, 8, (int)sizeof(DevReq), &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, L.smem, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, orig, n_pods, L,
                                          a.keys, a.counts, a.vfit, a.vscore, a.vwords, kNoFilter);
    } else {
        auto k = k1_score_single<DM, NUMS1, false, false>;
        if ((e = plan(k, dev, a.n_tiles, n_pods, 8, (int)sizeof(DevReq), &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, L.smem, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, orig, n_pods, L,
                                          a.keys, a.counts, nullptr, nullptr, 0, kNoFilter);
    }
    return cudaGetLastError();
}

template <int DM, bool NUMS1>
static cudaError_t launch_filter_single_t(int dev, const ScoreArgs &a, const FilterOne &fo, cudaStream_t st) {
    ScoreLaunch L;
    cudaError_t e;
    auto k = k1_score_single<DM, NUMS1, false, true>;
    if ((e = plan(k, dev, a.n_tiles, 1, 8, (int)sizeof(DevReq), &L)) != cudaSuccess) return e;
    k<<<L.grid, L.block, L.smem, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, nullptr, nullptr, 1, L, a.keys,
                                      a.counts, nullptr, nullptr, 0, fo);
    return cudaGetLastError();
}

// a Filter call whose pod is one request without a numa group to choose: the per-pair kernel streams 4 words per
// device where the walk kernel streams 6
cudaError_t launch_filter_single(int dm, bool nums1, int dev, const ScoreArgs &a, const FilterOne &fo,
                                 cudaStream_t st) {
    if (dm == 8)
        return nums1 ? launch_filter_single_t<8, true>(dev, a, fo, st) : launch_filter_single_t<8, false>(dev, a, fo, st);
    return nums1 ? launch_filter_single_t<12, true>(dev, a, fo, st) : launch_filter_single_t<12, false>(dev, a, fo, st);
}

cudaError_t launch_score_single(int dm, bool nums1, int dev, const ScoreArgs &a, const DevReq *reqs,
                                const int32_t *orig, int32_t n_pods, cudaStream_t st) {
    if (n_pods <= 0 || a.n_tiles <= 0) return cudaSuccess;
    if (dm == 8)
        return nums1 ? launch_single_t<8, true>(dev, a, reqs, orig, n_pods, st)
                     : launch_single_t<8, false>(dev, a, reqs, orig, n_pods, st);
    return nums1 ? launch_single_t<12, true>(dev, a, reqs, orig, n_pods, st)
                 : launch_single_t<12, false>(dev, a, reqs, orig, n_pods, st);
}

template <int DM>
static cudaError_t launch_general_t(int dev, const ScoreArgs &a, const DevReq *reqs, const GenPod *pods,
                                    int32_t n_pods, cudaStream_t st) {
    ScoreLaunch L;
    cudaError_t e;
    if (a.vfit) {
        auto k = k1_score_general<DM, true>;
        if ((e = plan(k, dev, a.n_tiles, n_pods, 4, 0, &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, 0, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, pods, n_pods, L,
                                     a.keys, a.counts, a.vfit, a.vscore, a.vwords);
    } else {
        auto k = k1_score_general<DM, false>;
        if ((e = plan(k, dev, a.n_tiles, n_pods, 4, 0, &L)) != cudaSuccess) return e;
        k<<<L.grid, L.block, 0, st>>>(a.walk, a.n_tiles, a.n_nodes, a.row_base, reqs, pods, n_pods, L,
                                     a.keys, a.counts, nullptr, nullptr, 0);
    }
    return cudaGetLastError();
}

cudaError_t launch_score_general(int dm, int dev, const ScoreArgs &a, const DevReq *reqs,
                                 const GenPod *pods, int32_t n_pods, cudaStream_t st) {
    if (n_pods <= 0 || a.n_tiles <= 0) return cudaSuccess;
    return dm == 8 ? launch_general_t<8>(dev, a, reqs, pods, n_pods, st)
                   : launch_general_t<12>(dev, a, reqs, pods, n_pods, st);
}


// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device), to the opt-in limit: the
// attribute never has to follow the batch, and two contexts on one GPU cannot lower it under each other
template <typename K>
static cudaError_t allow_big_smem(K kernel, int dev) {
    struct Entry {
        const void *k;
        int dev;
    };
    static std::mutex mu;
    static std::vector<Entry> done;
    const void *kp = reinterpret_cast<const void *>(kernel);
    std::lock_guard<std::mutex> lk(mu);
    for (const Entry &e : done)
        if (e.k == kp && e.dev == dev) return cudaSuccess;
    int optin = 0;
    cudaError_t e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin)) != cudaSuccess) return e;
    done.push_back(Entry{kp, dev});
    return cudaSuccess;
}

template <int NUMS>
static cudaError_t launch_sliced_t(int dev, const ScoreArgs &a, const SliceTab &tb_in, int n_types, uint32_t *status,
                                   cudaStream_t st, const vgs_tuning &tune, vgs_class_plan *plan, cudaStream_t st_ns,
                                   cudaEvent_t ev_ns, cudaEvent_t ev_sc, cudaStream_t st_sc) {
    const int sms = sm_count(dev);
    cudaError_t e;
    SliceTab tb = tb_in;
    // Split the class's pod words between the two kernels.  A pod-sliced warp costs ~0.37 warp
    // instructions per (pod, node) when all its 32 words are in use and the same total when only one is;
    // a node-sliced word costs ~1 per (pod, node).  So a few words beyond a multiple of 32 (or a class of
    // few words) go to the node-sliced kernel, the rest to pod-sliced warps.
    // Measured at 100k nodes: ~9 us per node-sliced word; a whole pod-sliced chain (sort, classify, scouts,
    // main) ~200 us, one more warp in an existing chain ~50 us.  vgs_tuning.node_words overrides both limits.
    const int W_all = tb.W, n_pods_all = tb.n_pods;
    // a class of its own (< 32 words): chain ~ (155 + 44 n) us against ~6.2 n us per node-sliced word,
    // n = nodes / 100k (measured at 100k and 400k nodes) -> break-even 7 + 25 / n words
    int small_limit = 7 + (int)(2500000ll / (a.n_nodes > 0 ? a.n_nodes : 1));
    if (small_limit > 24) small_limit = 24;
    const int node_words_max = tune.node_words >= 0 ? tune.node_words : (W_all < 32 ? small_limit : 6);
    int ns_words = W_all % 32;
    if (ns_words > node_words_max) ns_words = 0;
    const int W_main = W_all - ns_words;
    tb.W = W_main;
    tb.n_pods = n_pods_all < W_main * 32 ? n_pods_all : W_main * 32;
    int warps = (tb.W + 31) / 32;
    if (warps > 4) warps = 4;
    if (warps < 1) warps = 1;
    const int T = warps * 32;
    const int gy = (tb.W + T - 1) / T;
    if (n_types < 1) n_types = 1;
    if (n_types > VGS_MAX_TYPES) n_types = VGS_MAX_TYPES;
    const int rows_s = n_types + (tb.km + 1) + (tb.kp + 1) + 4 * (tb.kc + 1);
    // tables in shared memory up to tables_smem_kb per CTA (else global / L1); measured on cfg3:
    // 8 KB 0.498 ms, 16-24 KB 0.472, 32-64 KB 0.461
    const int tab_kb = tune.tables_smem_kb > 0 ? tune.tables_smem_kb : 48;
    bool tabsmem = tune.tables_in_smem >= 0 ? tune.tables_in_smem != 0 : (size_t)rows_s * T * 4 <= (size_t)tab_kb * 1024;
    if (slice_smem(T, NUMS, rows_s).total > 200 * 1024) tabsmem = false; // forced on, but it cannot fit
    tb.row_shift = tabsmem ? VGS_MAX_TYPES - n_types : 0;
    plan->words_pod_sliced = W_main;
    plan->words_node_sliced = ns_words;
    plan->tables_in_smem = W_main > 0 ? (tabsmem ? 1 : 0) : 0;

    // every scout-th tile is walked ahead of the main pass: about a hundred scout tiles (measured
    // optimum on 100k nodes: stride 24-48), vgs_tuning.scout_stride overrides
    int scout_stride = tune.scout_stride >= 2 ? tune.scout_stride : a.n_tiles / 96;
    if (scout_stride < 4 && tune.scout_stride < 2) scout_stride = 4;
    if (scout_stride > 32 && tune.scout_stride < 2) scout_stride = 32;
    // who walks them: the node-sliced kernel (default) or a cold pod-sliced launch
    // (bit 0: the Nums 1 class, bit 1: the Nums 2 class, bit 2: in order on the class's own stream)
    // (the caller clears the bits for batches with a class of more than kScoutNsMaxWords words: see there)
    const bool scouts_ns = ((tune.scouts_node_sliced >> (NUMS - 1)) & 1) != 0 && W_main > 0 && a.n_tiles >= 256;
    // (bit 3: on a stream of the class's own, beside sort / classify; else bit 2 clear: on st_ns)
    cudaStream_t sc_stream = st;
    if (scouts_ns && ev_ns && ev_sc) {
        if ((tune.scouts_node_sliced & 8) && st_sc && st_sc != st) sc_stream = st_sc;
        else if (!(tune.scouts_node_sliced & 4) && st_ns != st) sc_stream = st_ns;
    }

    // K1p: table rows + candidate scores per device, distinct scores -> classes
    const int nthreads = a.n_tiles * kWarp;
    if (W_main > 0) {
        if ((e = cudaMemsetAsync(tb.hset, 0, kSliceHash * sizeof(uint32_t), st)) != cudaSuccess) return e;
        if (NUMS == 1) k1p_collect1_dev<<<(nthreads * 8 + 255) / 256, 256, 0, st>>>(a.walk, a.n_tiles, tb, status, true);
        else k1p_collect<NUMS><<<(nthreads + 127) / 128, 128, 0, st>>>(a.walk, a.n_tiles, tb, status, true);
        // the node-sliced words need the records only (not the score classes): they can start now, on
        // another stream, instead of queueing behind the class's sort / classify / scout / main launches
        if (ns_words > 0 && st_ns != st && ev_ns && !(scouts_ns && sc_stream == st_ns)) { // (else the same wait is queued with the scouts)
            if ((e = cudaEventRecord(ev_ns, st)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(st_ns, ev_ns, 0)) != cudaSuccess) return e;
        }
        if (scouts_ns) {
            // Scouts: the bests of every scout_stride-th tile, which the main pass starts from.  They need K1p's
            // records only, so the node-sliced kernel walks those tiles for ALL the class's words while sort and
            // classify run (a pod-sliced scout launch starts cold -- every node improves for most pods -- and
            // took 84 us for 3 % of the nodes; this takes ~15)
            if (sc_stream != st) {
                if ((e = cudaEventRecord(ev_ns, st)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(sc_stream, ev_ns, 0)) != cudaSuccess) return e;
            }
            SliceTab ts = tb; // (tb.W / n_pods = the pod-sliced words)
            const int rows_sc = VGS_MAX_TYPES - ts.row_shift + (ts.km + 1) + (ts.kp + 1) + 4 * (ts.kc + 1);
            const int scsmem = rows_sc * 4;
            const int n_scout = (a.n_tiles + scout_stride - 1) / scout_stride;
            auto gos = [&](auto kernel) -> cudaError_t {
                if (scsmem > 48 * 1024 && (e = allow_big_smem(kernel, dev)) != cudaSuccess) return e;
                kernel<<<dim3((unsigned)((n_scout + 3) / 4), (unsigned)W_main), 128, scsmem, sc_stream>>>(
                    n_scout, scout_stride, a.n_nodes, a.row_base, ts, 0, a.keys, a.counts, a.vfit, a.vscore, a.vwords);
                return cudaGetLastError();
            };
            e = a.vfit ? gos(k1_score_nodesliced<NUMS, true>) : gos(k1_score_nodesliced<NUMS, false>);
            if (e != cudaSuccess) return e;
            if (sc_stream != st && (e = cudaEventRecord(ev_sc, sc_stream)) != cudaSuccess) return e;
            plan->launches += 1;
        }
        k1p_sort<<<1, 1024, 0, st>>>(tb);
        if (NUMS == 1) k1p_classify1_dev<<<(nthreads * 8 + 255) / 256, 256, 0, st>>>(a.n_tiles, tb);
        else k1p_classify<NUMS><<<(nthreads + 127) / 128, 128, 0, st>>>(a.n_tiles, tb);
        plan->launches += 3;
    } else { // a class of a few words: the node-sliced kernel needs the records only
        if (NUMS == 1) k1p_collect1_dev<<<(nthreads * 8 + 255) / 256, 256, 0, st>>>(a.walk, a.n_tiles, tb, status, false);
        else k1p_collect<NUMS><<<(nthreads + 127) / 128, 128, 0, st>>>(a.walk, a.n_tiles, tb, status, false);
        plan->launches += 1;
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;

    const int scout = scout_stride;
    // cfg3: 1 -> 0.459 ms, 2 -> 0.449, 4 -> 0.472, 8 -> 0.477 (set-up per CTA)
    int scout_sub_max = tune.scout_split;
    if (scout_sub_max != 1 && scout_sub_max != 2 && scout_sub_max != 4 && scout_sub_max != 8) scout_sub_max = 2;
    const int smem = slice_smem(T, NUMS, tabsmem ? rows_s : 0).total;
    auto go = [&](auto kernel) -> cudaError_t {
        if ((e = allow_big_smem(kernel, dev)) != cudaSuccess) return e;
        int occ = 1;
        if ((e = cached_occupancy(&occ, kernel, dev, T, smem)) != cudaSuccess) return e;
        if (occ < 1) occ = 1;
        // CTAs of one launch: one wave; at least 2 tiles each so the per-CTA set-up and flush stay
        // small, at most (2^planes - 1) nodes each (the bit-sliced fit counter)
        auto launch = [&](int n_virtual, int phase, int min_tiles, int32_t *tiles_out) -> cudaError_t {
            if (n_virtual <= 0) return cudaSuccess;
            long long want = (long long)sms * occ / gy;
            if (tune.max_ctas > 0 && want > tune.max_ctas) want = tune.max_ctas;
            if (want < 1) want = 1;
            int tiles_per = (int)((n_virtual + want - 1) / want);
            if (tiles_per < min_tiles) tiles_per = min_tiles;
            const int cap = ((1 << kSliceCountPlanes) - 1) / kWarp;
            if (tiles_per > cap) tiles_per = cap;
            int gx = (n_virtual + tiles_per - 1) / tiles_per;
            int sub = 1; // scouts: split each tile over a few CTAs while the launch is under one wave
            if (phase == 0)
                while (sub < scout_sub_max && (long long)gx * sub * 2 <= want) sub *= 2;
            gx *= sub;
            if (phase == 0) plan->scout_split = sub;
            *tiles_out = tiles_per;
            kernel<<<dim3(gx, gy), T, smem, st>>>(a.n_tiles, a.n_nodes, a.row_base, tb, n_types, tiles_per, phase,
                                                  scout, sub, a.keys, a.counts, a.vfit, a.vscore, a.vwords);
            plan->launches += 1;
            return cudaGetLastError();
        };
        if (a.n_tiles < 256) return launch(a.n_tiles, -1, 2, &plan->tiles_per_cta_main); // small snapshot (< 8k nodes): one launch
        const int scouts = (a.n_tiles + scout - 1) / scout;
        plan->scout_stride = scout;
        if (scouts_ns) { // the scout tiles were walked by the node-sliced kernel (above): the main pass inherits its bests
            if (sc_stream != st && (e = cudaStreamWaitEvent(st, ev_sc, 0)) != cudaSuccess) return e;
        } else if ((e = launch(scouts, 0, 1, &plan->tiles_per_cta_scout)) != cudaSuccess) {
            return e;
        }
        return launch(a.n_tiles - scouts, 1, 2, &plan->tiles_per_cta_main);
    };
    if (W_main > 0) {
        if (a.vfit) e = tabsmem ? go(k1_score_sliced<NUMS, true, true>) : go(k1_score_sliced<NUMS, false, true>);
        else e = tabsmem ? go(k1_score_sliced<NUMS, true, false>) : go(k1_score_sliced<NUMS, false, false>);
        if (e != cudaSuccess) return e;
    }
    if (ns_words > 0) { // needs K1p's records (row numbering included), not its score classes
        SliceTab tn = tb;
        tn.W = W_all;
        tn.n_pods = n_pods_all;
        const int rows_all = VGS_MAX_TYPES - tn.row_shift + (tn.km + 1) + (tn.kp + 1) + 4 * (tn.kc + 1);
        const int nsmem = rows_all * 4;
        auto gon = [&](auto kernel) -> cudaError_t {
            if (nsmem > 48 * 1024 && (e = allow_big_smem(kernel, dev)) != cudaSuccess) return e;
            int occ = 1;
            if ((e = cached_occupancy(&occ, kernel, dev, 128, nsmem)) != cudaSuccess) return e;
            if (occ < 1) occ = 1;
            long long gx = ((long long)sms * occ + ns_words - 1) / ns_words; // one wave over all words
            const long long need = (a.n_tiles + 3) / 4;
            if (gx > need) gx = need;
            if (gx < 1) gx = 1;
            cudaStream_t sn = (W_main > 0 && st_ns != st && ev_ns) ? st_ns : st;
            kernel<<<dim3((unsigned)gx, (unsigned)ns_words), 128, nsmem, sn>>>(a.n_tiles, 1, a.n_nodes, a.row_base, tn, W_main,
                                                                             a.keys, a.counts, a.vfit, a.vscore, a.vwords);
            plan->launches += 1;
            return cudaGetLastError();
        };
        e = a.vfit ? gon(k1_score_nodesliced<NUMS, true>) : gon(k1_score_nodesliced<NUMS, false>);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_score_sliced(int nums, int dev, const ScoreArgs &a, const SliceTab &tb, int n_types,
                                uint32_t *status, cudaStream_t st, const vgs_tuning &tune, vgs_class_plan *plan,
                                cudaStream_t st_ns, cudaEvent_t ev_ns, cudaEvent_t ev_sc, cudaStream_t st_sc) {
    plan->launches = 0;
    if (tb.n_pods <= 0 || a.n_tiles <= 0) return cudaSuccess;
    if (!st_ns) st_ns = st;
    return nums == 1 ? launch_sliced_t<1>(dev, a, tb, n_types, status, st, tune, plan, st_ns, ev_ns, ev_sc, st_sc)
                     : launch_sliced_t<2>(dev, a, tb, n_types, status, st, tune, plan, st_ns, ev_ns, ev_sc, st_sc);
}

// K2x: merge of whole exchange blocks.  Every shard has resolved its OWN best node per pod (K3 with
// keep_candidates) and the shards' blocks [keys | counts | picks] have been all-gathered; thread = pod
// picks the winning shard (max key = max score, then highest global row), sums the fit counts, writes the
// result record and copies the winner's picks.  One collective per pass instead of two.
__global__ void k2x_merge(const unsigned char *__restrict__ gathered, int32_t n_shards, size_t stride_bytes,
                          int32_t max_pods, size_t picks_off, const vgs_pod *__restrict__ pods,
                          const int32_t *__restrict__ pick_begin, int32_t n_pods,
                          unsigned long long *__restrict__ keys, int32_t *__restrict__ counts,
                          vgs_result *__restrict__ results, vgs_pick *__restrict__ picks) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    unsigned long long best = 0ull;
    int32_t total = 0, win = 0;
    for (int g = 0; g < n_shards; g++) {
        const unsigned char *blk = gathered + (size_t)g * stride_bytes;
        const unsigned long long k = reinterpret_cast<const unsigned long long *>(blk)[p];
        total += reinterpret_cast<const int32_t *>(blk + (size_t)max_pods * 8)[p];
        if (k > best) {
            best = k;
            win = g;
        }
    }
    keys[p] = 0ull; // the local block goes back to its resting state
    counts[p] = 0;
    const vgs_pod pod = pods[p];
    const int32_t lo = pick_begin[pod.req_begin], hi = pick_begin[pod.req_begin + pod.n_reqs];
    vgs_result res;
    res.fit_count = total;
    if (best == 0ull) {
        res.best_row = -1;
        res.score = 0.0f;
        res.n_picks = 0;
    } else {
        res.best_row = (int32_t)(uint32_t)best;
        res.score = __uint_as_float(unorder_bits((uint32_t)(best >> 32)));
        res.n_picks = hi - lo;
    }
    results[p] = res;
    // the winner's picks (all zero when nobody fits: every shard zero-filled its own)
    const vgs_pick *src = reinterpret_cast<const vgs_pick *>(gathered + (size_t)win * stride_bytes + picks_off);
    for (int32_t i = lo; i < hi; i++) picks[i] = src[i];
}

cudaError_t launch_exchange_merge(const void *gathered, int32_t n_shards, size_t stride_bytes, int32_t max_pods,
                                  size_t picks_off, const vgs_pod *pods, const int32_t *pick_begin, int32_t n_pods,
                                  unsigned long long *keys, int32_t *counts, vgs_result *results, vgs_pick *picks,
                                  cudaStream_t st) {
    if (n_pods <= 0) return cudaSuccess;
    k2x_merge<<<(n_pods + 127) / 128, 128, 0, st>>>(reinterpret_cast<const unsigned char *>(gathered), n_shards,
                                                   stride_bytes, max_pods, picks_off, pods, pick_begin, n_pods, keys,
                                                   counts, results, picks);
    return cudaGetLastError();
}

cudaError_t launch_merge(const unsigned long long *gk, const int32_t *gc, int32_t n_shards, int32_t stride,
                         int32_t cstride, int32_t n_pods, unsigned long long *keys, int32_t *counts,
                         cudaStream_t st) {
    if (n_pods <= 0) return cudaSuccess;
    k2_merge<<<(n_pods + 127) / 128, 128, 0, st>>>(gk, gc, n_shards, stride, cstride, n_pods, keys, counts);
    return cudaGetLastError();
}

cudaError_t launch_resolve(const RawPlanes &raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
                           const vgs_pod *pods, const DevReq *reqs, const int32_t *pick_begin,
                           int32_t n_pods, unsigned long long *keys, int32_t *counts,
                           vgs_result *results, vgs_pick *picks, uint32_t *status, cudaStream_t st,
                           int32_t *commit_rowlist, vgs_result *host_results, vgs_pick *host_picks,
                           bool keep_candidates, uint32_t *host_status, int4 *commit_walk, int commit_dm) {
    if (n_pods <= 0) return cudaSuccess;
    if (n_pods > 64 && !commit_rowlist && !host_results && !host_picks) { // a whole batch: one warp per pod
        k3_resolve_warp<<<(n_pods + 7) / 8, 256, 0, st>>>(raw, n_nodes, dmax, row_base, pods, reqs, pick_begin, n_pods, keys,
                                                         counts, results, picks, status, keep_candidates);
        return cudaGetLastError();
    }
    if (n_pods > 64) host_status = nullptr; // one CTA only (see k3_resolve)
    k3_resolve<<<(n_pods + 63) / 64, 64, 0, st>>>(raw, n_nodes, dmax, row_base, pods, reqs, pick_begin,
                                                 n_pods, keys, counts, results, picks, status, commit_rowlist,
                                                 host_results, host_picks, keep_candidates, host_status,
                                                 commit_walk, commit_dm);
    return cudaGetLastError();
}

cudaError_t launch_delta(const RawPlanesMut &raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
                         const vgs_delta *deltas, int32_t n, int32_t *rowlist, uint32_t *status,
                         cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    k4_apply_delta<<<(n + 127) / 128, 128, 0, st>>>(raw, n_nodes, dmax, row_base, deltas, n, rowlist, status);
    return cudaGetLastError();
}

} // namespace vgs
