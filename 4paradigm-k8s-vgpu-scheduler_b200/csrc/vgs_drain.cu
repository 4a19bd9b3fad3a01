// vgs_drain.cu -- the batch-scale drain: P sequential placements (Filter + addPod per pod,
// pkg/scheduler/scheduler.go:372-394 with the usage folding of :280-297) without a pass over the whole snapshot
// per pod.
//
// A commit changes ONE row.  So for a chunk of C consecutive pods:
//   k_drain_eval    every pod of the chunk against every node of the snapshot as the chunk finds it (thread =
//                   node, the pair or the walk evaluation per pod): the ordered score bits of every (pod, node),
//                   0 = does not fit, into a [pods][nodes] scratch matrix -- no atomics;
//   k_drain_select  per pod (and segment of its row of the matrix), one CTA: the number of fitting nodes, kDrainTop
//                   candidate keys (score, row) in descending order, and a bound: no row outside the list has a
//                   larger key than that.  (Every thread keeps its best few, the CTA merges them; what a thread or
//                   the merge drops only raises the bound.)
//   k_drain_commit  ONE CTA walks the chunk in pod order.  T = the rows touched inside the chunk so far (<= C).
//                   For pod q, every row outside T still has its chunk-start key, so
//                       winner(q) = max( best live key over T,  first candidate of q's list outside T )
//                   provided that maximum reaches the list's bound; the live keys over T are evaluated now (thread =
//                   touched row, from a copy of the row's tile words in shared memory).  Otherwise the list says
//                   nothing about the rest of the snapshot and the CTA scans the whole walk copy for that pod (rare:
//                   vgs_plan.drain_rescans).
//                       fit_count(q) = chunk-start count - #{T rows fitting at chunk start} + #{T rows fitting now},
//                   the middle term read from the matrix when the chunk is done.
//                   The winner is resolved by one warp from the raw planes (picks), folded into them, and its row
//                   prepared again -- in the walk copy and in the shared-memory copy.
// Every result equals what the per-pod chain (score kernel over all nodes per pod) produces.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vgpuscore.h"
#include "vgs_device.cuh"
#include "vgs_kernels.h"
#include "vgs_layout.h"
#include "vgs_pair_eval.cuh"
#include "vgs_prepare.cuh"
#include "vgs_resolve.cuh"
#include "vgs_walk_eval.cuh"

namespace vgs {

// what the kernels keep of the chunk's pods in shared memory
struct ChunkPods {
    GenPod pod[kDrainChunk];
    DevReq first[kDrainChunk];       // the pod's first request: all a pair pod has
    uint32_t pair[kDrainChunk / 32]; // bit k: pod k takes the pair evaluation (vgs_pair_eval.cuh)
};
__device__ __forceinline__ void chunk_pods_load(ChunkPods &cp, const GenPod *__restrict__ pods,
                                                const DevReq *__restrict__ reqs, int n_pods) {
    if (threadIdx.x < kDrainChunk / 32) cp.pair[threadIdx.x] = 0u;
    __syncthreads();
    for (int k = threadIdx.x; k < n_pods; k += blockDim.x) {
        const GenPod p = pods[k];
        const DevReq f = load_req(&reqs[p.req_begin]);
        cp.pod[k] = p;
        cp.first[k] = f;
        if (pair_pod(p.n_reqs, f)) atomicOr(&cp.pair[k >> 5], 1u << (k & 31));
    }
    __syncthreads();
}
constexpr unsigned long long kTouched = 1ull; // list entry whose row has been touched inside the chunk (no key is 1: score bits 0 = no fit)
__device__ __forceinline__ unsigned long long key_of(float score, uint32_t row) {
    return ((unsigned long long)order_bits(__float_as_uint(score)) << 32) | row;
}
// 64-bit max over a warp
__device__ __forceinline__ unsigned long long warp_max64(unsigned long long v) {
    const uint32_t hi = (uint32_t)(v >> 32);
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t lo = hi == mh ? (uint32_t)v : 0u;
    const uint32_t ml = __reduce_max_sync(0xffffffffu, lo);
    return ((unsigned long long)mh << 32) | ml;
}

// ------------------------------------------------------------------------------------
// chunk-start pass, part 1: the score bits of every (pod, node).  Warp = node tile, item = (tile, part of the pods).
// ------------------------------------------------------------------------------------
constexpr int kScoreThreads = 256;
template <int DM>
__global__ void __launch_bounds__(kScoreThreads)
    k_drain_eval(const int4 *__restrict__ walk, int32_t n_tiles, const DevReq *__restrict__ reqs,
                 const GenPod *__restrict__ pods, int32_t n_pods, int32_t psplit, uint32_t *__restrict__ ob) {
    __shared__ ChunkPods cp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kScoreThreads / 32;
    chunk_pods_load(cp, pods, reqs, n_pods);
    uint32_t one = 1u;
    asm volatile("" : "+r"(one));
    const size_t stride = (size_t)n_tiles * kWarp;
    const int per = (n_pods + psplit - 1) / psplit;
    const long long n_items = (long long)n_tiles * psplit;
    for (long long item = (long long)blockIdx.x * kWarps + warp; item < n_items; item += (long long)gridDim.x * kWarps) {
        const int tile = (int)(item / psplit), part = (int)(item % psplit);
        const int k0 = part * per, k1 = min(n_pods, k0 + per);
        const int4 *tp = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
        uint32_t *out = ob + (size_t)tile * kWarp + lane;
        bool any_pair = false, any_walk = false;
        for (int k = k0; k < k1; k++) {
            const bool pr = (cp.pair[k >> 5] >> (k & 31)) & 1u;
            any_pair |= pr, any_walk |= !pr;
        }
        if (any_pair) {
            PairNode<DM> nd;
            pair_node_load<DM, true>(tp, lane, nd);
            for (int k = k0; k < k1; k++) {
                if (!((cp.pair[k >> 5] >> (k & 31)) & 1u)) continue;
                bool fit;
                float score;
                pair_eval<DM>(nd, cp.first[k], fit, score);
                out[k * stride] = fit ? order_bits(__float_as_uint(score)) : 0u;
            }
        }
        if (any_walk) {
            WalkNode<DM> nd;
            walk_node_load<DM, true>(tp, lane, nd);
            for (int k = k0; k < k1; k++) {
                if ((cp.pair[k >> 5] >> (k & 31)) & 1u) continue;
                const GenPod pod = cp.pod[k];
                bool alive;
                float score;
                walk_eval<DM, true, false>(
                    nd, pod, [&](int r) { return load_req(&reqs[pod.req_begin + r]); }, one, nullptr, false, alive, score);
                out[k * stride] = alive ? order_bits(__float_as_uint(score)) : 0u;
            }
        }
    }
}

// ------------------------------------------------------------------------------------
// chunk-start pass, part 2: CTA (seg, pod) = candidates, bound and fit count of one segment of the pod's row of
// the matrix.  Output per (pod, seg): kDrainTop keys descending, then the bound.
// ------------------------------------------------------------------------------------
constexpr int kSelectThreads = 512;
constexpr int kSelectMine = 4;            // keys a thread keeps (the rows of a thread lie kSelectThreads apart)
constexpr int kListWords = kDrainTop + 1; // keys + bound
__global__ void __launch_bounds__(kSelectThreads)
    k_drain_select(const uint32_t *__restrict__ ob, int32_t n_tiles, int32_t row_base, int32_t n_segs,
                   unsigned long long *__restrict__ topk, int32_t *__restrict__ counts) {
    constexpr int kWarps = kSelectThreads / 32;
    __shared__ unsigned long long s_best[kWarps];
    __shared__ int32_t s_cnt[kWarps];
    const int seg = blockIdx.x, pod = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t stride = (size_t)n_tiles * kWarp;
    const int n = n_tiles * kWarp;
    const int r0 = (int)((long long)n * seg / n_segs), r1 = (int)((long long)n * (seg + 1) / n_segs);
    const uint32_t *src = ob + (size_t)pod * stride;
    unsigned long long t[kSelectMine]; // this thread's best, descending
#pragma unroll
    for (int i = 0; i < kSelectMine; i++) t[i] = 0ull;
    unsigned long long dropped = 0ull; // the largest key this thread has let go
    int32_t cnt = 0;
    for (int r = r0 + (int)threadIdx.x; r < r1; r += kSelectThreads) {
        const uint32_t bits = __ldg(&src[r]);
        if (bits == 0u) continue;
        cnt++;
        unsigned long long v = ((unsigned long long)bits << 32) | (uint32_t)(row_base + r);
#pragma unroll
        for (int i = 0; i < kSelectMine; i++) {
            const unsigned long long hi = v > t[i] ? v : t[i];
            v = v > t[i] ? t[i] : v;
            t[i] = hi;
        }
        dropped = v > dropped ? v : dropped;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) s_cnt[warp] = cnt;
    unsigned long long *list = topk + ((size_t)pod * n_segs + seg) * kListWords;
    // kDrainTop rounds of "largest head of all threads" (the owner moves on to its next), one more for the bound
    for (int r = 0; r <= kDrainTop; r++) {
        const unsigned long long mine = r < kDrainTop ? t[0] : (t[0] > dropped ? t[0] : dropped);
        const unsigned long long m = warp_max64(mine);
        __syncthreads(); // the previous round's readers of s_best (and, first round, the writers of s_cnt)
        if (lane == 0) s_best[warp] = m;
        __syncthreads();
        unsigned long long best = 0ull;
#pragma unroll
        for (int w = 0; w < kWarps; w++) best = s_best[w] > best ? s_best[w] : best;
        if (threadIdx.x == 0) list[r] = best;
        if (r < kDrainTop && t[0] == best && best != 0ull) { // keys are unique: one owner
#pragma unroll
            for (int i = 0; i + 1 < kSelectMine; i++) t[i] = t[i + 1];
            t[kSelectMine - 1] = 0ull;
        }
    }
    if (threadIdx.x == 0) {
        int32_t c = 0;
        for (int w = 0; w < kWarps; w++) c += s_cnt[w];
        counts[pod * n_segs + seg] = c;
    }
}

// pull a line towards this SM ahead of its use
__device__ __forceinline__ void warm(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

struct DrainChunk {
    int4 *walk;
    int32_t n_tiles, n_nodes, row_base, dmax;
    RawPlanes raw;
    const DevReq *reqs;
    const GenPod *pods;        // the chunk's pods, batch order
    const int32_t *pick_begin; // batch-wide, per request
    int32_t n_pods;
    const uint32_t *ob;             // [pod][n_tiles * 32] chunk-start score bits
    const unsigned long long *topk; // [pod][n_segs][kListWords]
    const int32_t *counts;          // [pod][n_segs]
    int32_t n_segs;
    vgs_result *results; // the chunk's
    vgs_pick *picks;     // the batch's
    uint32_t *status;
    int32_t *stats; // [0] += pods that needed the full scan
};

constexpr int kDrainThreads = 256;
constexpr int kDrainReqBuf = 384; // requests of the chunk's multi-request pods kept in shared memory

// ------------------------------------------------------------------------------------
// the chunk in pod order, one CTA
// ------------------------------------------------------------------------------------
template <int DM>
__global__ void __launch_bounds__(kDrainThreads) k_drain_commit(const __grid_constant__ DrainChunk a) {
    constexpr int kTile4 = kWalkWords * (DM / 4) * kWarp; // int4 per tile
    constexpr int kSlotTiles = kDrainChunk / kWarp;
    constexpr int kWarps = kDrainThreads / kWarp;
    extern __shared__ __align__(16) int4 s_cur[]; // the live words of the touched rows, kSlotTiles tiles (slot = column)
    __shared__ ChunkPods cp;
    __shared__ DevReq s_req[kDrainReqBuf];
    __shared__ int32_t s_roff[kDrainChunk]; // pod k's requests start at s_req[s_roff[k]]; -1 = read them from global
    __shared__ unsigned long long s_list[kDrainChunk][kDrainTop]; // kTouched = the row has been touched since
    __shared__ unsigned long long s_bound[kDrainChunk];           // no row outside pod k's list had a larger key
    __shared__ int s_newrow; // global row touched for the first time by the previous pod, or -1
    __shared__ int32_t s_count0[kDrainChunk], s_pick_lo[kDrainChunk], s_pick_hi[kDrainChunk];
    __shared__ int32_t s_trow[kDrainChunk];  // local row of slot s
    __shared__ int32_t s_nt_at[kDrainChunk]; // touched rows when pod k was placed; -1 = its fit count is final
    __shared__ unsigned long long s_wkey[kWarps];
    __shared__ int32_t s_wcnt[kWarps];
    __shared__ int s_nT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kSlotTiles * kTile4; i += kDrainThreads) s_cur[i] = make_int4(0, 0, 0, 0);
    if (tid < kDrainChunk) s_trow[tid] = -1;
    chunk_pods_load(cp, a.pods, a.reqs, a.n_pods);
    if (tid < a.n_pods) {
        const GenPod p = cp.pod[tid];
        // the segments' lists merged (what the merge drops goes into the bound), the segments' counts added up
        const unsigned long long *seg = a.topk + (size_t)tid * a.n_segs * kListWords;
        int head[kDrainMaxSegs];
        for (int g = 0; g < kDrainMaxSegs; g++) head[g] = 0;
        for (int r = 0; r < kDrainTop; r++) {
            unsigned long long best = 0ull;
            int from = 0;
            for (int g = 0; g < a.n_segs; g++) {
                const unsigned long long e = head[g] < kDrainTop ? seg[g * kListWords + head[g]] : 0ull;
                if (e > best) best = e, from = g;
            }
            head[from]++;
            s_list[tid][r] = best;
        }
        unsigned long long bound = 0ull;
        int32_t c0 = 0;
        for (int g = 0; g < a.n_segs; g++) {
            const unsigned long long rest = seg[g * kListWords + min(head[g], kDrainTop)]; // next key, or the bound
            const unsigned long long b = seg[g * kListWords + kDrainTop];
            bound = max(bound, max(rest, b));
            c0 += a.counts[tid * a.n_segs + g];
        }
        s_bound[tid] = bound;
        s_count0[tid] = c0;
        s_pick_lo[tid] = a.pick_begin[p.req_begin];
        s_pick_hi[tid] = a.pick_begin[p.req_begin + p.n_reqs];
    }
    if (tid == 0) { // where each multi-request pod's requests go
        s_newrow = -1;
        int off = 0;
        for (int k = 0; k < a.n_pods; k++) {
            const int n = cp.pod[k].n_reqs;
            const bool in = n > 1 && off + n <= kDrainReqBuf;
            s_roff[k] = in ? off : -1;
            off += in ? n : 0;
        }
    }
    __syncthreads();
    for (int k = warp; k < a.n_pods; k += kWarps) {
        const int off = s_roff[k];
        if (off >= 0)
            for (int r = lane; r < cp.pod[k].n_reqs; r += kWarp) s_req[off + r] = load_req(&a.reqs[cp.pod[k].req_begin + r]);
    }
    uint32_t one = 1u;
    asm volatile("" : "+r"(one));
    int nT = 0;
#ifdef VGS_DRAIN_TIMING
    long long tA = 0, tB = 0, tC = 0, tD = 0, tE = 0, t0 = 0, t1 = 0;
#define TICK(acc) t1 = clock64(), acc += t1 - t0, t0 = t1
#else
#define TICK(acc)
#endif
    __syncthreads();

    for (int q = 0; q < a.n_pods; q++) {
#ifdef VGS_DRAIN_TIMING
        t0 = clock64();
#endif
        const GenPod pod = cp.pod[q];
        const bool is_pair = (cp.pair[q >> 5] >> (q & 31)) & 1u;
        const int roff = s_roff[q];
        auto req_at = [&](int r) {
            return pod.n_reqs == 1 ? cp.first[q] : roff >= 0 ? s_req[roff + r] : load_req(&a.reqs[pod.req_begin + r]);
        };
        // the last warp pulls the next pod's likeliest winners (the head of its list) towards this SM: their raw planes
        if (warp == kWarps - 1 && q + 1 < a.n_pods) {
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const unsigned long long e = s_list[q + 1][c];
                if (e > kTouched) {
                    const int32_t row = (int32_t)(uint32_t)e - a.row_base;
                    if (lane < VGS_NUM_PLANES) warm(a.raw.p[lane] + (size_t)row * a.dmax);
                    else if (lane == VGS_NUM_PLANES) warm(a.raw.ndev + row);
                }
            }
        }
        // a row touched for the first time by the previous pod leaves every list (the entries stay, as kTouched)
        if (s_newrow >= 0) {
            const uint32_t nr = (uint32_t)s_newrow;
            for (int e = tid; e < a.n_pods * kDrainTop; e += kDrainThreads) {
                const unsigned long long x = (&s_list[0][0])[e];
                if (x > kTouched && (uint32_t)x == nr) (&s_list[0][0])[e] = kTouched;
            }
        }
        TICK(tA);

        // ---- the touched rows, live: warp w on slots [32 w, 32 w + 32)
        unsigned long long key = 0ull;
        int32_t dc = 0;
        if (warp < kSlotTiles && nT > warp * kWarp) { // warp-uniform
            const int slot = warp * kWarp + lane;
            const int4 *tp = s_cur + (size_t)warp * kTile4;
            bool alive;
            float score;
            if (is_pair) {
                PairNode<DM> nd;
                pair_node_load<DM, false>(tp, lane, nd);
                pair_eval<DM>(nd, cp.first[q], alive, score);
            } else {
                WalkNode<DM> nd;
                walk_node_load<DM, false>(tp, lane, nd);
                walk_eval<DM, false, false>(nd, pod, req_at, one, nullptr, false, alive, score);
            }
            if (slot < nT && alive) key = key_of(score, (uint32_t)(a.row_base + s_trow[slot])), dc = 1;
        }
        if (warp < kSlotTiles) {
            key = warp_max64(key);
            dc = __reduce_add_sync(0xffffffffu, dc);
            if (lane == 0) s_wkey[warp] = key, s_wcnt[warp] = dc;
        }
        __syncthreads();
        unsigned long long tkey = 0ull;
        int32_t live = 0;
#pragma unroll
        for (int w = 0; w < kSlotTiles; w++) {
            tkey = s_wkey[w] > tkey ? s_wkey[w] : tkey;
            live += s_wcnt[w];
        }
        // ---- the first candidate outside T (lane k looks at entry k)
        static_assert(kDrainTop <= 32, "one lane per list entry");
        const unsigned long long mine = lane < kDrainTop ? s_list[q][lane] : 0ull;
        const unsigned open_ = __ballot_sync(0xffffffffu, mine > kTouched);
        const unsigned long long lkey = open_ ? __shfl_sync(0xffffffffu, mine, __ffs((int)open_) - 1) : 0ull;
        unsigned long long best = tkey > lkey ? tkey : lkey;
        int32_t fit_count = live, nt_at = nT;
        if (best < s_bound[q]) {
            // some row the list does not name had a larger key at chunk start than anything in sight: scan the walk
            // copy (it holds the commits so far).  CTA-uniform branch.
            __syncthreads(); // s_w* are read above
            unsigned long long sk = 0ull;
            int32_t sc = 0;
            for (int tile = warp; tile < a.n_tiles; tile += kWarps) {
                const int4 *tp = a.walk + (size_t)tile * kTile4;
                bool alive;
                float score;
                if (is_pair) {
                    PairNode<DM> nd;
                    pair_node_load<DM, false>(tp, lane, nd);
                    pair_eval<DM>(nd, cp.first[q], alive, score);
                } else {
                    WalkNode<DM> nd;
                    walk_node_load<DM, false>(tp, lane, nd);
                    walk_eval<DM, false, false>(nd, pod, req_at, one, nullptr, false, alive, score);
                }
                if (alive) {
                    const unsigned long long v = key_of(score, (uint32_t)(a.row_base + tile * kWarp + lane));
                    sk = v > sk ? v : sk;
                    sc++;
                }
            }
            sk = warp_max64(sk);
            sc = __reduce_add_sync(0xffffffffu, sc);
            if (lane == 0) s_wkey[warp] = sk, s_wcnt[warp] = sc;
            __syncthreads();
            best = 0ull, fit_count = 0, nt_at = -1;
#pragma unroll
            for (int w = 0; w < kWarps; w++) {
                best = s_wkey[w] > best ? s_wkey[w] : best;
                fit_count += s_wcnt[w];
            }
            if (tid == 0) atomicAdd(&a.stats[0], 1);
        }
        TICK(tB);

        // ---- resolve, commit, prepare again: warp 0
        if (warp == 0) {
            const int32_t pick_lo = s_pick_lo[q], pick_hi = s_pick_hi[q];
            int newrow = -1;
            vgs_result res;
            res.fit_count = fit_count; // (nt_at >= 0: the live part; the chunk-start part is added when the chunk is done)
            if (best == 0ull) {
                vgs_pick z;
                z.slot = z.usedmem = z.usedcores = 0;
                for (int32_t i = pick_lo + lane; i < pick_hi; i += kWarp) a.picks[i] = z;
                res.best_row = -1, res.score = 0.0f, res.n_picks = 0;
            } else {
                const uint32_t kbits = unorder_bits((uint32_t)(best >> 32));
                const int32_t row = (int32_t)(uint32_t)best - a.row_base;
                bool alive;
                float score;
                int32_t ndv, used, usedmem, usedcores;
                RowStatic fx;
                resolve_row_warp(
                    a.raw, row, a.dmax, pod.n_reqs, req_at,
                    [&](int r, int t, const vgs_pick &pk) { a.picks[a.pick_begin[pod.req_begin + r] + t] = pk; }, alive,
                    score, ndv, used, usedmem, usedcores, &fx);
                if ((!alive || __float_as_uint(score) != kbits) && lane == 0) atomicOr(a.status, S_MISMATCH);
                TICK(tC);
                if (alive) {
                    if (lane < ndv) { // addPod: the placement becomes usage (scheduler.go:280-297)
                        const size_t at = (size_t)row * a.dmax + lane;
                        const_cast<int32_t *>(a.raw.p[VGS_PLANE_USED])[at] = used;
                        const_cast<int32_t *>(a.raw.p[VGS_PLANE_USEDMEM])[at] = usedmem;
                        const_cast<int32_t *>(a.raw.p[VGS_PLANE_USEDCORES])[at] = usedcores;
                    }
                    // the row's slot among the touched rows
                    unsigned hit = 0u;
#pragma unroll
                    for (int t = 0; t < kSlotTiles; t++) {
                        const unsigned b = __ballot_sync(0xffffffffu, t * kWarp + lane < nT && s_trow[t * kWarp + lane] == row);
                        if (b) hit = (unsigned)(t * kWarp + __ffs((int)b) - 1) + 1u;
                    }
                    int slot;
                    if (hit) {
                        slot = (int)hit - 1;
                    } else {
                        slot = nT++;
                        if (lane == 0) s_trow[slot] = row, newrow = (int)(uint32_t)best;
                    }
                    // its words again, from the columns the resolve holds in registers (lane = device slot)
                    prepare_lanes<DM>(row, ndv, 0u, used, fx.count, usedmem, fx.totalmem, usedcores, fx.totalcore, fx.numa,
                                      fx.type, a.raw.idle_types, a.walk, a.status, s_cur + (size_t)(slot / kWarp) * kTile4,
                                      slot % kWarp);
                    TICK(tD);
                }
                res.best_row = (int32_t)(uint32_t)best;
                res.score = __uint_as_float(kbits);
                res.n_picks = pick_hi - pick_lo;
            }
            if (lane == 0) {
                a.results[q] = res;
                s_nt_at[q] = nt_at;
                s_nT = nT;
                s_newrow = newrow;
            }
        }
        __syncthreads();
        nT = s_nT;
        TICK(tE);
    }
    // ---- fit counts: + chunk-start count - the touched rows (as of the pod's turn) that fitted at chunk start
    for (int k = warp; k < a.n_pods; k += kWarps) {
        const int n = s_nt_at[k];
        if (n < 0) continue; // scanned: final
        const uint32_t *mine = a.ob + (size_t)k * a.n_tiles * kWarp;
        int32_t was = 0;
        for (int s = lane; s < n; s += kWarp) was += mine[s_trow[s]] != 0u ? 1 : 0;
        was = __reduce_add_sync(0xffffffffu, was);
        if (lane == 0) a.results[k].fit_count += s_count0[k] - was;
    }
#ifdef VGS_DRAIN_TIMING
    if (tid == 0) {
        atomicAdd(&a.stats[1], (int)(tA >> 6)), atomicAdd(&a.stats[2], (int)(tB >> 6)), atomicAdd(&a.stats[3], (int)(tC >> 6));
        atomicAdd(&a.stats[4], (int)(tD >> 6)), atomicAdd(&a.stats[5], (int)(tE >> 6));
    }
#endif
}

template <int DM>
static cudaError_t drain_chunk_t(int dev, DrainChunk &c, uint32_t *ob, unsigned long long *topk, int32_t *counts,
                                 int sms, cudaStream_t st) {
    constexpr size_t smem = (size_t)(kDrainChunk / kWarp) * kWalkWords * (DM / 4) * kWarp * sizeof(int4);
    static bool attr[64] = {}; // (static + dynamic shared memory of the commit kernel passes 48 KB for DM = 12)
    if (dev >= 0 && dev < 64 && !attr[dev]) {
        cudaError_t ea = cudaFuncSetAttribute(k_drain_commit<DM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ea != cudaSuccess) return ea;
        attr[dev] = true;
    }
    // enough (tile, pod part) items for four rounds of the resident warps (two CTAs of 8 warps per SM)
    const long long ctas = (long long)sms * 2, warps = ctas * (kScoreThreads / 32);
    int psplit = (int)((4 * warps + c.n_tiles - 1) / (c.n_tiles > 0 ? c.n_tiles : 1));
    if (psplit < 1) psplit = 1;
    if (psplit > c.n_pods) psplit = c.n_pods;
    const long long items = (long long)c.n_tiles * psplit;
    long long grid = (items + kScoreThreads / 32 - 1) / (kScoreThreads / 32);
    if (grid > ctas) grid = ctas;
    if (grid < 1) grid = 1;
    k_drain_eval<DM><<<(int)grid, kScoreThreads, 0, st>>>(c.walk, c.n_tiles, c.reqs, c.pods, c.n_pods, psplit, ob);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // segments per pod: the select CTAs of a chunk fill the GPU about once, a segment is at least 64 tiles
    int segs = (int)((sms + c.n_pods - 1) / c.n_pods);
    if (segs > kDrainMaxSegs) segs = kDrainMaxSegs;
    if (segs > (c.n_tiles + 63) / 64) segs = (c.n_tiles + 63) / 64;
    if (segs < 1) segs = 1;
    k_drain_select<<<dim3(segs, c.n_pods), kSelectThreads, 0, st>>>(ob, c.n_tiles, c.row_base, segs, topk, counts);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    c.n_segs = segs;
    k_drain_commit<DM><<<1, kDrainThreads, smem, st>>>(c);
    return cudaGetLastError();
}

cudaError_t launch_drain_chunk(int dm, int dev, int sms, const RawPlanes &raw, int4 *walk, int32_t n_tiles,
                               int32_t n_nodes, int32_t row_base, int32_t dmax, const DevReq *reqs, const GenPod *pods,
                               const int32_t *pick_begin, int32_t n_pods, uint32_t *ob, unsigned long long *topk,
                               int32_t *counts, vgs_result *results, vgs_pick *picks, uint32_t *status, int32_t *stats,
                               cudaStream_t st) {
    if (n_pods <= 0 || n_pods > kDrainChunk || n_tiles <= 0) return cudaErrorInvalidValue;
    DrainChunk c;
    c.walk = walk, c.n_tiles = n_tiles, c.n_nodes = n_nodes, c.row_base = row_base, c.dmax = dmax;
    c.raw = raw;
    c.reqs = reqs, c.pods = pods, c.pick_begin = pick_begin, c.n_pods = n_pods;
    c.ob = ob, c.topk = topk, c.counts = counts, c.n_segs = 1;
    c.results = results, c.picks = picks, c.status = status, c.stats = stats;
    return dm == 8 ? drain_chunk_t<8>(dev, c, ob, topk, counts, sms, st) : drain_chunk_t<12>(dev, c, ob, topk, counts, sms, st);
}

} // namespace vgs
