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
//   k_drain_commit  ONE CTA walks the chunk in pod order.  T = the rows the previous chunk changed (see below) +
//                   the rows this chunk has changed so far (<= 2 C), their tile words in shared memory.  For pod q
//                   every row outside T still has the key the matrix holds, so
//                       winner(q) = max( best live key over T,  first candidate of q's list outside T )
//                   provided that maximum reaches the list's bound; the live keys over T are evaluated now (pair
//                   pods: lane = (row, device); others: lane = row).  Otherwise the list says nothing about the rest
//                   of the snapshot and the CTA scans the walk copy for the rows outside T (rare:
//                   vgs_plan.drain_rescans).
//                       fit_count(q) = matrix count - #{T rows fitting in the matrix} + #{T rows fitting now},
//                   the middle term read from the matrix when the chunk is done.
//                   The winner is committed by warp 0: a pair pod straight from the row's tile words (picks, usage
//                   into the raw planes, the row's words after the take, re-ranked as K0 would); other pods by the
//                   warp resolve from the raw planes + the warp preparation.  The walk copy is NOT written: the rows
//                   the chunk changed leave in a carry block,
//   k_drain_apply   which this kernel stores into the walk copy.
// Two streams: the pass (eval + select) of chunk k+1 runs while chunk k is placed, over a walk copy that holds the
// chunks up to k-1 -- which is why the rows chunk k changed are carried into T of chunk k+1 (vgs_capi.cu queues it).
// Every result equals what the per-pod chain (one score kernel over all nodes per pod) produces.
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
// WALK = false: the chunk's pair pods (fewer registers, three CTAs per SM); true: the others.  A chunk with both kinds
// launches both; each writes its pods' rows of the matrix.
template <int DM, bool WALK>
__global__ void __launch_bounds__(kScoreThreads, WALK ? 2 : 3)
    k_drain_eval(const int4 *__restrict__ walk, int32_t n_tiles, const DevReq *__restrict__ reqs,
                 const GenPod *__restrict__ pods, int32_t n_pods, int32_t psplit, uint32_t *__restrict__ ob,
                 const int32_t *__restrict__ stop) {
    if (*reinterpret_cast<const volatile int32_t *>(stop) != 0) return; // the commit kernels have given the chunks up (see there)
    __shared__ ChunkPods cp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kScoreThreads / 32;
    chunk_pods_load(cp, pods, reqs, n_pods);
    const size_t stride = (size_t)n_tiles * kWarp;
    const int per = (n_pods + psplit - 1) / psplit;
    const long long n_items = (long long)n_tiles * psplit;
    for (long long item = (long long)blockIdx.x * kWarps + warp; item < n_items; item += (long long)gridDim.x * kWarps) {
        const int tile = (int)(item / psplit), part = (int)(item % psplit);
        const int k0 = part * per, k1 = min(n_pods, k0 + per);
        const int4 *tp = walk + (size_t)tile * kWalkWords * (DM / 4) * kWarp;
        uint32_t *out = ob + (size_t)tile * kWarp + lane;
        if constexpr (!WALK) {
            PairNode<DM> nd;
            pair_node_load<DM, true>(tp, lane, nd);
            for (int k = k0; k < k1; k++) {
                if (!((cp.pair[k >> 5] >> (k & 31)) & 1u)) continue;
                bool fit;
                float score;
                pair_eval<DM>(nd, cp.first[k], fit, score);
                out[k * stride] = fit ? order_bits(__float_as_uint(score)) : 0u;
            }
        } else {
            uint32_t one = 1u;
            asm volatile("" : "+r"(one));
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
constexpr int kListWords = kDrainListWords; // keys + bound
__global__ void __launch_bounds__(kSelectThreads)
    k_drain_select(const uint32_t *__restrict__ ob, int32_t n_tiles, int32_t row_base, int32_t n_segs,
                   unsigned long long *__restrict__ seglist, int32_t *__restrict__ segcount, unsigned int *__restrict__ done,
                   unsigned long long *__restrict__ topk, int32_t *__restrict__ counts, const int32_t *__restrict__ stop) {
    if (*reinterpret_cast<const volatile int32_t *>(stop) != 0) return;
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
    constexpr int kFly = 8; // loads in flight per thread
    for (int rb = r0 + (int)threadIdx.x; rb < r1; rb += kFly * kSelectThreads) {
        uint32_t bits[kFly];
#pragma unroll
        for (int u = 0; u < kFly; u++) {
            const int r = rb + u * kSelectThreads;
            bits[u] = r < r1 ? __ldg(&src[r]) : 0u;
        }
#pragma unroll
        for (int u = 0; u < kFly; u++) {
            if (bits[u] == 0u) continue;
            cnt++;
            unsigned long long v = ((unsigned long long)bits[u] << 32) | (uint32_t)(row_base + rb + u * kSelectThreads);
            if (v > t[kSelectMine - 1]) {
#pragma unroll
                for (int i = 0; i < kSelectMine; i++) {
                    const unsigned long long hi = v > t[i] ? v : t[i];
                    v = v > t[i] ? t[i] : v;
                    t[i] = hi;
                }
            }
            dropped = v > dropped ? v : dropped;
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) s_cnt[warp] = cnt;
    unsigned long long *list = seglist + ((size_t)pod * n_segs + seg) * kListWords;
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
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        int32_t c = 0;
        for (int w = 0; w < kWarps; w++) c += s_cnt[w];
        segcount[pod * n_segs + seg] = c;
        __threadfence();
        s_last = atomicAdd(&done[pod], 1u) == (unsigned)n_segs - 1u;
    }
    __syncthreads();
    if (!s_last) return;
    // the pod's last segment merges: a key's place in the pod's list = the number of larger keys among the segments'
    __threadfence();
    __shared__ unsigned long long s_keys[kDrainMaxSegs * kDrainTop];
    const int n_keys = n_segs * kDrainTop;
    const unsigned long long *mine_list = seglist + (size_t)pod * n_segs * kListWords;
    for (int i = threadIdx.x; i < n_keys; i += kSelectThreads)
        s_keys[i] = __ldcg(&mine_list[(i / kDrainTop) * kListWords + i % kDrainTop]);
    __syncthreads();
    unsigned long long out_of_list = 0ull; // goes into the bound
    for (int i = threadIdx.x; i < n_keys; i += kSelectThreads) {
        const unsigned long long v = s_keys[i];
        int rank = 0;
        for (int j = 0; j < n_keys; j++) rank += s_keys[j] > v ? 1 : 0;
        if (v != 0ull && rank < kDrainTop) topk[(size_t)pod * kListWords + rank] = v;
        else out_of_list = v;
    }
    if (threadIdx.x < n_segs) {
        const unsigned long long b = __ldcg(&mine_list[threadIdx.x * kListWords + kDrainTop]);
        out_of_list = b > out_of_list ? b : out_of_list;
    }
    // slots no key claimed: zero (fewer than kDrainTop fitting rows)
    int n_nonzero = 0;
    for (int j = 0; j < n_keys; j++) n_nonzero += s_keys[j] != 0ull ? 1 : 0;
    for (int r = n_nonzero + (int)threadIdx.x; r < kDrainTop; r += kSelectThreads) topk[(size_t)pod * kListWords + r] = 0ull;
    const unsigned long long m = warp_max64(out_of_list);
    __syncthreads();
    if (lane == 0) s_best[warp] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long bound = 0ull;
        for (int w = 0; w < kWarps; w++) bound = s_best[w] > bound ? s_best[w] : bound;
        topk[(size_t)pod * kListWords + kDrainTop] = bound;
        int32_t c = 0;
        for (int g = 0; g < n_segs; g++) c += __ldcg(&segcount[pod * n_segs + g]);
        counts[pod] = c;
        done[pod] = 0u; // zero at rest
    }
}

// Address of word k of the device at walk position w of column col of a tile (int32 view of the int4 layout)
template <int DM>
__device__ __forceinline__ int tile_at(int k, int w, int col) {
    return ((k * (DM / 4) + (w >> 2)) * kWarp + col) * 4 + (w & 3);
}

// pull a line towards this SM ahead of its use
__device__ __forceinline__ void warm(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// What one chunk's commit kernel hands to the next and to k_drain_apply: the rows it changed, with their words as the
// chunk leaves them (tile layout, column = index in rows[]).
constexpr int kDrainSlots = 2 * kDrainChunk; // rows a commit kernel can hold: carried in + touched by its own pods
struct DrainCarry {
    int32_t n;
    int32_t rows[kDrainChunk]; // local rows
};

struct DrainChunk {
    const int4 *walk; // the walk copy: commits up to the previous chunk (k_drain_apply), not this one's
    int32_t n_tiles, n_nodes, row_base, dmax;
    RawPlanes raw;
    const DevReq *reqs;
    const GenPod *pods;        // the chunk's pods, batch order
    const int32_t *pick_begin; // batch-wide, per request
    int32_t n_pods;
    const uint32_t *ob;             // [pod][n_tiles * 32] score bits, as of the start of the PREVIOUS chunk
    const unsigned long long *topk; // [pod][kListWords]: candidate keys descending, then the bound
    const int32_t *counts;          // [pod]
    int32_t first; // 1 = first chunk of the drain: nothing carried in, the matrix is as of this chunk's start
    const DrainCarry *carry_in; // rows the previous chunk changed ...
    const int4 *words_in;       // ... and their words, kDrainChunk / 32 tiles
    DrainCarry *carry_out;      // the same for this chunk
    int4 *words_out;
    vgs_result *results; // the chunk's
    vgs_pick *picks;     // the batch's
    uint32_t *status;
    int32_t *stats; // [0] += pods that needed the full scan
    // Re-scans are what this scheme cannot afford on a big snapshot (one CTA over the whole walk copy: ~0.2 ms at 100k
    // nodes).  Near-identical pods on nodes that take one or two of them exhaust each other's lists and need one per
    // pod.  A chunk that would pass max_rescans stops there: *stop = the first pod not placed (0 = none so far; pod 0
    // never re-scans), every later kernel of the drain returns at once, and the host places the rest pod by pod.
    int32_t first_pod, max_rescans;
    int32_t *stop;
};

constexpr int kDrainThreads = 512; // (with its registers, nothing of the concurrent pass fits beside it on the SM)
constexpr int kDrainReqBuf = 384; // requests of the chunk's multi-request pods kept in shared memory

// ------------------------------------------------------------------------------------
// the chunk in pod order, one CTA.  T = the rows the previous chunk changed (the matrix and the lists do not know
// that: they were worked out while that chunk ran) + the rows this chunk has changed so far.
// ------------------------------------------------------------------------------------
template <int DM>
__global__ void __launch_bounds__(kDrainThreads) k_drain_commit(const __grid_constant__ DrainChunk a) {
    constexpr int kTile4 = kWalkWords * (DM / 4) * kWarp; // int4 per tile
    constexpr int kRow4 = kWalkWords * (DM / 4);          // int4 per row
    constexpr int kSlotTiles = kDrainSlots / kWarp;
    constexpr int kWarps = kDrainThreads / kWarp;
    static_assert(kSlotTiles < kWarps, "helper warps needed");
    extern __shared__ __align__(16) int4 s_cur[]; // the live words of T's rows, kSlotTiles tiles (slot = column)
    __shared__ ChunkPods cp;
    __shared__ DevReq s_req[kDrainReqBuf];
    __shared__ int32_t s_roff[kDrainChunk]; // pod k's requests start at s_req[s_roff[k]]; -1 = read them from global
    __shared__ unsigned long long s_list[kDrainChunk][kDrainTop]; // kTouched = the row is in T
    __shared__ unsigned long long s_bound[kDrainChunk];           // no row outside pod k's list had a larger key
    __shared__ int s_newrow; // global row that entered T with the previous pod, or -1
    __shared__ int32_t s_count0[kDrainChunk], s_pick_lo[kDrainChunk], s_pick_hi[kDrainChunk];
    __shared__ int32_t s_trow[kDrainSlots];  // local row of slot s
    __shared__ uint32_t s_mod[kSlotTiles];   // bit: the slot's row was changed by this chunk
    __shared__ int32_t s_nt_at[kDrainChunk]; // |T| when pod k was placed; -1 = its fit count is final
    __shared__ unsigned long long s_wkey[kWarps];
    __shared__ int32_t s_wcnt[kWarps];
    constexpr int kHashBits = 8, kHash = 1 << kHashBits; // >= 4 x the carried rows
    __shared__ int32_t s_hash[kHash];
    __shared__ uint32_t s_filter[32]; // rescan: bit (tile & 1023) for every tile holding a row of T
    __shared__ int s_nT;
    // the tile words of the next pod's two likeliest winners outside T, fetched one pod ahead (row-major: kRow4 int4)
    __shared__ __align__(16) int4 s_cand[2][2][kRow4];
    __shared__ int32_t s_cand_row[2][2]; // local rows, -1 = none
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int32_t st0 = *reinterpret_cast<volatile int32_t *>(a.stop);
        if (st0 != 0 && st0 <= a.first_pod) return; // given up in an earlier chunk
    }
    int n_done = a.n_pods, n_rescans = 0;
    // (the last warp) start the copies for pod k into buffer k & 1; `skip` = global row that is about to enter T
    auto fetch_candidates = [&](int k, int skip) {
        const unsigned long long e = lane < kDrainTop ? s_list[k][lane] : 0ull;
        unsigned open_ = __ballot_sync(0xffffffffu, e > kTouched && (int)(uint32_t)e != skip);
#pragma unroll
        for (int c = 0; c < 2; c++) {
            int32_t row = -1;
            if (open_) {
                const int at = __ffs((int)open_) - 1;
                open_ &= open_ - 1u;
                row = (int32_t)(uint32_t)__shfl_sync(0xffffffffu, e, at) - a.row_base;
                if (lane < kRow4) {
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(&s_cand[k & 1][c][lane]);
                    const int4 *src = a.walk + (size_t)(row / kWarp) * kTile4 + lane * kWarp + row % kWarp;
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
                }
            }
            if (lane == 0) s_cand_row[k & 1][c] = row;
        }
    };
    for (int i = tid; i < kSlotTiles * kTile4; i += kDrainThreads) s_cur[i] = make_int4(0, 0, 0, 0);
    if (tid < kDrainSlots) s_trow[tid] = -1;
    if (tid < kSlotTiles) s_mod[tid] = 0u;
    chunk_pods_load(cp, a.pods, a.reqs, a.n_pods);
    const int n_in = a.first ? 0 : a.carry_in->n;
    // carried in: rows and live words
    if (tid < n_in) s_trow[tid] = a.carry_in->rows[tid];
    for (int i = tid; i < ((n_in + kWarp - 1) / kWarp) * kTile4; i += kDrainThreads) s_cur[i] = a.words_in[i];
    for (int i = tid; i < a.n_pods * kListWords; i += kDrainThreads) {
        const unsigned long long x = a.topk[i];
        const int k = i / kListWords, r = i % kListWords;
        if (r < kDrainTop) s_list[k][r] = x;
        else s_bound[k] = x;
    }
    if (tid < a.n_pods) {
        const GenPod p = cp.pod[tid];
        s_count0[tid] = a.counts[tid];
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
    // the carried rows leave the lists (a small open-addressing set of them)
    for (int i = tid; i < kHash; i += kDrainThreads) s_hash[i] = -1;
    __syncthreads();
    if (tid < n_in) {
        const int32_t r = s_trow[tid];
        for (uint32_t h = ((uint32_t)r * 2654435761u) >> (32 - kHashBits);; h = (h + 1) & (kHash - 1))
            if (atomicCAS(&s_hash[h], -1, r) == -1) break;
    }
    __syncthreads();
    for (int e = tid; e < a.n_pods * kDrainTop; e += kDrainThreads) {
        const unsigned long long x = (&s_list[0][0])[e];
        if (x == 0ull) continue;
        const int32_t r = (int32_t)(uint32_t)x - a.row_base;
        for (uint32_t h = ((uint32_t)r * 2654435761u) >> (32 - kHashBits); s_hash[h] != -1; h = (h + 1) & (kHash - 1))
            if (s_hash[h] == r) {
                (&s_list[0][0])[e] = kTouched;
                break;
            }
    }
    __syncthreads(); // the lists are final
    if (warp == kWarps - 1) fetch_candidates(0, -1);
    uint32_t one = 1u;
    asm volatile("" : "+r"(one));
    int nT = n_in;
#ifdef VGS_DRAIN_TIMING
    long long tA = 0, tB = 0, tB1 = 0, tB2 = 0, tC = 0, tD = 0, tE = 0, t0 = 0, t1 = 0;
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
        auto eval_on = [&](const int4 *tp, bool &alive, float &score) { // this lane's column of the tile
            if (is_pair) {
                PairNode<DM> nd;
                pair_node_load<DM, false>(tp, lane, nd);
                pair_eval<DM>(nd, cp.first[q], alive, score);
            } else {
                WalkNode<DM> nd;
                walk_node_load<DM, false>(tp, lane, nd);
                walk_eval<DM, false, false>(nd, pod, req_at, one, nullptr, false, alive, score);
            }
        };
        if (warp >= kSlotTiles) { // the helper warps
            // this pod's candidates have arrived (started one pod ago); the next pod's are started
            if (warp == kWarps - 1) {
                asm volatile("cp.async.wait_all;" ::: "memory");
                if (q + 1 < a.n_pods) fetch_candidates(q + 1, s_newrow);
            }
            // a row that entered T with the previous pod leaves every list (the entries stay, as kTouched)
            if (s_newrow >= 0) {
                const uint32_t nr = (uint32_t)s_newrow;
                for (int e = tid - kSlotTiles * kWarp; e < a.n_pods * kDrainTop; e += kDrainThreads - kSlotTiles * kWarp) {
                    const unsigned long long x = (&s_list[0][0])[e];
                    if (x > kTouched && (uint32_t)x == nr) (&s_list[0][0])[e] = kTouched;
                }
            }
        }
        TICK(tA);

        // ---- the first candidate outside T (lane k looks at entry k; the row that entered T with the previous pod
        // is being struck out by the helper warps right now: it is excluded by name)
        static_assert(kDrainTop <= 32, "one lane per list entry");
        const unsigned long long mine = lane < kDrainTop ? s_list[q][lane] : 0ull;
        const unsigned open_ = __ballot_sync(0xffffffffu, mine > kTouched && (int)(uint32_t)mine != s_newrow);
        const unsigned long long lkey = open_ ? __shfl_sync(0xffffffffu, mine, __ffs((int)open_) - 1) : 0ull;
        const unsigned long long bound = s_bound[q];
        // ---- T's rows, live
        if (is_pair) {
            // pair pods: lane = (slot, device) -- a group of DM-rounded-to-8/16 lanes per row, the take decided by
            // ballots and the sums by shuffles inside the group; every warp takes part
            constexpr int kGrp = DM <= 8 ? 8 : 16, kPerWarp = kWarp / kGrp;
            const DevReq rq = cp.first[q];
            const int j = lane % kGrp, g = lane / kGrp;
            unsigned long long key = 0ull;
            int32_t dc = 0;
            constexpr int kRounds = kDrainSlots / (kWarps * kPerWarp);
            // (the rounds are independent: unrolled, their instruction streams interleave)
#pragma unroll
            for (int rd = 0; rd < kRounds; rd++) {
                const int s0 = rd * kWarps * kPerWarp + warp * kPerWarp;
                if (rd * kWarps * kPerWarp >= nT) break; // CTA-uniform
                const int slot = s0 + g;
                const bool on = slot < nT && j < DM;
                const int32_t *src = reinterpret_cast<const int32_t *>(s_cur + (size_t)((on ? slot : 0) / kWarp) * kTile4);
                const int col = (on ? slot : 0) % kWarp, jj = on ? j : 0;
                const int32_t tm = src[tile_at<DM>(W_TOTALMEM, jj, col)], fm = src[tile_at<DM>(W_FREEMEM, jj, col)];
                const int32_t fc = src[tile_at<DM>(W_FREECORES, jj, col)], cn = src[tile_at<DM>(W_COUNT, jj, col)];
                const int32_t fs = src[tile_at<DM>(W_FREESLOTS, jj, col)];
                const uint32_t mt = (uint32_t)src[tile_at<DM>(W_META, jj, col)];
                const int32_t mem = (rq.flags & R_USEPCT) ? pct_mem(tm, rq.pct) : rq.memfix;
                const bool f = (int)on & (int)!(mt & rq.blockmask) & (int)type_ok(rq.mask_lo, rq.mask_hi, mt) & (int)(fm >= mem) &
                               (int)(fc >= rq.coresreq);
                const unsigned grp = (__ballot_sync(0xffffffffu, f) >> (g * kGrp)) & ((1u << kGrp) - 1u);
                const bool taken = f && __popc(grp & ((1u << j) - 1u)) < rq.nums;
                uint32_t total = taken ? (uint32_t)cn : 0u, free_ = taken ? (uint32_t)fs : 0u;
#pragma unroll
                for (int d = 1; d < kGrp; d <<= 1) {
                    total += __shfl_xor_sync(0xffffffffu, total, d);
                    free_ += __shfl_xor_sync(0xffffffffu, free_, d);
                }
                const bool fit = slot < nT && __popc(grp) >= rq.nums;
                const int32_t ndv = (int32_t)((__shfl_sync(0xffffffffu, mt, g * kGrp) >> M_NDEV_SHIFT) & 15u);
                // (rows that are out divide 1 by 1: a zero operand would take the whole warp through the division's slow path)
                const float score = container_score(fit ? (int32_t)total : 1, fit ? (int32_t)free_ : 1, ndv - rq.nums);
                if (fit) {
                    const unsigned long long v = key_of(score, (uint32_t)(a.row_base + s_trow[slot]));
                    key = v > key ? v : key;
                    dc += j == 0 ? 1 : 0;
                }
            }
            key = warp_max64(key);
            dc = __reduce_add_sync(0xffffffffu, dc);
            if (lane == 0) s_wkey[warp] = key, s_wcnt[warp] = dc;
        } else {
            // the others: warp w on slots [32 w, 32 w + 32), lane = row
            unsigned long long key = 0ull;
            int32_t dc = 0;
            if (warp < kSlotTiles && nT > warp * kWarp) { // warp-uniform
                const int slot = warp * kWarp + lane;
                bool alive;
                float score;
                eval_on(s_cur + (size_t)warp * kTile4, alive, score);
                if (slot < nT && alive) key = key_of(score, (uint32_t)(a.row_base + s_trow[slot])), dc = 1;
                key = warp_max64(key);
                dc = __reduce_add_sync(0xffffffffu, dc);
            }
            if (lane == 0) s_wkey[warp] = key, s_wcnt[warp] = dc;
        }
        TICK(tB1);
        __syncthreads();
        TICK(tB2);
        unsigned long long tkey = 0ull;
        int32_t live = 0;
        { // lane w reads warp w's part
            static_assert(kWarps <= kWarp, "one lane per warp");
            tkey = warp_max64(lane < kWarps ? s_wkey[lane] : 0ull);
            live = __reduce_add_sync(0xffffffffu, lane < kWarps ? s_wcnt[lane] : 0);
        }
        unsigned long long best = tkey > lkey ? tkey : lkey;
        int32_t fit_count = live, nt_at = nT;
        if (best < bound && n_rescans >= a.max_rescans) { // CTA-uniform: stop here, the host goes on pod by pod
            n_done = q;
            break;
        }
        if (best < bound) {
            n_rescans++;
            // some row the list does not name had a larger key than anything in sight: scan the walk copy for the
            // rows outside T (T's rows are stale there; their live keys and count are in hand).  CTA-uniform branch.
            __syncthreads(); // s_w* are read above
            if (tid < 32) s_filter[tid] = 0u;
            __syncthreads();
            for (int s = tid; s < nT; s += kDrainThreads) atomicOr(&s_filter[((s_trow[s] / kWarp) & 1023) >> 5], 1u << ((s_trow[s] / kWarp) & 31));
            __syncthreads();
            // The scan does not only name the winner: it leaves this pod a fresh candidate list (the kDrainTop best rows
            // outside T and the bound behind them, selected as k_drain_select does), and the pods after it in the chunk
            // that ask for exactly the same inherit it -- replicas use up each other's candidates, one scan serves the
            // next kDrainTop of them.  (Outside T a row's key is the same for every pod with that request.)
            unsigned long long t4[kSelectMine], dropped = 0ull;
#pragma unroll
            for (int i = 0; i < kSelectMine; i++) t4[i] = 0ull;
            int32_t sc = 0;
            for (int tile = warp; tile < a.n_tiles; tile += kWarps) {
                bool alive;
                float score;
                eval_on(a.walk + (size_t)tile * kTile4, alive, score);
                if ((s_filter[(tile & 1023) >> 5] >> (tile & 31)) & 1u) { // may hold rows of T: they do not count here
                    const int32_t r = tile * kWarp + lane;
                    for (int s = 0; s < nT; s++) alive = alive && s_trow[s] != r;
                }
                if (alive) {
                    unsigned long long v = key_of(score, (uint32_t)(a.row_base + tile * kWarp + lane));
                    sc++;
#pragma unroll
                    for (int i = 0; i < kSelectMine; i++) {
                        const unsigned long long hi = v > t4[i] ? v : t4[i];
                        v = v > t4[i] ? t4[i] : v;
                        t4[i] = hi;
                    }
                    dropped = v > dropped ? v : dropped;
                }
            }
            sc = __reduce_add_sync(0xffffffffu, sc);
            if (lane == 0) s_wcnt[warp] = sc;
            unsigned long long scan_best = 0ull;
            for (int r = 0; r <= kDrainTop; r++) { // kDrainTop rounds of "largest head of all threads", one for the bound
                const unsigned long long mine_r = r < kDrainTop ? t4[0] : (t4[0] > dropped ? t4[0] : dropped);
                const unsigned long long m = warp_max64(mine_r);
                __syncthreads();
                if (lane == 0) s_wkey[warp] = m;
                __syncthreads();
                unsigned long long top = 0ull;
#pragma unroll
                for (int w = 0; w < kWarps; w++) top = s_wkey[w] > top ? s_wkey[w] : top;
                if (r == 0) scan_best = top;
                if (tid == 0) {
                    if (r < kDrainTop) s_list[q][r] = top;
                    else s_bound[q] = top;
                }
                if (r < kDrainTop && t4[0] == top && top != 0ull) {
#pragma unroll
                    for (int i = 0; i + 1 < kSelectMine; i++) t4[i] = t4[i + 1];
                    t4[kSelectMine - 1] = 0ull;
                }
            }
            best = tkey > scan_best ? tkey : scan_best, nt_at = -1;
#pragma unroll
            for (int w = 0; w < kWarps; w++) fit_count += s_wcnt[w];
            __syncthreads(); // the new list and bound of pod q
            for (int o = q + 1 + warp; o < a.n_pods; o += kWarps) { // warp per later pod: the same request?
                const bool both = is_pair && ((cp.pair[o >> 5] >> (o & 31)) & 1u);
                const uint32_t *x = reinterpret_cast<const uint32_t *>(&cp.first[q]), *y = reinterpret_cast<const uint32_t *>(&cp.first[o]);
                const bool same = both && __all_sync(0xffffffffu, lane >= (int)(sizeof(DevReq) / 4) || x[lane] == y[lane]);
                if (same) {
                    if (lane < kDrainTop) s_list[o][lane] = s_list[q][lane];
                    if (lane == 0) s_bound[o] = s_bound[q];
                }
            }
            if (tid == 0) atomicAdd(&a.stats[0], 1);
        }
        TICK(tB);

        // ---- resolve, commit, prepare again: warp 0
        if (warp == 0) {
            const int32_t pick_lo = s_pick_lo[q], pick_hi = s_pick_hi[q];
            int newrow = -1;
            vgs_result res;
            res.fit_count = fit_count; // (nt_at >= 0: the live part; the rest is added when the chunk is done)
            if (best == 0ull) {
                vgs_pick z;
                z.slot = z.usedmem = z.usedcores = 0;
                for (int32_t i = pick_lo + lane; i < pick_hi; i += kWarp) a.picks[i] = z;
                res.best_row = -1, res.score = 0.0f, res.n_picks = 0;
            } else {
                const uint32_t kbits = unorder_bits((uint32_t)(best >> 32));
                const int32_t row = (int32_t)(uint32_t)best - a.row_base;
                // the row's slot in T, a new one if it enters now
                unsigned hit = 0u;
                if (best != lkey) { // (a row out of the list is not in T)
#pragma unroll
                    for (int t = 0; t < kSlotTiles; t++) {
                        const unsigned b = __ballot_sync(0xffffffffu, t * kWarp + lane < nT && s_trow[t * kWarp + lane] == row);
                        if (b) hit = (unsigned)(t * kWarp + __ffs((int)b) - 1) + 1u;
                    }
                }
                auto take_slot = [&]() {
                    int slot = (int)hit - 1;
                    if (!hit) {
                        slot = nT++;
                        if (lane == 0) s_trow[slot] = row, newrow = (int)(uint32_t)best;
                    }
                    if (lane == 0) s_mod[slot / kWarp] |= 1u << (slot % kWarp);
                    return slot;
                };
                if (is_pair) {
                    // One request, the first Nums fitting devices in walk order: everything comes out of the row's
                    // tile words (lane = walk position) -- the picks, the usage for the raw planes (Used = Count -
                    // free slots, ...), and the row's words after the commit, devices re-ranked by (numa rank, free
                    // slots, slot) as K0 would.  W_SCORE1 is put right when the chunk hands its rows on.
                    const DevReq rq = cp.first[q];
                    // its words: from T, from the candidates fetched ahead (row-major), or from the walk copy
                    const int32_t *src;
                    int scol, cstride = kWarp;
                    if (hit) {
                        src = reinterpret_cast<const int32_t *>(s_cur + (size_t)((hit - 1) / kWarp) * kTile4), scol = (int)((hit - 1) % kWarp);
                    } else if (row == s_cand_row[q & 1][0] || row == s_cand_row[q & 1][1]) {
                        src = reinterpret_cast<const int32_t *>(&s_cand[q & 1][row == s_cand_row[q & 1][0] ? 0 : 1][0]);
                        scol = 0, cstride = 1;
                    } else {
                        src = reinterpret_cast<const int32_t *>(a.walk + (size_t)(row / kWarp) * kTile4), scol = row % kWarp;
                    }
                    const int j = lane < DM ? lane : 0;
                    auto word = [&](int k) { return src[((k * (DM / 4) + (j >> 2)) * cstride + scol) * 4 + (j & 3)]; };
                    int32_t tm = word(W_TOTALMEM), fm = word(W_FREEMEM), fc = word(W_FREECORES), cn = word(W_COUNT);
                    int32_t fs = word(W_FREESLOTS), tc = word(W_TOTALCORE);
                    uint32_t mt = (uint32_t)word(W_META);
                    const int32_t ndv = (int32_t)((mt >> M_NDEV_SHIFT) & 15u);
                    const int32_t mem = (rq.flags & R_USEPCT) ? pct_mem(tm, rq.pct) : rq.memfix;
                    const bool f = (int)(lane < DM) & (int)!(mt & rq.blockmask) & (int)type_ok(rq.mask_lo, rq.mask_hi, mt) &
                                   (int)(fm >= mem) & (int)(fc >= rq.coresreq);
                    const unsigned fits = __ballot_sync(0xffffffffu, f);
                    const int t = __popc(fits & ((1u << lane) - 1u));
                    const bool taken = f && t < rq.nums;
                    const int dslot = (int)((mt >> M_SLOT_SHIFT) & 15u);
                    if (__popc(fits) < rq.nums) {
                        if (lane == 0) atomicOr(a.status, S_MISMATCH);
                    } else {
                        TICK(tC);
                        const int slot = take_slot();
                        if (taken) { // score.go:135-145,168-174
                            vgs_pick pk;
                            pk.slot = dslot, pk.usedmem = mem, pk.usedcores = rq.coresreq;
                            a.picks[pick_lo + t] = pk;
                            fs = wrap_sub(fs, 1), fm = wrap_sub(fm, mem), fc = wrap_sub(fc, rq.coresreq);
                            const int32_t used = wrap_sub(cn, fs);
                            const size_t at = (size_t)row * a.dmax + dslot;
                            const_cast<int32_t *>(a.raw.p[VGS_PLANE_USED])[at] = used;
                            const_cast<int32_t *>(a.raw.p[VGS_PLANE_USEDMEM])[at] = wrap_sub(tm, fm);
                            const_cast<int32_t *>(a.raw.p[VGS_PLANE_USEDCORES])[at] = wrap_sub(tc, fc);
                            if (used >= VGS_MAX_SHARE) atomicOr(a.status, S_RANGE);
                            mt &= ~(M_NOFREE | M_EXCL | M_FULL | M_BUSY370);
                            if (fs <= 0) mt |= M_NOFREE;
                            if (mt & M_IS100) mt |= M_EXCL;
                            if ((mt & M_CORENZ) && fc == 0) mt |= M_FULL;
                            if (mt & M_T370) mt |= M_BUSY370;
                        }
                        // the device's place in the walk order now
                        const bool dev = lane < ndv;
                        const uint32_t sk = (((mt >> M_NUMA_SHIFT) & 15u) << 25) | ((uint32_t)(fs + (int32_t)kFsBias) << 4) | (uint32_t)dslot;
                        int rank = 0;
#pragma unroll
                        for (int i = 0; i < DM; i++) {
                            const uint32_t ski = __shfl_sync(0xffffffffu, sk, i);
                            rank += (i < ndv && ski < sk) ? 1 : 0;
                        }
                        const int w = dev ? ndv - 1 - rank : lane;
                        int32_t *dst = reinterpret_cast<int32_t *>(s_cur + (size_t)(slot / kWarp) * kTile4);
                        const int dcol = slot % kWarp;
                        __syncwarp(); // every lane has read its words
                        if (lane < DM) {
                            dst[tile_at<DM>(W_TOTALMEM, w, dcol)] = tm, dst[tile_at<DM>(W_FREEMEM, w, dcol)] = fm;
                            dst[tile_at<DM>(W_FREECORES, w, dcol)] = fc, dst[tile_at<DM>(W_COUNT, w, dcol)] = cn;
                            dst[tile_at<DM>(W_FREESLOTS, w, dcol)] = fs, dst[tile_at<DM>(W_META, w, dcol)] = (int32_t)mt;
                            dst[tile_at<DM>(W_TOTALCORE, w, dcol)] = tc;
                        }
                        TICK(tD);
                    }
                } else {
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
                        const int slot = take_slot();
                        // its words again, from the columns the resolve holds in registers (lane = device slot), into T
                        prepare_lanes<DM>(row, ndv, 0u, used, fx.count, usedmem, fx.totalmem, usedcores, fx.totalcore, fx.numa,
                                          fx.type, a.raw.idle_types, nullptr, a.status, s_cur + (size_t)(slot / kWarp) * kTile4,
                                          slot % kWarp);
                        TICK(tD);
                    }
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
    // ---- fit counts: + the matrix's count - T's rows (as of the pod's turn) that fitted in the matrix
    {
        constexpr int kPer = kDrainThreads / kDrainChunk; // threads per pod
        const int k = tid / kPer, part = tid % kPer;
        const int n = k < n_done ? s_nt_at[k] : -1;
        const uint32_t *mine = a.ob + (size_t)k * a.n_tiles * kWarp;
        int32_t was = 0;
        for (int s = part; s < n; s += kPer) was += mine[s_trow[s]] != 0u ? 1 : 0;
#pragma unroll
        for (int d = 1; d < kPer; d <<= 1) was += __shfl_xor_sync(0xffffffffu, was, d);
        // (n < 0: the pod scanned -- it counted the rows outside T itself, and T live)
        if (part == 0 && n >= 0) a.results[k].fit_count += s_count0[k] - was;
    }
    if (warp == kWarps - 1) asm volatile("cp.async.wait_all;" ::: "memory"); // (a fetch for a pod that was never reached)
    if (n_done < a.n_pods && tid == 0) *a.stop = a.first_pod + n_done;
    // ---- the rows this chunk changed go on to k_drain_apply (walk copy) and to the next chunk (its T), W_SCORE1 put right
    __syncthreads();
    if (warp == 0) {
        int n_out = 0;
        for (int t = 0; t < kSlotTiles; t++) {
            const uint32_t m = s_mod[t];
            const bool mine = (m >> lane) & 1u;
            const int at = n_out + __popc(m & ((1u << lane) - 1u));
            if (mine) a.carry_out->rows[at] = s_trow[t * kWarp + lane];
            n_out += __popc(m);
        }
        if (lane == 0) a.carry_out->n = n_out;
    }
    // word copy: one warp per changed slot in turn
    {
        // position of slot s among the changed slots
        for (int s = warp; s < nT; s += kWarps) {
            if (!((s_mod[s / kWarp] >> (s % kWarp)) & 1u)) continue;
            int at = 0;
            for (int t = 0; t < s / kWarp; t++) at += __popc(s_mod[t]);
            at += __popc(s_mod[s / kWarp] & ((1u << (s % kWarp)) - 1u));
            if (lane < DM) {
                const int32_t *src = reinterpret_cast<const int32_t *>(s_cur + (size_t)(s / kWarp) * kTile4);
                int32_t *dst = reinterpret_cast<int32_t *>(a.words_out + (size_t)(at / kWarp) * kTile4);
                int32_t wd[kWalkWords];
#pragma unroll
                for (int k = 0; k < kWalkWords; k++) wd[k] = src[tile_at<DM>(k, lane, s % kWarp)];
                const uint32_t mt = (uint32_t)wd[W_META];
                const float s1 = (!(mt & M_PAD) && wd[W_FREESLOTS] > 0)
                                     ? container_score(wd[W_COUNT], wd[W_FREESLOTS], (int32_t)((mt >> M_NDEV_SHIFT) & 15u) - 1)
                                     : 0.0f;
                wd[W_SCORE1] = __float_as_int(s1);
#pragma unroll
                for (int k = 0; k < kWalkWords; k++) dst[tile_at<DM>(k, lane, at % kWarp)] = wd[k];
            }
        }
    }
#ifdef VGS_DRAIN_TIMING
    if (tid == 0) {
        atomicAdd(&a.stats[1], (int)(tA >> 6)), atomicAdd(&a.stats[2], (int)(tB >> 6)), atomicAdd(&a.stats[3], (int)(tC >> 6));
        atomicAdd(&a.stats[4], (int)(tD >> 6)), atomicAdd(&a.stats[5], (int)(tE >> 6));
        atomicAdd(&a.stats[6], (int)(tB1 >> 6)), atomicAdd(&a.stats[7], (int)(tB2 >> 6));
    }
#endif
}

// the rows a chunk changed, into the walk copy (after the pass that was reading it has finished)
template <int DM>
__global__ void __launch_bounds__(256) k_drain_apply(const DrainCarry *__restrict__ carry, const int4 *__restrict__ words,
                                                     int4 *__restrict__ walk, int32_t first_pod,
                                                     const int32_t *__restrict__ stop) {
    {
        const int32_t st0 = *reinterpret_cast<const volatile int32_t *>(stop);
        if (st0 != 0 && st0 <= first_pod) return; // that chunk was never placed: its carry block is an older chunk's
    }
    constexpr int kTile4 = kWalkWords * (DM / 4) * kWarp, kRow4 = kWalkWords * (DM / 4);
    const int n = carry->n;
    for (int i = threadIdx.x; i < n * kRow4; i += blockDim.x) {
        const int s = i / kRow4, k = i % kRow4, row = carry->rows[s];
        walk[(size_t)(row / kWarp) * kTile4 + k * kWarp + row % kWarp] = words[(size_t)(s / kWarp) * kTile4 + k * kWarp + s % kWarp];
    }
}

// geometry shared by the launchers
static int drain_segs(int sms, int n_pods, int n_tiles) {
    // segments per pod: the select CTAs of a chunk fill the GPU about once, a segment is at least 64 tiles
    int segs = (sms + n_pods - 1) / n_pods;
    if (segs > kDrainMaxSegs) segs = kDrainMaxSegs;
    if (segs > (n_tiles + 63) / 64) segs = (n_tiles + 63) / 64;
    return segs < 1 ? 1 : segs;
}

template <int DM, bool WALK>
static cudaError_t drain_eval_launch(int sms, const DrainPass &p, cudaStream_t st) {
    // one SM is left to the commit kernel of the chunk before (a persistent grid that counted on it would run a
    // second wave); items = (tile, part of the pods): at least one per resident warp
    const long long ctas = (long long)(sms > 1 ? sms - 1 : 1) * (WALK ? 2 : 3), warps = ctas * (kScoreThreads / 32);
    int psplit = (int)((warps + p.n_tiles - 1) / p.n_tiles);
    if (psplit < 1) psplit = 1;
    if (psplit > p.n_pods) psplit = p.n_pods;
    const long long items = (long long)p.n_tiles * psplit;
    long long grid = (items + kScoreThreads / 32 - 1) / (kScoreThreads / 32);
    if (grid > ctas) grid = ctas;
    if (grid < 1) grid = 1;
    k_drain_eval<DM, WALK><<<(int)grid, kScoreThreads, 0, st>>>(p.walk, p.n_tiles, p.reqs, p.pods, p.n_pods, psplit, p.ob, p.stop);
    return cudaGetLastError();
}

template <int DM>
static cudaError_t drain_eval_t(int sms, const DrainPass &p, cudaStream_t st) {
    cudaError_t e;
    if (p.any_pair && (e = drain_eval_launch<DM, false>(sms, p, st)) != cudaSuccess) return e;
    if (p.any_walk && (e = drain_eval_launch<DM, true>(sms, p, st)) != cudaSuccess) return e;
    const int segs = drain_segs(sms, p.n_pods, p.n_tiles);
    k_drain_select<<<dim3(segs, p.n_pods), kSelectThreads, 0, st>>>(p.ob, p.n_tiles, p.row_base, segs, p.seglist, p.segcount,
                                                                     p.done, p.topk, p.counts, p.stop);
    return cudaGetLastError();
}

cudaError_t launch_drain_pass(int dm, int sms, const DrainPass &p, cudaStream_t st) {
    if (p.n_pods <= 0 || p.n_pods > kDrainChunk || p.n_tiles <= 0) return cudaErrorInvalidValue;
    return dm == 8 ? drain_eval_t<8>(sms, p, st) : drain_eval_t<12>(sms, p, st);
}

template <int DM>
static cudaError_t drain_commit_t(int dev, const DrainChunk &c, cudaStream_t st) {
    constexpr size_t smem = (size_t)(kDrainSlots / kWarp) * kWalkWords * (DM / 4) * kWarp * sizeof(int4);
    static bool attr[64] = {}; // (static + dynamic shared memory of the commit kernel passes 48 KB)
    if (dev >= 0 && dev < 64 && !attr[dev]) {
        cudaError_t ea = cudaFuncSetAttribute(k_drain_commit<DM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ea != cudaSuccess) return ea;
        attr[dev] = true;
    }
    k_drain_commit<DM><<<1, kDrainThreads, smem, st>>>(c);
    return cudaGetLastError();
}

cudaError_t launch_drain_commit(int dm, int dev, int sms, const RawPlanes &raw, int32_t n_nodes, int32_t dmax,
                                const DrainPass &p, const int32_t *pick_begin, bool first, const void *carry_in,
                                const int4 *words_in, void *carry_out, int4 *words_out, int32_t first_pod,
                                int32_t max_rescans, vgs_result *results, vgs_pick *picks, uint32_t *status, int32_t *stats, cudaStream_t st) {
    if (p.n_pods <= 0 || p.n_pods > kDrainChunk || p.n_tiles <= 0) return cudaErrorInvalidValue;
    DrainChunk c;
    c.walk = p.walk, c.n_tiles = p.n_tiles, c.n_nodes = n_nodes, c.row_base = p.row_base, c.dmax = dmax;
    c.raw = raw;
    c.reqs = p.reqs, c.pods = p.pods, c.pick_begin = pick_begin, c.n_pods = p.n_pods;
    c.ob = p.ob, c.topk = p.topk, c.counts = p.counts;
    c.first = first ? 1 : 0;
    c.carry_in = static_cast<const DrainCarry *>(carry_in), c.words_in = words_in;
    c.carry_out = static_cast<DrainCarry *>(carry_out), c.words_out = words_out;
    c.results = results, c.picks = picks, c.status = status, c.stats = stats;
    c.first_pod = first_pod, c.max_rescans = max_rescans, c.stop = p.stop;
    return dm == 8 ? drain_commit_t<8>(dev, c, st) : drain_commit_t<12>(dev, c, st);
}

cudaError_t launch_drain_apply(int dm, const void *carry, const int4 *carry_words, int4 *walk, int32_t first_pod,
                               const int32_t *stop, cudaStream_t st) {
    if (dm == 8) k_drain_apply<8><<<1, 256, 0, st>>>(static_cast<const DrainCarry *>(carry), carry_words, walk, first_pod, stop);
    else k_drain_apply<12><<<1, 256, 0, st>>>(static_cast<const DrainCarry *>(carry), carry_words, walk, first_pod, stop);
    return cudaGetLastError();
}

size_t drain_carry_bytes() { return sizeof(DrainCarry); }

} // namespace vgs
