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
#include "vgs_walk_eval.cuh"

namespace vgs {

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
        WalkNode<DM> nd; // per node, for the whole item: Count, the static meta bits, the initial keys
        walk_node_load<DM, true>(tp, lane, nd);
        const uint32_t row0 = (uint32_t)(row_base + tile * kWarp);
        const int row = tile * kWarp + lane;

        for (int pp = 0; pp < npod; pp += kWarp) {
            unsigned long long mykey = 0ull;
            int32_t mycnt = 0;
            const int kend = min(kWarp, npod - pp);
            for (int k = 0; k < kend; k++) {
                const GenPod pod = pods[pod0 + pp + k];
                bool alive;
                float score;
                // (pp + k == 0: first pod of this warp's run over the tile, nothing parked yet)
                walk_eval<DM, true, true>(
                    nd, pod,
                    [&](int r) { return FILTER ? reqs[pod.req_begin + r] : load_req(&reqs[pod.req_begin + r]); }, one,
                    s_save, (pp + k) > 0, alive, score);
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
