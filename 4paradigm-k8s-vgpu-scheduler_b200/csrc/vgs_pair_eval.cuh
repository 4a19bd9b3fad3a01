// vgs_pair_eval.cuh -- one (pod, node) pair for a pod that is ONE request without a numa group to choose (numa-bind
// only with Nums == 1): the first Nums fitting devices of the K0 walk order decide (score.go:92-149,180).  Shared by
// k1_score_single (thread = node of a tile, loop over pods) and the batch-scale drain (vgs_drain.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vgs_device.cuh"
#include "vgs_layout.h"

namespace vgs {

// the pod classes the pair evaluation covers (what vgs_capi.cu routes to classes A, B and C)
__host__ __device__ inline bool pair_pod(int32_t n_reqs, const DevReq &first) {
    return n_reqs == 1 && (!(first.flags & R_NUMA) || first.nums == 1);
}

template <int DM>
struct PairNode {
    int32_t totalmem[DM], freemem[DM], freecores[DM], meta[DM], count[DM], freeslots[DM];
    int32_t ndev;
};

// NC: through the read-only path (see tile_word in vgs_walk_eval.cuh); plain loads otherwise
template <int DM, bool NC>
__device__ __forceinline__ void pair_node_load(const int4 *tp, int lane, PairNode<DM> &nd) {
    auto word = [&](int w, int32_t(&dst)[DM]) {
        if constexpr (NC) {
            load_word<DM>(tp, w, lane, dst);
        } else {
#pragma unroll
            for (int q = 0; q < DM / 4; q++) {
                const int4 x = tp[(w * (DM / 4) + q) * kWarp + lane];
                dst[4 * q + 0] = x.x, dst[4 * q + 1] = x.y, dst[4 * q + 2] = x.z, dst[4 * q + 3] = x.w;
            }
        }
    };
    word(W_TOTALMEM, nd.totalmem);
    word(W_FREEMEM, nd.freemem);
    word(W_FREECORES, nd.freecores);
    word(W_META, nd.meta);
    word(W_COUNT, nd.count);
    word(W_FREESLOTS, nd.freeslots);
    nd.ndev = (nd.meta[0] >> M_NDEV_SHIFT) & 15;
}

template <int DM>
__device__ __forceinline__ void pair_eval(const PairNode<DM> &nd, const DevReq &rq, bool &fit, float &score) {
    int32_t cnt = 0, total = 0, free_ = 0;
    const bool usepct = rq.flags & R_USEPCT;
#pragma unroll
    for (int j = 0; j < DM; j++) {
        const int32_t mem = usepct ? pct_mem(nd.totalmem[j], rq.pct) : rq.memfix;
        // (& on purpose: no short-circuit branches)
        const bool f = (int)!(nd.meta[j] & rq.blockmask) & (int)type_ok(rq.mask_lo, rq.mask_hi, nd.meta[j]) &
                       (int)(nd.freemem[j] >= mem) & (int)(nd.freecores[j] >= rq.coresreq);
        const bool take = (int)f & (int)(cnt < rq.nums);
        cnt += f ? 1 : 0;
        total += take ? nd.count[j] : 0;
        free_ += take ? nd.freeslots[j] : 0;
    }
    fit = cnt >= rq.nums;
    // (a node that is out divides 1 by 1: a zero on either side of the division sends the whole warp through its slow path)
    score = container_score(fit ? total : 1, fit ? free_ : 1, nd.ndev - rq.nums);
    score = fit ? score : 0.0f;
}

} // namespace vgs
