// vgs_kernels.h -- launch interface between the C ABI (vgs_capi.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vgpuscore.h"
#include "vgs_layout.h"

namespace vgs {

struct RawPlanes {
    const int32_t *p[VGS_NUM_PLANES];
    const int32_t *ndev;
    uint64_t idle_types; // bit t: type id t carries VGS_TYPE_IDLE_SHARE
};
struct RawPlanesMut {
    int32_t *p[VGS_NUM_PLANES];
    int32_t *ndev;
};

struct ScoreArgs {
    const int4 *walk;
    int32_t n_tiles, n_nodes, row_base;
    unsigned long long *keys;
    int32_t *counts;
    uint32_t *vfit; // verify mode only, else nullptr
    float *vscore;
    int32_t vwords;
};

inline int walk_dm(int dmax) { return dmax <= 8 ? 8 : 12; }
inline size_t walk_int4_per_tile(int dm) { return (size_t)kWalkWords * (dm / 4) * kWarp; }

// zero-copy upload: src planes are host-pinned (device-accessible) memory, 16-byte aligned
cudaError_t launch_copy_planes(const int32_t *const src[VGS_NUM_PLANES], const int32_t *src_ndev,
                               int32_t *const dst[VGS_NUM_PLANES], int32_t *dst_ndev, size_t cells, size_t n_nodes,
                               cudaStream_t st);
// the same from the narrow wire format: src[f] / width[f] for the 8 planes and (index VGS_NUM_PLANES) ndev; elements of
// 1, 2 or 4 bytes, signed; src in device-accessible memory (pinned host or device), 16-byte aligned
cudaError_t launch_copy_planes_narrow(const void *const src[VGS_NUM_PLANES + 1], const uint8_t width[VGS_NUM_PLANES + 1],
                                      int32_t *const dst[VGS_NUM_PLANES], int32_t *dst_ndev, size_t cells, size_t n_nodes,
                                      cudaStream_t st);
cudaError_t launch_prepare(int dm, const RawPlanes &raw, int32_t n_nodes, int32_t dmax, int4 *walk,
                           int32_t n_tiles, uint32_t *status, const int32_t *rowlist,
                           int32_t n_list, cudaStream_t st);
cudaError_t launch_score_single(int dm, bool nums1, int dev, const ScoreArgs &a, const DevReq *reqs,
                                const int32_t *orig, int32_t n_pods, cudaStream_t st);
cudaError_t launch_score_general(int dm, int dev, const ScoreArgs &a, const DevReq *reqs,
                                 const GenPod *pods, int32_t n_pods, cudaStream_t st);
// the walk kernel (vgs_general.cu): multi-request pods by ordering keys; pods flagged G_LEGACY must not come here
cudaError_t launch_score_walk(int dm, int dev, const ScoreArgs &a, const DevReq *reqs, const GenPod *pods,
                              int32_t n_pods, cudaStream_t st);
// The single-pod Filter call as ONE launch: the pod by value, the outputs it writes (device arrays and the pinned
// host mirrors), a CTA counter that is zero at rest.
constexpr int kFilterMaxReqs = 8;
struct FilterOne {
    GenPod pod; // req_begin = 0, orig = 0
    DevReq reqs[kFilterMaxReqs];
    int32_t pick_begin[kFilterMaxReqs + 1];
    RawPlanes raw;
    int32_t dmax;
    vgs_result *results, *host_results;
    vgs_pick *picks, *host_picks;
    uint32_t *status, *host_status;
    unsigned int *done;
    int4 *commit_walk; // sequential drain: not null = fold the winner's picks into the raw planes and re-prepare its row
};
cudaError_t launch_filter_single(int dm, bool nums1, int dev, const ScoreArgs &a, const FilterOne &fo,
                                 cudaStream_t st);
cudaError_t launch_filter_one(int dm, int dev, const ScoreArgs &a, const FilterOne &fo, cudaStream_t st);
// Batch-scale drain (vgs_drain.cu), one chunk of <= kDrainChunk pods.
//   launch_drain_pass    the chunk's pods against the whole walk copy: score bits of every (pod, node) into
//                        ob[kDrainChunk][n_tiles * 32], then per pod the fit count and the best keys into
//                        counts[kDrainChunk] / topk[kDrainChunk][kDrainListWords]
//   launch_drain_commit  the chunk in pod order by one CTA: results, picks, usage into the raw planes; the rows it
//                        changed go into the carry block (and NOT into the walk copy: the next chunk's pass may be
//                        reading it) -- that pass may therefore be one chunk stale: `first` = it is not
//   launch_drain_apply   the carry block into the walk copy
constexpr int kDrainMaxSegs = 8;
constexpr int kDrainListWords = kDrainTop + 1;
struct DrainPass {
    const int4 *walk;
    int32_t n_tiles, row_base;
    const DevReq *reqs;
    const GenPod *pods; // the chunk's, batch order
    int32_t n_pods;
    bool any_pair, any_walk; // pods of the chunk that take the pair / the walk evaluation (vgs_pair_eval.cuh)
    uint32_t *ob;
    unsigned long long *seglist; // scratch of the select kernel: [pod][kDrainMaxSegs][kDrainListWords]
    int32_t *segcount;           // [pod][kDrainMaxSegs]
    unsigned int *done;          // [kDrainChunk], zero at rest
    unsigned long long *topk;    // [pod][kDrainListWords]
    int32_t *counts;             // [pod]
    int32_t *stop;               // one word, 0 at the start of a drain: first pod the chunks gave up on (vgs_drain.cu)
};
cudaError_t launch_drain_pass(int dm, int sms, const DrainPass &p, cudaStream_t st);
cudaError_t launch_drain_commit(int dm, int dev, int sms, const RawPlanes &raw, int32_t n_nodes, int32_t dmax,
                                const DrainPass &p, const int32_t *pick_begin, bool first, const void *carry_in,
                                const int4 *words_in, void *carry_out, int4 *words_out, int32_t first_pod,
                                int32_t max_rescans, vgs_result *results, vgs_pick *picks, uint32_t *status, int32_t *stats, cudaStream_t st);
cudaError_t launch_drain_apply(int dm, const void *carry, const int4 *carry_words, int4 *walk, int32_t first_pod,
                               const int32_t *stop, cudaStream_t st);
size_t drain_carry_bytes();
// launch geometry of a thread-per-node score kernel (persistent grid over (tile group, pod chunk) items)
cudaError_t plan_score_launch(const void *kernel, int dev, int32_t n_tiles, int32_t n_pods, int max_warps,
                              int smem_per_pod, ScoreLaunch *L);
// K1t: fill the bit tables of a sliced class on the device (table region zeroed, value arrays uploaded)
cudaError_t launch_slice_tables(const SliceTab &tb, const DevReq *dreq_class, cudaStream_t st);
// pod-sliced K1 for one class (all pods: one request, Nums == nums in {1,2}, walk DM == 8)
// (K1p collect / sort / classify, then the score kernel as scouts + main pass, or one pass on a small
// snapshot; *plan receives the split of the class's words, the launch geometry and the number of kernels launched)
cudaError_t launch_score_sliced(int nums, int dev, const ScoreArgs &a, const SliceTab &tb, int n_types,
                                uint32_t *status, cudaStream_t st, const vgs_tuning &tune, vgs_class_plan *plan,
                                cudaStream_t st_ns = nullptr, cudaEvent_t ev_ns = nullptr, cudaEvent_t ev_sc = nullptr,
                                cudaStream_t st_sc = nullptr);
// (st_ns / ev_ns: optional second stream + event on which the node-sliced words of the class may run as soon
// as K1p's records exist; the caller joins st_ns with whatever consumes the keys.  ev_sc: event for "the node-sliced
// scout launch on st_ns is done", which the class's main launch on st waits for)
cudaError_t launch_merge(const unsigned long long *gk, const int32_t *gc, int32_t n_shards,
                         int32_t stride, int32_t cstride, int32_t n_pods, unsigned long long *keys,
                         int32_t *counts, cudaStream_t st);
// K3 also clears keys[p] / counts[p] (it is their last reader) and writes every pick of every pod,
// zeros where the pod has no winner on this shard: neither needs a memset before the next pass
cudaError_t launch_resolve(const RawPlanes &raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
                           const vgs_pod *pods, const DevReq *reqs, const int32_t *pick_begin,
                           int32_t n_pods, unsigned long long *keys, int32_t *counts,
                           vgs_result *results, vgs_pick *picks, uint32_t *status, cudaStream_t st,
                           int32_t *commit_rowlist = nullptr, vgs_result *host_results = nullptr,
                           vgs_pick *host_picks = nullptr, bool keep_candidates = false,
                           uint32_t *host_status = nullptr, int4 *commit_walk = nullptr, int commit_dm = 8);
// K2x: winner shard per pod out of all-gathered exchange blocks [keys | counts | picks] (see k2x_merge)
cudaError_t launch_exchange_merge(const void *gathered, int32_t n_shards, size_t stride_bytes, int32_t max_pods,
                                  size_t picks_off, const vgs_pod *pods, const int32_t *pick_begin, int32_t n_pods,
                                  unsigned long long *keys, int32_t *counts, vgs_result *results, vgs_pick *picks,
                                  cudaStream_t st);
cudaError_t launch_delta(const RawPlanesMut &raw, int32_t n_nodes, int32_t dmax, int32_t row_base,
                         const vgs_delta *deltas, int32_t n, int32_t *rowlist, uint32_t *status,
                         cudaStream_t st);

} // namespace vgs
