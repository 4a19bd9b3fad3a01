// vgs_capi.cu -- the C ABI of libvgpuscore.so (include/vgpuscore.h): context, device
// memory, request packing, kernel sequencing.  No CPU fallback anywhere: every scoring
// entry point ends in a kernel launch or an error code.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/vgpuscore.h"
#include "vgs_kernels.h"
#include "vgs_pair_eval.cuh"
#include "vgs_layout.h"

using namespace vgs;

namespace {

enum { EV_UP0, EV_UP1, EV_DR1, EV_PH0, EV_PH1, EV_SC0, EV_SC1, EV_RS0, EV_RS1, EV_DH0, EV_DH1, EV_N };

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == dev) { // the usual case: nothing to switch, nothing to restore
            ok = true;
            prev = -1;
        } else {
            ok = cudaSetDevice(dev) == cudaSuccess;
        }
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

} // namespace

constexpr int32_t kDirectOutPods = 64;

struct vgs_ctx {
    vgs_config cfg{};
    int dm = 8;
    cudaStream_t st = nullptr;
    bool own_stream = false;
    std::mutex mu;
    std::string err;

    // snapshot
    int32_t *d_raw[VGS_NUM_PLANES] = {};
    int32_t *d_ndev = nullptr;
    int4 *d_walk = nullptr;
    char *d_nstage = nullptr;     // narrow planes from pageable memory land here first (allocated at the first such upload)
    uint32_t *d_status = nullptr;
    uint32_t *h_status = nullptr; // pinned
    // the small arrays every score launch touches (request staging, bit tables of the two sliced classes, candidate
    // block) are carved from ONE allocation: their addresses relative to each other decide how the K1 phase runs (a
    // 512 KB shift of this group cost 11 % before), so they must not depend on what else the context allocates
    unsigned char *d_arena = nullptr;
    unsigned int *d_done = nullptr; // CTA counter of the single-launch Filter kernel, zero at rest
    // batch-scale drain: per-chunk scratch -- candidate lists, counts, and (allocated by the first chunked drain) the
    // score bits of every (pod of the chunk, node)
    unsigned long long *d_topk[2] = {}, *d_seglist = nullptr;
    int32_t *d_dcounts[2] = {}, *d_segcount = nullptr;
    unsigned int *d_seldone = nullptr;
    uint32_t *d_drain_ob[2] = {};
    void *d_carry[2] = {}; // rows a chunk changed + their words (vgs_drain.cu)
    int4 *d_carry_words[2] = {};
    cudaEvent_t ev_pass[2] = {}, ev_commit[2] = {};
    int32_t *d_drain_stats = nullptr, *h_drain_stats = nullptr;
    int sms = 0;
    int32_t n_nodes = 0, dmax = 0, n_tiles = 0;
    bool have_snapshot = false;
    bool snap_pending = false;    // vgs_snapshot_upload_async: copy + K0 queued, status not looked at yet
    cudaEvent_t ev_snap = nullptr;
    bool stage_in_flight = false; // the pinned staging block may still be the source of a queued copy
    uint64_t version = 0;
    uint64_t idle_types = 0; // vgs_snapshot_type_flags: types carrying VGS_TYPE_IDLE_SHARE
    vgs_delta *d_deltas = nullptr, *h_deltas = nullptr;
    int32_t *d_rowlist = nullptr;
    int32_t max_deltas = 0;
    cudaEvent_t ev_delta = nullptr; // the pinned delta staging buffer has been consumed
    bool delta_in_flight = false;

    // pod batch
    unsigned char *h_stage = nullptr, *d_stage = nullptr; // one block, carved below
    size_t stage_bytes = 0;
    size_t off_reqs = 0, off_pods = 0, off_pick = 0, off_reqsA = 0, off_origA = 0, off_reqsB = 0,
           off_origB = 0, off_podsG = 0;
    size_t off_reqsC = 0, off_origC = 0;
    size_t off_genall = 0; // GenPod of EVERY pod, batch order (drain: one pod at a time)
    int32_t n_pods = 0, n_reqs = 0, nA = 0, nB = 0, nC = 0, nG = 0, n_picks = 0;
    int32_t nGL = 0; // of the nG general pods, those on the all-pairs kernel (stored from the END of the podsG region)
    // pod-sliced classes (A: Nums == 1, B: Nums == 2): bit tables, see SliceTab
    uint32_t *h_slice[2] = {}, *d_slice[2] = {};
    unsigned char *d_prep[2] = {}; // K1p products: SliceDev tiles, pair classes, score set, classes
    size_t slice_words = 0;
    SliceTab tab[2] = {};
    bool slice_on[2] = {false, false};
    vgs_tuning tune{};           // copied from vgs_config.tuning at vgs_init (defaults when NULL)
    vgs_plan plan{};             // how the last score pass ran (vgs_last_plan)
    bool snap_slice_ok = false;  // snapshot satisfies the sliced kernel's preconditions
    bool snap_bigmem = false;    // some Totalmem of the resident snapshot is outside the sliced kernels' range;
                                 // sticky until the next full upload (deltas never touch Totalmem)
    int32_t n_types_used = VGS_MAX_TYPES; // largest type id of the snapshot + 1
    bool have_pods = false, scored = false, resolved = false;
    bool direct_out = false; // K3 of the current batch wrote h_results / h_picks itself
    size_t picks_off = 0, xblock_bytes = 0; // exchange block = [keys | counts | pad | picks], one allocation
    int32_t dirty_pods = 0; // leading slots of d_keys / d_counts that may be non-zero (K3 clears behind itself)

    // results
    unsigned long long *d_keys = nullptr;
    int32_t *d_counts = nullptr;
    vgs_result *d_results = nullptr, *h_results = nullptr;
    vgs_pick *d_picks = nullptr, *h_picks = nullptr;
    uint32_t *d_vfit = nullptr;
    float *d_vscore = nullptr;
    int32_t vwords = 0;

    // the pod classes of one batch are independent: their kernel chains run on side streams
    // forked from / joined to `st` with events
    cudaStream_t side[2] = {}, side_sc[2] = {};
    // a pod batch that arrives while an asynchronous snapshot upload is still crossing PCIe goes up (and has its
    // bit tables built) on this stream, beside the copy; the score pass joins the two
    cudaStream_t st_pods = nullptr;
    cudaEvent_t ev_idle = nullptr, ev_pods = nullptr; // `st` as it stood before the upload was queued / pod work done
    bool pods_on_side = false;
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {}, ev_ns[2] = {}, ev_sc[2] = {};
    cudaEvent_t ev[EV_N] = {};
    bool ev_set[EV_N] = {};
    int ev_alias[EV_N] = {};  // which recorded event stands for this boundary (see rec)
    bool chained = false;     // inside vgs_score_batch: one phase's end IS the next one's start
    long long launches = 0;
};

namespace {

int fail_cuda(vgs_ctx *c, cudaError_t e, const char *what) {
    char buf[512];
    snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    if (c) c->err = buf;
    return VGS_ECUDA;
}
int fail(vgs_ctx *c, int code, const char *what) {
    if (c) c->err = what;
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) return fail_cuda(c, e__, #call);                                   \
    } while (0)

void rec(vgs_ctx *c, int which) {
    // back-to-back phases of one call share their boundary event (nothing is queued in between)
    if (c->chained && (which == EV_SC0 || which == EV_RS0 || which == EV_DH0) && c->ev_set[which - 1]) {
        c->ev_alias[which] = c->ev_alias[which - 1];
        c->ev_set[which] = true;
        return;
    }
    cudaEventRecord(c->ev[which], c->st);
    c->ev_alias[which] = which;
    c->ev_set[which] = true;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

RawPlanes raw_const(const vgs_ctx *c) {
    RawPlanes r;
    for (int f = 0; f < VGS_NUM_PLANES; f++) r.p[f] = c->d_raw[f];
    r.ndev = c->d_ndev;
    r.idle_types = c->idle_types;
    return r;
}

int check_status(vgs_ctx *c) {
    // d_status -> pinned word; stream must be synchronised by the caller afterwards
    uint32_t s = *c->h_status;
    if (s & S_NDEV) return fail(c, VGS_EDMAX, "ndev outside [0, dmax]");
    if (s & S_TYPE) return fail(c, VGS_ERANGE, "type id outside [0, VGS_MAX_TYPES)");
    if (s & S_RANGE) return fail(c, VGS_ERANGE, "Count/Used outside [0, VGS_MAX_SHARE)");
    if (s & S_DELTA) return fail(c, VGS_EINVAL, "delta names a slot the node does not have");
    if (s & S_MISMATCH) return fail(c, VGS_EINTERNAL, "resolve kernel disagrees with score kernel");
    return VGS_OK;
}

// h_status holds the two status words of the resident snapshot: errors, and what the score path may assume
int apply_status(vgs_ctx *c) {
    int rc = check_status(c);
    // S_BIGMEM is a property of the planes, not an error: K0 raises it only for rows it prepares, so it is
    // remembered on the host -- clearing the error bits below (or a delta that re-prepares other rows) must
    // not make the sliced kernels believe Totalmem * pct cannot wrap
    if (*c->h_status & S_BIGMEM) c->snap_bigmem = true;
    if (*c->h_status & S_DICTFULL) c->plan.dict_full = 1;
    c->snap_slice_ok = rc == VGS_OK && !c->snap_bigmem && c->dm == 8;
    c->n_types_used = (int32_t)c->h_status[1] + 1;
    if (rc != VGS_OK) cudaMemsetAsync(c->d_status, 0, sizeof(uint32_t), c->st);
    return rc;
}

// an asynchronous upload is completed by the first call that needs the snapshot
int finish_snapshot(vgs_ctx *c) {
    if (!c->snap_pending) return VGS_OK;
    c->snap_pending = false;
    CU(cudaEventSynchronize(c->ev_snap));
    int rc = apply_status(c);
    if (rc != VGS_OK) return rc;
    c->have_snapshot = true;
    return VGS_OK;
}

int pull_status(vgs_ctx *c) {
    CU(cudaMemcpyAsync(c->h_status, c->d_status, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    c->stage_in_flight = false;
    int rc = apply_status(c);
    return rc;
}

int prepare_all(vgs_ctx *c) {
    CU(launch_prepare(c->dm, raw_const(c), c->n_nodes, c->dmax, c->d_walk, c->n_tiles, c->d_status,
                      nullptr, 0, c->st));
    c->launches++;
    return VGS_OK;
}

void free_all(vgs_ctx *c) {
    for (int f = 0; f < VGS_NUM_PLANES; f++) cudaFree(c->d_raw[f]);
    cudaFree(c->d_ndev);
    cudaFree(c->d_walk);
    cudaFree(c->d_nstage);
    if (c->ev_snap) cudaEventDestroy(c->ev_snap);
    if (c->ev_delta) cudaEventDestroy(c->ev_delta);
    cudaFree(c->d_status);
    cudaFree(c->d_done);
    cudaFree(c->d_arena);
    for (int k = 0; k < 2; k++) {
        cudaFree(c->d_topk[k]);
        cudaFree(c->d_dcounts[k]);
        cudaFree(c->d_drain_ob[k]);
        if (c->ev_pass[k]) cudaEventDestroy(c->ev_pass[k]);
        if (c->ev_commit[k]) cudaEventDestroy(c->ev_commit[k]);
        cudaFree(c->d_carry[k]);
        cudaFree(c->d_carry_words[k]);
    }
    cudaFree(c->d_seglist);
    cudaFree(c->d_segcount);
    cudaFree(c->d_seldone);
    cudaFree(c->d_drain_stats);
    if (c->h_drain_stats) cudaFreeHost(c->h_drain_stats);
    cudaFreeHost(c->h_status);
    cudaFree(c->d_deltas);
    cudaFreeHost(c->h_deltas);
    cudaFree(c->d_rowlist);
    for (int k = 0; k < 2; k++) {
        cudaFreeHost(c->h_slice[k]);
        cudaFree(c->d_prep[k]);
    }
    cudaFreeHost(c->h_stage);
    cudaFree(c->d_results);
    cudaFreeHost(c->h_results);
    cudaFreeHost(c->h_picks);
    cudaFree(c->d_vfit);
    cudaFree(c->d_vscore);
    for (int i = 0; i < EV_N; i++)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int k = 0; k < 2; k++) {
        if (c->side[k]) cudaStreamDestroy(c->side[k]);
        if (c->side_sc[k]) cudaStreamDestroy(c->side_sc[k]);
        if (c->ev_join[k]) cudaEventDestroy(c->ev_join[k]);
        if (c->ev_ns[k]) cudaEventDestroy(c->ev_ns[k]);
        if (c->ev_sc[k]) cudaEventDestroy(c->ev_sc[k]);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->st_pods) cudaStreamDestroy(c->st_pods);
    if (c->ev_idle) cudaEventDestroy(c->ev_idle);
    if (c->ev_pods) cudaEventDestroy(c->ev_pods);
    if (c->own_stream && c->st) cudaStreamDestroy(c->st);
}

} // namespace

extern "C" {

int vgs_abi_version(void) { return VGS_ABI_VERSION; }

const char *vgs_strerror(int status) {
    switch (status) {
    case VGS_OK: return "ok";
    case VGS_EINVAL: return "invalid argument";
    case VGS_ECUDA: return "CUDA error";
    case VGS_ETOOBIG: return "exceeds the capacity given to vgs_init";
    case VGS_EDMAX: return "more devices per node than supported";
    case VGS_ERANGE: return "snapshot value out of range";
    case VGS_ESTATE: return "call sequence error";
    case VGS_EINTERNAL: return "internal self-check failed";
    case VGS_ELIMIT: return "larger than the fixed VGS_NOALLOC_MAX_* limits of vgs_score_batch_noalloc";
    case VGS_ENCCL: return "NCCL unavailable or collective failed";
    default: return "unknown status";
    }
}

const char *vgs_last_error(vgs_ctx *c) { return c ? c->err.c_str() : ""; }

int vgs_init(const vgs_config *cfg, vgs_ctx **out) {
    if (!cfg || !out) return VGS_EINVAL;
    *out = nullptr;
    if (cfg->max_nodes < 1 || cfg->max_pods < 1 || cfg->max_reqs < cfg->max_pods ||
        cfg->max_picks < cfg->max_reqs || cfg->max_devices < 1)
        return VGS_EINVAL;
    if (cfg->max_devices > VGS_MAX_DEVICES) return VGS_EDMAX;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return VGS_ECUDA; // no GPU, no fallback
    if (cfg->device < 0 || cfg->device >= ndev) return VGS_EINVAL;
    vgs_ctx *c = new (std::nothrow) vgs_ctx();
    if (!c) return VGS_EINVAL;
    c->cfg = *cfg;
    vgs_tuning_defaults(&c->tune);
    if (cfg->tuning) c->tune = *cfg->tuning;
    c->cfg.tuning = nullptr; // the caller's struct is not kept
    if (c->tune.slice_min < 1) c->tune.slice_min = 1;
    if (c->tune.slice_small < 1) c->tune.slice_small = 1;
    c->dm = walk_dm(cfg->max_devices);
    DeviceGuard g(cfg->device);
    auto bail = [&](int rc) {
        free_all(c);
        delete c;
        return rc;
    };
#define CUI(call)                                                                                  \
    do {                                                                                           \
        if ((call) != cudaSuccess) return bail(VGS_ECUDA);                                         \
    } while (0)
    if (!g.ok) return bail(VGS_ECUDA);
    if (cfg->stream) {
        c->st = (cudaStream_t)cfg->stream;
    } else {
        CUI(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
        c->own_stream = true;
    }
    for (int i = 0; i < EV_N; i++) CUI(cudaEventCreate(&c->ev[i]));
    CUI(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    CUI(cudaEventCreateWithFlags(&c->ev_snap, cudaEventDisableTiming));
    CUI(cudaStreamCreateWithFlags(&c->st_pods, cudaStreamNonBlocking));
    CUI(cudaEventCreateWithFlags(&c->ev_idle, cudaEventDisableTiming));
    CUI(cudaEventCreateWithFlags(&c->ev_pods, cudaEventDisableTiming));
    for (int k = 0; k < 2; k++) {
        CUI(cudaStreamCreateWithFlags(&c->side[k], cudaStreamNonBlocking));
        CUI(cudaStreamCreateWithFlags(&c->side_sc[k], cudaStreamNonBlocking));
        CUI(cudaEventCreateWithFlags(&c->ev_join[k], cudaEventDisableTiming));
        CUI(cudaEventCreateWithFlags(&c->ev_ns[k], cudaEventDisableTiming));
        CUI(cudaEventCreateWithFlags(&c->ev_sc[k], cudaEventDisableTiming));
    }
    const size_t cells = (size_t)cfg->max_nodes * cfg->max_devices;
    for (int f = 0; f < VGS_NUM_PLANES; f++) CUI(cudaMalloc(&c->d_raw[f], cells * sizeof(int32_t)));
    CUI(cudaMalloc(&c->d_ndev, (size_t)cfg->max_nodes * sizeof(int32_t)));
    const size_t max_tiles = ((size_t)cfg->max_nodes + kWarp - 1) / kWarp;
    CUI(cudaMalloc(&c->d_walk, max_tiles * walk_int4_per_tile(c->dm) * sizeof(int4)));
    CUI(cudaMalloc(&c->d_status, 2 * sizeof(uint32_t))); // [0] S_* flags, [1] largest type id in use
    CUI(cudaMemset(c->d_status, 0, 2 * sizeof(uint32_t)));
    CUI(cudaMallocHost(&c->h_status, 2 * sizeof(uint32_t)));
    CUI(cudaMalloc(&c->d_done, sizeof(unsigned int)));
    CUI(cudaMemset(c->d_done, 0, sizeof(unsigned int)));
    CUI(cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, cfg->device));
    c->max_deltas = 16384;
    CUI(cudaEventCreateWithFlags(&c->ev_delta, cudaEventDisableTiming));
    CUI(cudaMalloc(&c->d_deltas, c->max_deltas * sizeof(vgs_delta)));
    CUI(cudaMallocHost(&c->h_deltas, c->max_deltas * sizeof(vgs_delta)));
    CUI(cudaMalloc(&c->d_rowlist, c->max_deltas * sizeof(int32_t)));

    // pod staging block: [reqs][pods][pick_begin][reqsA][origA][reqsB][origB][podsG]
    const size_t P = cfg->max_pods, R = cfg->max_reqs;
    size_t o = 0;
    // worst-case size of the staging block; the offsets inside it are laid out per batch
    // (stage_layout) so that one batch is ONE host-to-device copy of its used prefix
    o = align_up(R * sizeof(DevReq), 256) + align_up(P * sizeof(vgs_pod), 256) +
        align_up((R + 1) * sizeof(int32_t), 256) + 3 * align_up(P * sizeof(DevReq), 256) +
        3 * align_up(P * sizeof(int32_t), 256) + 2 * align_up(P * sizeof(GenPod), 256);
    c->stage_bytes = o;
    CUI(cudaMallocHost(&c->h_stage, o));
    // arena layout: [stage | pad | slice 0 | slice 1 | candidate block], every piece on a 4 KB boundary
    {
        const size_t wp0 = align_up((P + 31) / 32, 32);
        const size_t rows0 = VGS_MAX_TYPES + 5 * (kSliceMaxDistinct + 1) + (kSliceMaxPct + 1);
        const size_t slice_b = (rows0 * wp0 + 2 * kSliceMaxDistinct + kSliceMaxPct + P + 64) * sizeof(uint32_t);
        const size_t picks_off0 = align_up(P * (sizeof(unsigned long long) + sizeof(int32_t)) + 8, 16);
        const size_t xblock_b = align_up(picks_off0 + (size_t)cfg->max_picks * sizeof(vgs_pick), 16);
        const size_t pad_b = (size_t)(c->tune.alloc_pad_kb > 0 ? c->tune.alloc_pad_kb : 0) * 1024;
        const size_t o_slice0 = align_up(o, 4096) + pad_b, o_slice1 = o_slice0 + align_up(slice_b, 4096);
        const size_t o_keys = o_slice1 + align_up(slice_b, 4096);
        CUI(cudaMalloc(&c->d_arena, o_keys + align_up(xblock_b, 4096)));
        c->d_stage = c->d_arena;
        c->d_slice[0] = reinterpret_cast<uint32_t *>(c->d_arena + o_slice0);
        c->d_slice[1] = reinterpret_cast<uint32_t *>(c->d_arena + o_slice1);
        c->d_keys = reinterpret_cast<unsigned long long *>(c->d_arena + o_keys);
    }

    {
        const size_t wp = align_up((P + 31) / 32, 32);
        const size_t rows = VGS_MAX_TYPES + 5 * (kSliceMaxDistinct + 1) + (kSliceMaxPct + 1);
        c->slice_words = rows * wp + 2 * kSliceMaxDistinct + kSliceMaxPct + P + 64;
        const size_t dinfo_b = max_tiles * 32 * 8 * sizeof(SliceDev), pcls_b = max_tiles * 32 * 32;
        const size_t hset_b = kSliceHash * sizeof(uint32_t), gvals_b = (kSliceDict + 4) * sizeof(uint32_t);
        for (int k = 0; k < 2; k++) {
            CUI(cudaMallocHost(&c->h_slice[k], c->slice_words * sizeof(uint32_t)));
            CUI(cudaMalloc(&c->d_prep[k], dinfo_b + pcls_b + hset_b + gvals_b));
            c->tab[k].dinfo = reinterpret_cast<uint4 *>(c->d_prep[k]);
            c->tab[k].pcls = c->d_prep[k] + dinfo_b;
            c->tab[k].hset = reinterpret_cast<uint32_t *>(c->d_prep[k] + dinfo_b + pcls_b);
            c->tab[k].gvals = reinterpret_cast<uint32_t *>(c->d_prep[k] + dinfo_b + pcls_b + hset_b);
        }
    }
    // keys and counts are one allocation ("candidate block": max_pods u64 keys, then max_pods
    // int32 counts) so that shards can exchange both with a single all-gather
    // ... and the picks follow in the same allocation ("exchange block"), so that a shard can also ship
    // the picks of its own best nodes in that one collective (vgs_exchange_block / vgs_exchange_merge)
    c->picks_off = align_up(P * (sizeof(unsigned long long) + sizeof(int32_t)) + 8, 16);
    c->xblock_bytes = align_up(c->picks_off + (size_t)cfg->max_picks * sizeof(vgs_pick), 16);
    c->d_counts = reinterpret_cast<int32_t *>(c->d_keys + P);
    CUI(cudaMemset(c->d_keys, 0, c->xblock_bytes));
    c->dirty_pods = 0;
    CUI(cudaMalloc(&c->d_results, P * sizeof(vgs_result)));
    CUI(cudaMallocHost(&c->h_results, P * sizeof(vgs_result)));
    c->d_picks = reinterpret_cast<vgs_pick *>(reinterpret_cast<unsigned char *>(c->d_keys) + c->picks_off);
    CUI(cudaMallocHost(&c->h_picks, (size_t)cfg->max_picks * sizeof(vgs_pick)));
    if (cfg->verify) {
        c->vwords = (int32_t)max_tiles;
        CUI(cudaMalloc(&c->d_vfit, P * max_tiles * sizeof(uint32_t)));
        CUI(cudaMalloc(&c->d_vscore, P * (size_t)cfg->max_nodes * sizeof(float)));
    }
    // the drain's small scratch blocks, after everything the score path uses (their addresses stay what they were)
    for (int k = 0; k < 2; k++) {
        CUI(cudaMalloc(&c->d_topk[k], (size_t)kDrainChunk * kDrainListWords * sizeof(unsigned long long)));
        CUI(cudaMalloc(&c->d_dcounts[k], (size_t)kDrainChunk * sizeof(int32_t)));
        CUI(cudaEventCreateWithFlags(&c->ev_pass[k], cudaEventDisableTiming));
        CUI(cudaEventCreateWithFlags(&c->ev_commit[k], cudaEventDisableTiming));
        CUI(cudaMalloc(&c->d_carry[k], drain_carry_bytes()));
        CUI(cudaMalloc(&c->d_carry_words[k], (size_t)(kDrainChunk / kWarp) * walk_int4_per_tile(c->dm) * sizeof(int4)));
    }
    CUI(cudaMalloc(&c->d_seglist, (size_t)kDrainChunk * kDrainMaxSegs * kDrainListWords * sizeof(unsigned long long)));
    CUI(cudaMalloc(&c->d_segcount, (size_t)kDrainChunk * kDrainMaxSegs * sizeof(int32_t)));
    CUI(cudaMalloc(&c->d_seldone, kDrainChunk * sizeof(unsigned int)));
    CUI(cudaMemset(c->d_seldone, 0, kDrainChunk * sizeof(unsigned int)));
    CUI(cudaMalloc(&c->d_drain_stats, 12 * sizeof(int32_t)));
    CUI(cudaMallocHost(&c->h_drain_stats, 12 * sizeof(int32_t)));
#undef CUI
    *out = c;
    return VGS_OK;
}

void vgs_destroy(vgs_ctx *c) {
    if (!c) return;
    {
        DeviceGuard g(c->cfg.device);
        cudaStreamSynchronize(c->st);
        free_all(c);
    }
    delete c;
}

// widths != nullptr: the narrow wire format -- planes[f] / ndev hold elements of widths[f] / widths[VGS_NUM_PLANES]
// bytes (host memory only)
static int snapshot_common(vgs_ctx *c, int32_t n_nodes, int32_t dmax,
                           const int32_t *const planes[VGS_NUM_PLANES], const int32_t *ndev,
                           cudaMemcpyKind kind, uint64_t *out_version, bool async = false,
                           const uint8_t *widths = nullptr) {
    if (!planes || (!ndev && !widths)) return fail(c, VGS_EINVAL, "null planes");
    for (int f = 0; widths && f <= VGS_NUM_PLANES; f++)
        if (widths[f] != 1 && widths[f] != 2 && widths[f] != 4) return fail(c, VGS_EINVAL, "plane width must be 1, 2 or 4 bytes");
    bool partial = false; // narrow format: a null plane keeps what is resident
    for (int f = 0; widths && f <= VGS_NUM_PLANES; f++) partial |= (f < VGS_NUM_PLANES ? (const void *)planes[f] : (const void *)ndev) == nullptr;
    if (partial) {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
        if (!c->have_snapshot || n_nodes != c->n_nodes || dmax != c->dmax)
            return fail(c, VGS_ESTATE, "planes left out need a resident snapshot of the same shape");
    }
    if (n_nodes < 0 || dmax < 1) return fail(c, VGS_EINVAL, "bad snapshot shape");
    if (dmax > VGS_MAX_DEVICES) return fail(c, VGS_EDMAX, "dmax > VGS_MAX_DEVICES");
    if (n_nodes > c->cfg.max_nodes || dmax > c->cfg.max_devices)
        return fail(c, VGS_ETOOBIG, "snapshot larger than vgs_init capacity");
    for (int f = 0; f < VGS_NUM_PLANES; f++)
        if (!planes[f] && !partial) return fail(c, VGS_EINVAL, "null plane");
    if (c->snap_pending) { // superseded before anybody looked at it
        c->snap_pending = false;
        CU(cudaEventSynchronize(c->ev_snap));
    }
    c->have_snapshot = false;
    c->snap_bigmem = false; // every row is prepared again below
    c->scored = c->resolved = false;
    const size_t cells = (size_t)n_nodes * dmax;
    if (async) CU(cudaEventRecord(c->ev_idle, c->st)); // everything queued so far: what a pod upload beside the copy waits for
    CU(cudaMemsetAsync(c->d_status, 0, 2 * sizeof(uint32_t), c->st));
    rec(c, EV_UP0);
    if (cells) {
        // pinned, 16-byte aligned host planes are read by a kernel straight over PCIe (faster
        // than the copy engine for this size); anything else goes through cudaMemcpyAsync
        bool zero_copy = kind == cudaMemcpyHostToDevice && c->tune.zero_copy_upload != 0;
        for (int f = 0; f <= VGS_NUM_PLANES && zero_copy; f++) {
            const void *p = f < VGS_NUM_PLANES ? (const void *)planes[f] : (const void *)ndev;
            if (!p) continue; // (left out)
            cudaPointerAttributes at;
            if (((uintptr_t)p & 15) || cudaPointerGetAttributes(&at, p) != cudaSuccess || at.type != cudaMemoryTypeHost)
                zero_copy = false;
        }
        cudaGetLastError(); // cudaPointerGetAttributes on pageable memory may leave an error behind
        if (widths) {
            const void *src[VGS_NUM_PLANES + 1];
            for (int f = 0; f <= VGS_NUM_PLANES; f++) src[f] = f < VGS_NUM_PLANES ? (const void *)planes[f] : (const void *)ndev;
            if (!zero_copy) { // pageable or unaligned: the narrow bytes go to a staging block, the kernel widens from there
                const size_t max_cells = (size_t)c->cfg.max_nodes * c->cfg.max_devices;
                if (!c->d_nstage)
                    CU(cudaMalloc(&c->d_nstage, (max_cells * VGS_NUM_PLANES + (size_t)c->cfg.max_nodes) * sizeof(int32_t) +
                                                    16 * (VGS_NUM_PLANES + 1)));
                size_t off = 0;
                for (int f = 0; f <= VGS_NUM_PLANES; f++) {
                    const size_t b = (f < VGS_NUM_PLANES ? cells : (size_t)n_nodes) * widths[f];
                    if (!src[f]) continue;
                    CU(cudaMemcpyAsync(c->d_nstage + off, src[f], b, cudaMemcpyHostToDevice, c->st));
                    src[f] = c->d_nstage + off;
                    off += align_up(b, 16);
                }
            }
            CU(launch_copy_planes_narrow(src, widths, c->d_raw, c->d_ndev, cells, (size_t)n_nodes, c->st));
            c->launches++;
        } else if (zero_copy) {
            CU(launch_copy_planes(planes, ndev, c->d_raw, c->d_ndev, cells, (size_t)n_nodes, c->st));
            c->launches++;
        } else {
            for (int f = 0; f < VGS_NUM_PLANES; f++)
                CU(cudaMemcpyAsync(c->d_raw[f], planes[f], cells * sizeof(int32_t), kind, c->st));
            CU(cudaMemcpyAsync(c->d_ndev, ndev, (size_t)n_nodes * sizeof(int32_t), kind, c->st));
        }
    }
    rec(c, EV_UP1);
    c->n_nodes = n_nodes;
    c->dmax = dmax;
    c->n_tiles = (n_nodes + kWarp - 1) / kWarp;
    int rc = prepare_all(c);
    if (rc != VGS_OK) return rc;
    rec(c, EV_DR1);
    if (async) { // the status words come back behind K0; finish_snapshot looks at them
        CU(cudaMemcpyAsync(c->h_status, c->d_status, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st));
        CU(cudaEventRecord(c->ev_snap, c->st));
        c->snap_pending = true;
        c->version++;
        if (out_version) *out_version = c->version;
        return VGS_OK;
    }
    rc = pull_status(c);
    if (rc != VGS_OK) return rc;
    c->have_snapshot = true;
    c->version++;
    if (out_version) *out_version = c->version;
    return VGS_OK;
}

int vgs_snapshot_upload(vgs_ctx *c, int32_t n_nodes, int32_t dmax,
                        const int32_t *const planes[VGS_NUM_PLANES], const int32_t *ndev,
                        uint64_t *out_version) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return snapshot_common(c, n_nodes, dmax, planes, ndev, cudaMemcpyHostToDevice, out_version);
}

int vgs_snapshot_upload_async(vgs_ctx *c, int32_t n_nodes, int32_t dmax,
                              const int32_t *const planes[VGS_NUM_PLANES], const int32_t *ndev,
                              uint64_t *out_version) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return snapshot_common(c, n_nodes, dmax, planes, ndev, cudaMemcpyHostToDevice, out_version, true);
}

int vgs_snapshot_upload_narrow(vgs_ctx *c, int32_t n_nodes, int32_t dmax, const void *const planes[VGS_NUM_PLANES],
                               const void *ndev, const uint8_t widths[VGS_NUM_PLANES + 1], int32_t asynchronous,
                               uint64_t *out_version) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (!widths) return fail(c, VGS_EINVAL, "null widths");
    return snapshot_common(c, n_nodes, dmax, reinterpret_cast<const int32_t *const *>(planes),
                           static_cast<const int32_t *>(ndev), cudaMemcpyHostToDevice, out_version, asynchronous != 0, widths);
}

int vgs_snapshot_set_device(vgs_ctx *c, int32_t n_nodes, int32_t dmax,
                            const int32_t *const dev_planes[VGS_NUM_PLANES], const int32_t *dev_ndev,
                            uint64_t *out_version) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return snapshot_common(c, n_nodes, dmax, dev_planes, dev_ndev, cudaMemcpyDeviceToDevice, out_version);
}

int vgs_snapshot_type_flags(vgs_ctx *c, int32_t n_types, const uint32_t *flags) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (n_types < 0 || n_types > VGS_MAX_TYPES || (n_types > 0 && !flags))
        return fail(c, VGS_EINVAL, "type flags: n_types outside [0, VGS_MAX_TYPES]");
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    uint64_t m = 0;
    for (int32_t t = 0; t < n_types; t++)
        if (flags[t] & VGS_TYPE_IDLE_SHARE) m |= 1ull << t;
    if (m == c->idle_types) return VGS_OK;
    c->idle_types = m;
    if (!c->have_snapshot) return VGS_OK;
    c->scored = c->resolved = false; // the walk copy changes under a resident batch
    return prepare_all(c);
}

int vgs_snapshot_refresh(vgs_ctx *c) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    if (!c->have_snapshot) return fail(c, VGS_ESTATE, "refresh before snapshot upload");
    rec(c, EV_UP1);
    int rc = prepare_all(c);
    if (rc != VGS_OK) return rc;
    rec(c, EV_DR1);
    return VGS_OK;
}

int vgs_snapshot_delta(vgs_ctx *c, int32_t n, const vgs_delta *deltas) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    if (!c->have_snapshot) return fail(c, VGS_ESTATE, "delta before snapshot upload");
    if (n < 0 || (n > 0 && !deltas)) return fail(c, VGS_EINVAL, "bad delta list");
    c->scored = c->resolved = false;
    RawPlanesMut rm;
    for (int f = 0; f < VGS_NUM_PLANES; f++) rm.p[f] = c->d_raw[f];
    rm.ndev = c->d_ndev;
    for (int32_t done = 0; done < n;) {
        int32_t m = n - done < c->max_deltas ? n - done : c->max_deltas;
        if (c->delta_in_flight) { // staging buffer reuse: only the previous delta copy has to be through
            CU(cudaEventSynchronize(c->ev_delta));
            c->delta_in_flight = false;
        }
        memcpy(c->h_deltas, deltas + done, (size_t)m * sizeof(vgs_delta));
        CU(cudaMemcpyAsync(c->d_deltas, c->h_deltas, (size_t)m * sizeof(vgs_delta),
                           cudaMemcpyHostToDevice, c->st));
        CU(launch_delta(rm, c->n_nodes, c->dmax, c->cfg.row_base, c->d_deltas, m, c->d_rowlist,
                        c->d_status, c->st));
        CU(launch_prepare(c->dm, raw_const(c), c->n_nodes, c->dmax, c->d_walk, c->n_tiles, c->d_status,
                          c->d_rowlist, m, c->st));
        CU(cudaEventRecord(c->ev_delta, c->st));
        c->delta_in_flight = true;
        c->launches += 2;
        done += m;
    }
    // Nothing is waited for: a delta is a few microseconds of GPU work in front of the next score pass.  What
    // the kernels find wrong with it (a slot the node does not have, usage out of range) is in the status word
    // and comes out of the next call that reads it -- vgs_results_download / vgs_score_batch / vgs_sync.
    c->version++;
    return VGS_OK;
}

int vgs_snapshot_download(vgs_ctx *c, int32_t *const planes[VGS_NUM_PLANES], int32_t *ndev) {
    if (!c || !planes || !ndev) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    if (!c->have_snapshot) return fail(c, VGS_ESTATE, "no snapshot");
    const size_t cells = (size_t)c->n_nodes * c->dmax;
    if (cells) {
        for (int f = 0; f < VGS_NUM_PLANES; f++)
            CU(cudaMemcpyAsync(planes[f], c->d_raw[f], cells * sizeof(int32_t), cudaMemcpyDeviceToHost, c->st));
        CU(cudaMemcpyAsync(ndev, c->d_ndev, (size_t)c->n_nodes * sizeof(int32_t), cudaMemcpyDeviceToHost, c->st));
    }
    CU(cudaStreamSynchronize(c->st));
    return VGS_OK;
}

// ---- pod batch ---------------------------------------------------------------------

// a ContainerDeviceRequest as the kernels see it
static DevReq to_dev_req(const vgs_req &q) {
    DevReq d;
    d.nums = q.nums;
    d.memfix = q.memreq > 0 ? q.memreq : 0;                            // score.go:114-116
    d.pct = q.mempct;
    d.coresreq = q.coresreq;
    d.flags = 0;
    if (q.mempct != VGS_MEMPCT_UNSET && q.memreq == 0) d.flags |= R_USEPCT; // score.go:117
    if (q.flags & VGS_REQ_NUMA_BIND) d.flags |= R_NUMA;
    d.blockmask = M_NOFREE | M_PAD;
    if (q.flags & VGS_REQ_IDLE_SHARE) { // cambricon/device.go:97-99
        d.flags |= R_IDLE;
        d.blockmask |= M_BUSY370;
    }
    if (q.coresreq == 100) d.blockmask |= M_EXCL; // score.go:128
    if (q.coresreq == 0) d.blockmask |= M_FULL;   // score.go:132
    if (q.coresreq > 100) d.blockmask |= M_ALWAYS; // score.go:110-113: never fits
    d.mask_lo = (uint32_t)q.typemask;
    d.mask_hi = (uint32_t)(q.typemask >> 32);
    return d;
}

// lay the staging block out for a batch of n_pods / n_reqs; returns the bytes in use
static size_t stage_layout(vgs_ctx *c, size_t P, size_t R) {
    size_t o = 0;
    c->off_reqs = o, o = align_up(o + R * sizeof(DevReq), 256);
    c->off_pods = o, o = align_up(o + P * sizeof(vgs_pod), 256);
    c->off_pick = o, o = align_up(o + (R + 1) * sizeof(int32_t), 256);
    c->off_reqsA = o, o = align_up(o + P * sizeof(DevReq), 256);
    c->off_origA = o, o = align_up(o + P * sizeof(int32_t), 256);
    c->off_reqsB = o, o = align_up(o + P * sizeof(DevReq), 256);
    c->off_origB = o, o = align_up(o + P * sizeof(int32_t), 256);
    c->off_podsG = o, o = align_up(o + P * sizeof(GenPod), 256);
    c->off_reqsC = o, o = align_up(o + P * sizeof(DevReq), 256);
    c->off_origC = o, o = align_up(o + P * sizeof(int32_t), 256);
    c->off_genall = o, o = align_up(o + P * sizeof(GenPod), 256);
    return o;
}

// Build the bit tables of one pod-sliced class (see SliceTab) into pinned staging and queue the
// upload.  Returns false when the class does not qualify (then the per-pair kernel runs).
static bool build_slice_class(vgs_ctx *c, int k, const DevReq *req, const DevReq *dreq_class,
                              const int32_t *orig, int32_t n, cudaError_t *err, cudaStream_t sp) {
    *err = cudaSuccess;
    c->slice_on[k] = false;
    // >= slice_min pods: always.  From slice_small pods upward when the pass is big enough for the
    // class set-up (tables, K1p) to pay: measured 100k nodes x 256 pods 0.224 -> 0.176 ms through the
    // node-sliced kernel, but 100k x 64 and 10k x 256 lose to the per-pair kernel.
    // Small snapshots stay on the per-pair kernel whatever the class size.  Wall clock of the resident call, 4096 / 1024
    // pods (tools/small_probe.py): 64-512 nodes 400 / 117 us sliced against 152 / 104 us per pair, 4096 nodes 422 / 127
    // against 262 / 124, 8192 nodes 379 / 138 against 362 / 149.
    if (c->n_nodes > 0 && c->n_nodes < c->tune.slice_min_nodes) return false;
    const bool big = n >= c->tune.slice_min;
    const bool mid = n >= c->tune.slice_small && (int64_t)n * c->n_nodes >= c->tune.slice_small_pairs;
    if (n < 1 || !(big || mid)) return false;
    // sorted distinct thresholds (few in practice): percentages through a 101-bit map, the rest through a small hash
    // set (slots tagged with the call's generation, so nothing is cleared) and sorted once they are few -- sorting
    // all n values first cost 0.19 ms of host time per 4096-pod batch, on the critical path of every call
    static thread_local std::vector<int32_t> mv, pv, cv;
    constexpr uint32_t kSlots = 4096;
    static_assert(kSlots >= 2 * (kSliceMaxDistinct + 1), "hash set too small for the distinct values a class may have");
    static thread_local int32_t set_val[2][kSlots];
    static thread_local uint32_t set_gen[2][kSlots], gen = 0;
    if (++gen == 0) { // (wrapped)
        memset(set_gen, 0, sizeof set_gen);
        gen = 1;
    }
    mv.clear(), pv.clear(), cv.clear();
    auto add = [&](int w, std::vector<int32_t> &out, int32_t x) {
        uint32_t h = ((uint32_t)x * 2654435761u) >> 20;
        while (set_gen[w][h] == gen && set_val[w][h] != x) h = (h + 1) & (kSlots - 1);
        if (set_gen[w][h] != gen) {
            set_gen[w][h] = gen, set_val[w][h] = x;
            out.push_back(x);
        }
    };
    bool pct_seen[kSliceMaxPct] = {};
    for (int32_t i = 0; i < n; i++) {
        const DevReq &d = req[i];
        if (d.flags & R_USEPCT) {
            if (d.pct < 0 || d.pct > 100) return false; // monotone in pct only on [0,100]
            pct_seen[d.pct] = true;
        } else {
            add(0, mv, d.memfix);
        }
        add(1, cv, d.coresreq);
        if ((int)mv.size() > kSliceMaxDistinct || (int)cv.size() > kSliceMaxDistinct) return false;
    }
    std::sort(mv.begin(), mv.end());
    std::sort(cv.begin(), cv.end());
    for (int32_t q = 0; q < kSliceMaxPct; q++)
        if (pct_seen[q]) pv.push_back(q);
    const int32_t km = (int32_t)mv.size(), kp = (int32_t)pv.size(), kc = (int32_t)cv.size();
    const int32_t W = (n + 31) / 32, Wp = (int32_t)align_up((size_t)W, 32);
    uint32_t *h = c->h_slice[k];
    size_t o = 0;
    const size_t o_type = o; o += (size_t)VGS_MAX_TYPES * Wp;
    const size_t o_mem = o;  o += (size_t)(km + 1) * Wp;
    const size_t o_pct = o;  o += (size_t)(kp + 1) * Wp;
    const size_t o_core = o; o += (size_t)4 * (kc + 1) * Wp;
    const size_t o_memv = o; o += km;
    const size_t o_pctv = o; o += kp;
    const size_t o_corev = o; o += kc;
    const size_t o_orig = o; o += n;
    if (o > c->slice_words) return false;
    // host: only the small value arrays; the bit tables are filled on the GPU (K1t) from the
    // request records that are already on their way to the device
    int32_t *hi = reinterpret_cast<int32_t *>(h);
    std::copy(mv.begin(), mv.end(), hi + o_memv);
    std::copy(pv.begin(), pv.end(), hi + o_pctv);
    std::copy(cv.begin(), cv.end(), hi + o_corev);
    std::copy(orig, orig + n, hi + o_orig);
    (void)o_mem, (void)o_pct, (void)o_core;
    *err = cudaMemcpyAsync(c->d_slice[k] + o_memv, h + o_memv, (o - o_memv) * sizeof(uint32_t),
                           cudaMemcpyHostToDevice, sp);
    if (*err != cudaSuccess) return false;
    *err = cudaMemsetAsync(c->d_slice[k], 0, o_memv * sizeof(uint32_t), sp);
    if (*err != cudaSuccess) return false;
    const uint32_t *d = c->d_slice[k];
    SliceTab &t = c->tab[k];
    t.base = d + o_type;
    (void)o_mem, (void)o_pct, (void)o_core;
    t.memv = reinterpret_cast<const int32_t *>(d + o_memv);
    t.pctv = reinterpret_cast<const int32_t *>(d + o_pctv);
    t.corev = reinterpret_cast<const int32_t *>(d + o_corev);
    t.orig = reinterpret_cast<const int32_t *>(d + o_orig);
    t.km = km, t.kp = kp, t.kc = kc, t.n_pods = n, t.W = W, t.Wp = Wp;
    *err = launch_slice_tables(t, dreq_class, sp);
    if (*err != cudaSuccess) return false;
    c->launches += 2;
    c->slice_on[k] = true;
    return true;
}

// want_all: also list EVERY pod as a GenPod in batch order (the drain scores one pod at a time)
static int pods_upload_locked(vgs_ctx *c, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs,
                              const vgs_req *reqs, bool want_all = false) {
    if (n_pods < 0 || n_reqs < 0 || (n_pods > 0 && (!pods || !reqs)))
        return fail(c, VGS_EINVAL, "bad pod batch");
    if (n_pods > c->cfg.max_pods || n_reqs > c->cfg.max_reqs)
        return fail(c, VGS_ETOOBIG, "pod batch larger than vgs_init capacity");
    c->have_pods = false;
    c->scored = c->resolved = false;
    // the staging block may still be the source of an earlier async copy
    if (c->stage_in_flight) {
        CU(cudaStreamSynchronize(c->pods_on_side ? c->st_pods : c->st));
        c->stage_in_flight = false;
    }
    const size_t stage_used = stage_layout(c, (size_t)n_pods, (size_t)n_reqs);
    DevReq *hreq = reinterpret_cast<DevReq *>(c->h_stage + c->off_reqs);
    vgs_pod *hpod = reinterpret_cast<vgs_pod *>(c->h_stage + c->off_pods);
    int32_t *hpick = reinterpret_cast<int32_t *>(c->h_stage + c->off_pick);
    DevReq *hA = reinterpret_cast<DevReq *>(c->h_stage + c->off_reqsA);
    int32_t *hoA = reinterpret_cast<int32_t *>(c->h_stage + c->off_origA);
    DevReq *hB = reinterpret_cast<DevReq *>(c->h_stage + c->off_reqsB);
    int32_t *hoB = reinterpret_cast<int32_t *>(c->h_stage + c->off_origB);
    GenPod *hG = reinterpret_cast<GenPod *>(c->h_stage + c->off_podsG);
    DevReq *hC = reinterpret_cast<DevReq *>(c->h_stage + c->off_reqsC);
    int32_t *hoC = reinterpret_cast<int32_t *>(c->h_stage + c->off_origC);

    int64_t picks = 0;
    for (int32_t r = 0; r < n_reqs; r++) {
        const vgs_req &q = reqs[r];
        if (q.nums < 1) return fail(c, VGS_EINVAL, "request with Nums < 1 (k8sutil/pod.go:33 drops those)");
        hreq[r] = to_dev_req(q);
        hpick[r] = (int32_t)picks;
        picks += q.nums;
        if (picks > c->cfg.max_picks) return fail(c, VGS_ETOOBIG, "sum of nums exceeds max_picks");
    }
    hpick[n_reqs] = (int32_t)picks;

    int32_t nA = 0, nB = 0, nC = 0, nG = 0, nGL = 0;
    for (int32_t p = 0; p < n_pods; p++) {
        const vgs_pod &pd = pods[p];
        if (pd.n_reqs < 1 || pd.n_reqs > VGS_MAX_REQS_PER_POD || pd.req_begin < 0 ||
            (int64_t)pd.req_begin + pd.n_reqs > n_reqs)
            return fail(c, VGS_EINVAL,
                        "pod without requests or with a bad request range (Filter short-circuits "
                        "pods that request nothing, scheduler.go:357-370)");
        int32_t prev = -1;
        for (int32_t r = pd.req_begin; r < pd.req_begin + pd.n_reqs; r++) {
            if (reqs[r].ctr < prev || reqs[r].ctr < 0) return fail(c, VGS_EINVAL, "container index not monotone");
            if (reqs[r].ctr != prev) hreq[r].flags |= R_NEWCTR;
            prev = reqs[r].ctr;
        }
        hpod[p] = pd;
        GenPod *hall = reinterpret_cast<GenPod *>(c->h_stage + c->off_genall);
        auto as_gen = [&]() {
            GenPod g;
            g.req_begin = pd.req_begin, g.n_reqs = pd.n_reqs, g.orig = p;
            g.flags = gen_pod_flags(&hreq[pd.req_begin], pd.n_reqs);
            if (c->tune.general_legacy) g.flags |= G_LEGACY;
            return g;
        };
        if (want_all) hall[p] = as_gen();
        const DevReq &first = hreq[pd.req_begin];
        // the bit tables of the sliced classes know nothing of M_BUSY370: idle-share requests
        // take the per-pair kernel, which tests the whole blockmask
        const bool idle = first.flags & R_IDLE;
        if (pd.n_reqs == 1 && first.nums == 1 && !idle) {
            hA[nA] = first; // numa-bind cannot matter for a single device
            hoA[nA++] = p;
        } else if (pd.n_reqs == 1 && !(first.flags & R_NUMA) && first.nums == 2 && !idle) {
            hB[nB] = first;
            hoB[nB++] = p;
        } else if (pd.n_reqs == 1 && (!(first.flags & R_NUMA) || first.nums == 1)) {
            hC[nC] = first;
            hoC[nC++] = p;
        } else { // walk-kernel pods from the front of the region, all-pairs pods from its end
            const GenPod gp = want_all ? hall[p] : as_gen();
            GenPod &g = (gp.flags & G_LEGACY) ? hG[n_pods - 1 - nGL++] : hG[nG - nGL];
            g = gp;
            nG++;
        }
    }
    // (a sliced class keeps all its pods: the launcher gives the words beyond a multiple of 32 to the
    // node-sliced kernel)
    c->n_pods = n_pods;
    c->n_reqs = n_reqs;
    c->nA = nA;
    c->nB = nB;
    c->nC = nC;
    c->nG = nG;
    c->nGL = nGL;
    // The walk kernel shares the work of a common first request between consecutive pods: order its class by
    // (pod flags, first request) and mark the runs.  Results are scattered by GenPod.orig, so the order is free.
    if (nG - nGL > 1) {
        GenPod *w = hG;
        const int32_t nw = nG - nGL;
        auto first_key = [&](const GenPod &g, int32_t (&k)[9]) {
            const DevReq &d = hreq[g.req_begin];
            k[0] = g.flags & ~(G_SAME_FIRST | G_SAVE_FIRST), k[1] = d.nums, k[2] = d.memfix, k[3] = d.pct, k[4] = d.coresreq, k[5] = (int32_t)d.blockmask;
            k[6] = (int32_t)d.flags, k[7] = (int32_t)d.mask_lo, k[8] = (int32_t)d.mask_hi;
        };
        std::sort(w, w + nw, [&](const GenPod &a, const GenPod &b) {
            int32_t ka[9], kb[9];
            first_key(a, ka), first_key(b, kb);
            for (int i = 0; i < 9; i++)
                if (ka[i] != kb[i]) return ka[i] < kb[i];
            return a.orig < b.orig;
        });
        for (int32_t i = 1; i < nw; i++) {
            int32_t ka[9], kb[9];
            first_key(w[i - 1], ka), first_key(w[i], kb);
            bool same = (w[i].flags & G_NO_USED_TRACK) && w[i].n_reqs > 1; // (the parked state has no usage flags)
            for (int q = 0; q < 9 && same; q++) same = ka[q] == kb[q];
            if (same) {
                w[i].flags |= G_SAME_FIRST;
                w[i - 1].flags |= G_SAVE_FIRST;
            }
        }
    }
    c->n_picks = (int32_t)picks;
    // Everything above was host work that overlaps an asynchronous snapshot upload.  So does the device side of
    // the batch: the request records and the bit tables of the sliced classes depend on the pods alone (what the
    // snapshot allows -- sliced kernels, type rows -- is looked at when the score pass is queued), so while a
    // snapshot copy is pending they go up on a stream of their own, behind what `st` held before the copy.
    cudaStream_t sp = c->st;
    c->pods_on_side = c->snap_pending;
    if (c->pods_on_side) {
        sp = c->st_pods;
        CU(cudaStreamWaitEvent(sp, c->ev_idle, 0));
        cudaEventRecord(c->ev[EV_PH0], sp);
        c->ev_alias[EV_PH0] = EV_PH0, c->ev_set[EV_PH0] = true;
    } else {
        rec(c, EV_PH0);
    }
    if (stage_used) {
        CU(cudaMemcpyAsync(c->d_stage, c->h_stage, stage_used, cudaMemcpyHostToDevice, sp));
        c->stage_in_flight = true;
    }
    if (!want_all) {
        cudaError_t e;
        build_slice_class(c, 0, hA, reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqsA), hoA, nA, &e, sp);
        CU(e);
        build_slice_class(c, 1, hB, reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqsB), hoB, nB, &e, sp);
        CU(e);
    } else { // the drain scores with the pair / walk evaluations: no bit tables (and the batch is not kept for a score pass)
        c->slice_on[0] = c->slice_on[1] = false;
    }
    if (c->pods_on_side) {
        cudaEventRecord(c->ev[EV_PH1], sp);
        c->ev_alias[EV_PH1] = EV_PH1, c->ev_set[EV_PH1] = true;
        CU(cudaEventRecord(c->ev_pods, sp));
        CU(cudaStreamWaitEvent(c->st, c->ev_pods, 0)); // whatever is queued on `st` next sees the batch
    } else {
        rec(c, EV_PH1);
    }
    c->have_pods = !want_all;
    return VGS_OK;
}

static int score_locked(vgs_ctx *c) {
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    if (!c->have_snapshot) return fail(c, VGS_ESTATE, "score before snapshot upload");
    if (!c->have_pods) return fail(c, VGS_ESTATE, "score before pod upload");
    const int32_t P = c->n_pods;
    rec(c, EV_SC0);
    // the candidate block is zero at rest: allocated zeroed, and K3 clears what it reads.  Only a pass
    // that was scored but never resolved leaves slots dirty.
    if (c->dirty_pods > 0) {
        const int32_t n = c->dirty_pods > P ? c->dirty_pods : P;
        if (n * 4 >= c->cfg.max_pods) { // keys and counts are one block: clear it in one go
            CU(cudaMemsetAsync(c->d_keys, 0, (size_t)c->cfg.max_pods * 12, c->st));
        } else {
            CU(cudaMemsetAsync(c->d_keys, 0, (size_t)n * sizeof(unsigned long long), c->st));
            CU(cudaMemsetAsync(c->d_counts, 0, (size_t)n * sizeof(int32_t), c->st));
        }
    }
    c->dirty_pods = P;
    ScoreArgs a;
    a.walk = c->d_walk;
    a.n_tiles = c->n_tiles;
    a.n_nodes = c->n_nodes;
    a.row_base = c->cfg.row_base;
    a.keys = c->d_keys;
    a.counts = c->d_counts;
    a.vfit = nullptr;
    a.vscore = nullptr;
    a.vwords = 0;
    if (c->cfg.verify && P > 0) {
        a.vfit = c->d_vfit;
        a.vscore = c->d_vscore;
        a.vwords = c->vwords;
        CU(cudaMemsetAsync(c->d_vfit, 0, (size_t)P * c->vwords * sizeof(uint32_t), c->st));
        if (c->n_nodes)
            CU(cudaMemsetAsync(c->d_vscore, 0, (size_t)P * c->n_nodes * sizeof(float), c->st));
    }
    const DevReq *dreq = reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqs);
    const bool sliced = c->snap_slice_ok && c->dm == 8;
    // fork: class A and class B chains on side streams, the rest on `st` -- only when there is
    // more than one chain worth overlapping (a single-pod Filter stays on one stream)
    const bool fork = (c->nA > 0) + (c->nB > 0) + (c->nC + c->nG > 0) > 1 && P >= 64;
    cudaStream_t sA = fork ? c->side[0] : c->st, sB = fork ? c->side[1] : c->st;
    if (fork) {
        CU(cudaEventRecord(c->ev_fork, c->st));
        for (int k = 0; k < 2; k++) CU(cudaStreamWaitEvent(c->side[k], c->ev_fork, 0));
    }
    // The Nums 2 chain is issued first: it is the longer one on the mixes measured (pair enumeration in its
    // scouts), and whichever chain is issued first gets the SMs first (cfg3, alternating runs on one box:
    // 0.4123 ms with class A first, 0.4083 ms with class B first).  vgs_tuning.class_order = 0 issues class A first.
    const int order_env = c->tune.class_order != 0 ? 1 : 0;
    {
        const int32_t df = c->plan.dict_full;
        memset(&c->plan, 0, sizeof c->plan);
        c->plan.dict_full = df;
        c->plan.cls[0].pods = c->nA, c->plan.cls[1].pods = c->nB, c->plan.cls[2].pods = c->nC, c->plan.cls[3].pods = c->nG;
    }
    // Node-sliced scouts pay for classes of up to ~160 pod words and only when every sliced class of the batch uses
    // them (measured at 100k-125k nodes: 2 x 64 words 0.402 -> 0.369 ms per step; 385 + 128 words 0.948 -> 0.989 with
    // both, 0.966 with the smaller class alone; one class of two on cfg3 0.43-0.47): a node-sliced word costs three
    // times a pod-sliced one, the cold start it avoids does not grow with the words.  Bit 4 lifts the limit.
    vgs_tuning tune_k = c->tune;
    if (!(tune_k.scouts_node_sliced & 16)) {
        const int32_t kScoutNsMaxWords = 160;
        const bool a_sl = c->nA && sliced && c->slice_on[0], b_sl = c->nB && sliced && c->slice_on[1];
        if ((a_sl && (c->nA + 31) / 32 > kScoutNsMaxWords) || (b_sl && (c->nB + 31) / 32 > kScoutNsMaxWords))
            tune_k.scouts_node_sliced &= ~3;
    }
    for (int turn = 0; turn < 2; turn++) {
    const bool doA = (turn == 0) != (order_env == 1);
    if (doA) {
    if (c->nA) {
        if (sliced && c->slice_on[0]) {
            c->plan.cls[0].kernel = VGS_KERNEL_SLICED;
            CU(launch_score_sliced(1, c->cfg.device, a, c->tab[0], c->n_types_used, c->d_status, sA, tune_k,
                                   &c->plan.cls[0], fork ? c->st : nullptr, c->ev_ns[0], c->ev_sc[0], fork ? c->side_sc[0] : nullptr));
            c->launches += c->plan.cls[0].launches;
        } else {
            c->plan.cls[0].kernel = VGS_KERNEL_PAIR, c->plan.cls[0].launches = c->n_tiles ? 1 : 0;
            CU(launch_score_single(c->dm, true, c->cfg.device, a,
                                   reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqsA),
                                   reinterpret_cast<const int32_t *>(c->d_stage + c->off_origA), c->nA, sA));
            if (c->n_tiles) c->launches++;
        }
    }
    } else {
    if (c->nB) {
        if (sliced && c->slice_on[1]) {
            c->plan.cls[1].kernel = VGS_KERNEL_SLICED;
            CU(launch_score_sliced(2, c->cfg.device, a, c->tab[1], c->n_types_used, c->d_status, sB, tune_k,
                                   &c->plan.cls[1], fork ? c->st : nullptr, c->ev_ns[1], c->ev_sc[1], fork ? c->side_sc[1] : nullptr));
            c->launches += c->plan.cls[1].launches;
        } else {
            c->plan.cls[1].kernel = VGS_KERNEL_PAIR, c->plan.cls[1].launches = c->n_tiles ? 1 : 0;
            CU(launch_score_single(c->dm, false, c->cfg.device, a,
                                   reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqsB),
                                   reinterpret_cast<const int32_t *>(c->d_stage + c->off_origB), c->nB, sB));
            if (c->n_tiles) c->launches++;
        }
    }
    }
    }
    if (c->nC) {
        CU(launch_score_single(c->dm, false, c->cfg.device, a,
                               reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqsC),
                               reinterpret_cast<const int32_t *>(c->d_stage + c->off_origC), c->nC, c->st));
        c->plan.cls[2].kernel = VGS_KERNEL_PAIR, c->plan.cls[2].launches = c->n_tiles ? 1 : 0;
        if (c->n_tiles) c->launches++;
    }
    if (c->nG) {
        const GenPod *gp = reinterpret_cast<const GenPod *>(c->d_stage + c->off_podsG);
        const int32_t n_walk = c->nG - c->nGL;
        CU(launch_score_walk(c->dm, c->cfg.device, a, dreq, gp, n_walk, c->st));
        CU(launch_score_general(c->dm, c->cfg.device, a, dreq, gp + c->n_pods - c->nGL, c->nGL, c->st));
        c->plan.cls[3].kernel = VGS_KERNEL_GENERAL;
        c->plan.cls[3].launches = c->n_tiles ? (n_walk > 0) + (c->nGL > 0) : 0;
        c->launches += c->plan.cls[3].launches;
    }
    for (int k = 0; k < 2 && fork; k++) { // join
        CU(cudaEventRecord(c->ev_join[k], c->side[k]));
        CU(cudaStreamWaitEvent(c->st, c->ev_join[k], 0));
    }
    rec(c, EV_SC1);
    c->scored = true;
    c->resolved = false;
    return VGS_OK;
}

static int resolve_locked(vgs_ctx *c) {
    if (!c->scored) return fail(c, VGS_ESTATE, "resolve before score");
    rec(c, EV_RS0);
    // a Filter-sized batch: K3 stores its records into the pinned host mirrors as well (the pinned
    // allocations are device-accessible under unified addressing), so the download is one status word
    c->direct_out = c->n_pods <= kDirectOutPods;
    CU(launch_resolve(raw_const(c), c->n_nodes, c->dmax, c->cfg.row_base,
                      reinterpret_cast<const vgs_pod *>(c->d_stage + c->off_pods),
                      reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqs),
                      reinterpret_cast<const int32_t *>(c->d_stage + c->off_pick), c->n_pods, c->d_keys,
                      c->d_counts, c->d_results, c->d_picks, c->d_status, c->st, nullptr,
                      c->direct_out ? c->h_results : nullptr, c->direct_out ? c->h_picks : nullptr, false,
                      c->direct_out ? c->h_status : nullptr));
    if (c->n_pods) c->launches++;
    rec(c, EV_RS1);
    c->dirty_pods = 0;
    c->resolved = true;
    return VGS_OK;
}

static int download_locked(vgs_ctx *c, vgs_result *results, vgs_pick *picks) {
    if (!c->resolved) return fail(c, VGS_ESTATE, "download before resolve");
    if (c->n_pods > 0 && !results) return fail(c, VGS_EINVAL, "null results");
    rec(c, EV_DH0);
    if (c->n_pods && !c->direct_out)
        CU(cudaMemcpyAsync(c->h_results, c->d_results, (size_t)c->n_pods * sizeof(vgs_result),
                           cudaMemcpyDeviceToHost, c->st));
    if (c->n_picks && picks && !c->direct_out)
        CU(cudaMemcpyAsync(c->h_picks, c->d_picks, (size_t)c->n_picks * sizeof(vgs_pick),
                           cudaMemcpyDeviceToHost, c->st));
    if (!(c->direct_out && c->n_pods > 0)) // (K3 of a Filter-sized batch forwarded the status word itself)
        CU(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st));
    rec(c, EV_DH1);
    CU(cudaStreamSynchronize(c->st));
    c->stage_in_flight = false;
    if (*c->h_status & S_DICTFULL) c->plan.dict_full = 1;
    int rc = check_status(c);
    if (rc != VGS_OK) {
        // a delta that went wrong leaves planes that may hold out-of-range values: upload again
        if (*c->h_status & (S_DELTA | S_RANGE | S_TYPE | S_NDEV)) c->have_snapshot = false;
        cudaMemsetAsync(c->d_status, 0, sizeof(uint32_t), c->st);
        return rc;
    }
    if (c->n_pods) memcpy(results, c->h_results, (size_t)c->n_pods * sizeof(vgs_result));
    if (c->n_picks && picks) memcpy(picks, c->h_picks, (size_t)c->n_picks * sizeof(vgs_pick));
    return VGS_OK;
}

int vgs_pods_upload(vgs_ctx *c, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return pods_upload_locked(c, n_pods, pods, n_reqs, reqs);
}

int vgs_score_resident(vgs_ctx *c) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return score_locked(c);
}

int vgs_resolve_resident(vgs_ctx *c) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return resolve_locked(c);
}

int vgs_results_download(vgs_ctx *c, vgs_result *results, vgs_pick *picks) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    return download_locked(c, results, picks);
}

// The Filter call proper (one pod, a handful of requests, snapshot resident): ONE kernel launch and one stream
// synchronisation -- the pod travels as launch parameters, the last CTA of the score kernel resolves the winner
// and writes the pinned host mirrors.  Returns 1 when the call is not of that shape (the phased path takes it).
static int filter_one_locked(vgs_ctx *c, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs,
                             vgs_result *results, vgs_pick *picks) {
    if (n_pods != 1 || !pods || !reqs || !results || n_reqs < 1 || n_reqs > kFilterMaxReqs || c->cfg.verify ||
        !c->tune.filter_one_launch || c->n_tiles <= 0)
        return 1;
    if (pods[0].req_begin != 0 || pods[0].n_reqs != n_reqs) return 1;
    FilterOne fo;
    memset(&fo, 0, sizeof fo);
    int32_t prev = -1, total = 0;
    for (int32_t r = 0; r < n_reqs; r++) {
        if (reqs[r].nums < 1 || reqs[r].ctr < prev || reqs[r].ctr < 0) return 1; // the phased path words the error
        fo.reqs[r] = to_dev_req(reqs[r]);
        if (reqs[r].ctr != prev) fo.reqs[r].flags |= R_NEWCTR;
        prev = reqs[r].ctr;
        fo.pick_begin[r] = total;
        total += reqs[r].nums;
        if (total > c->cfg.max_picks || total > kWalkMaxPicks) return 1;
    }
    if (total > 0 && !picks) return 1;
    fo.pick_begin[n_reqs] = total;
    fo.pod.req_begin = 0, fo.pod.n_reqs = n_reqs, fo.pod.orig = 0;
    fo.pod.flags = gen_pod_flags(fo.reqs, n_reqs);
    fo.raw = raw_const(c);
    fo.dmax = c->dmax;
    fo.results = c->d_results, fo.host_results = c->h_results;
    fo.picks = c->d_picks, fo.host_picks = c->h_picks;
    fo.status = c->d_status, fo.host_status = c->h_status;
    fo.done = c->d_done;
    c->have_pods = false; // the staging block does not hold this pod
    c->scored = c->resolved = false;
    if (c->dirty_pods > 0) {
        CU(cudaMemsetAsync(c->d_keys, 0, (size_t)c->cfg.max_pods * 12, c->st));
        c->dirty_pods = 0;
    }
    ScoreArgs a;
    a.walk = c->d_walk, a.n_tiles = c->n_tiles, a.n_nodes = c->n_nodes, a.row_base = c->cfg.row_base;
    a.keys = c->d_keys, a.counts = c->d_counts;
    a.vfit = nullptr, a.vscore = nullptr, a.vwords = 0;
    for (int i = 0; i < EV_N; i++) c->ev_set[i] = false;
    rec(c, EV_SC0);
    CU(launch_filter_one(c->dm, c->cfg.device, a, fo, c->st));
    rec(c, EV_SC1);
    c->launches++;
    {
        const int32_t df = c->plan.dict_full;
        memset(&c->plan, 0, sizeof c->plan);
        c->plan.dict_full = df;
        // the same classes as a batch: one request without a numa group to choose takes the per-pair kernel
        const DevReq &q = fo.reqs[0];
        const bool pair = n_reqs == 1 && !(fo.pod.flags & G_LEGACY) && (!(q.flags & R_NUMA) || q.nums == 1);
        const int k = !pair ? 3 : q.nums == 1 ? 0 : q.nums == 2 ? 1 : 2;
        c->plan.cls[k].pods = 1, c->plan.cls[k].launches = 1;
        c->plan.cls[k].kernel = pair ? VGS_KERNEL_PAIR : VGS_KERNEL_GENERAL;
    }
    CU(cudaStreamSynchronize(c->st));
    c->stage_in_flight = false;
    int rc = check_status(c);
    if (rc != VGS_OK) {
        if (*c->h_status & (S_DELTA | S_RANGE | S_TYPE | S_NDEV)) c->have_snapshot = false;
        cudaMemsetAsync(c->d_status, 0, sizeof(uint32_t), c->st);
        return rc;
    }
    results[0] = c->h_results[0];
    if (total) memcpy(picks, c->h_picks, (size_t)total * sizeof(vgs_pick));
    return VGS_OK;
}

int vgs_score_batch(vgs_ctx *c, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs,
                    vgs_result *results, vgs_pick *picks) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    int rc;
    if (n_pods == 1) {
        if ((rc = finish_snapshot(c)) != VGS_OK) return rc;
        if (c->have_snapshot && (rc = filter_one_locked(c, n_pods, pods, n_reqs, reqs, results, picks)) != 1) return rc;
    }
    c->chained = true;
    if ((rc = pods_upload_locked(c, n_pods, pods, n_reqs, reqs)) == VGS_OK && (rc = score_locked(c)) == VGS_OK &&
        (rc = resolve_locked(c)) == VGS_OK)
        rc = download_locked(c, results, picks);
    c->chained = false;
    return rc;
}

int vgs_drain_batch(vgs_ctx *c, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs,
                    vgs_result *results, vgs_pick *picks) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    {
        int frc = finish_snapshot(c);
        if (frc != VGS_OK) return frc;
    }
    if (!c->have_snapshot) return fail(c, VGS_ESTATE, "drain before snapshot upload");
    int rc;
    if ((rc = pods_upload_locked(c, n_pods, pods, n_reqs, reqs, true)) != VGS_OK) return rc;
    rec(c, EV_SC0);
    if (c->dirty_pods > 0) CU(cudaMemsetAsync(c->d_keys, 0, (size_t)c->cfg.max_pods * 12, c->st));
    c->dirty_pods = n_pods;
    ScoreArgs a;
    a.walk = c->d_walk, a.n_tiles = c->n_tiles, a.n_nodes = c->n_nodes, a.row_base = c->cfg.row_base;
    a.keys = c->d_keys, a.counts = c->d_counts;
    a.vfit = nullptr, a.vscore = nullptr, a.vwords = 0;
    const DevReq *dreq = reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqs);
    const vgs_pod *dpod = reinterpret_cast<const vgs_pod *>(c->d_stage + c->off_pods);
    const int32_t *dpick = reinterpret_cast<const int32_t *>(c->d_stage + c->off_pick);
    const GenPod *dall = reinterpret_cast<const GenPod *>(c->d_stage + c->off_genall);
    // one pod at a time, everything chained on the stream: score (the key/count slot of pod p is
    // addressed through GenPod.orig) -> resolve + commit into the raw planes + re-prepare the row (K3)
    const DevReq *hreq_all = reinterpret_cast<const DevReq *>(c->h_stage + c->off_reqs);
    const int32_t *hpick_all = reinterpret_cast<const int32_t *>(c->h_stage + c->off_pick);
    // batch-scale drain (vgs_drain.cu): chunks of pods, each scored against the whole snapshot once and then placed
    // in order from the rows the chunk has touched
    int chunk = c->tune.drain_chunk == 0 ? kDrainChunk : c->tune.drain_chunk;
    if (chunk > kDrainChunk) chunk = kDrainChunk;
    bool chunked = chunk > 0 && n_pods >= 32 && c->n_tiles > 0;
    for (int32_t p = 0; p < n_pods && chunked; p++)
        if (reinterpret_cast<const GenPod *>(c->h_stage + c->off_genall)[p].flags & G_LEGACY) chunked = false;
    c->plan.drain_chunks = 0, c->plan.drain_rescans = 0, c->plan.drain_chain_pods = 0;
    int32_t p_first = 0; // first pod of the per-pod chain (after chunks that gave up)
    if (chunked && !c->d_drain_ob[0]) { // 2 x 4 bytes per (pod of a chunk, node): 51 MB at 100k nodes
        const size_t max_tiles = ((size_t)c->cfg.max_nodes + kWarp - 1) / kWarp;
        for (int k = 0; k < 2 && chunked; k++)
            if (cudaMalloc(&c->d_drain_ob[k], (size_t)kDrainChunk * max_tiles * kWarp * sizeof(uint32_t)) != cudaSuccess) {
                cudaGetLastError(); // no room: the per-pod chain needs none
                cudaFree(c->d_drain_ob[0]);
                c->d_drain_ob[0] = c->d_drain_ob[1] = nullptr;
                chunked = false;
            }
    }
    if (chunked) {
        // Two streams.  st: chunk k is placed pod by pod (k_drain_commit); it hands the rows it changed to the next
        // chunk's commit kernel and to k_drain_apply in a carry block and leaves the walk copy alone.  side[0],
        // meanwhile: the carry block of chunk k-1 goes into the walk copy, then the pass of chunk k+1 runs over it
        // (score-bit matrix, candidate lists) -- one chunk stale, which the commit kernel of chunk k+1 makes up for
        // by starting from the rows chunk k changed.
        cudaStream_t ps = c->side[0];
        CU(cudaMemsetAsync(c->d_drain_stats, 0, 12 * sizeof(int32_t), c->st));
        CU(cudaMemsetAsync(c->d_seldone, 0, kDrainChunk * sizeof(unsigned int), c->st)); // (a drain that gave up may leave counts)
        // re-scans a chunk may spend before the drain gives the chunks up: about what its pods would cost pod by pod
        // (~11 us each) over what one re-scan costs (~0.064 us per tile: 0.2 ms at 100k nodes)
        int32_t max_rescans = c->n_tiles > 0 ? (int32_t)(11000 / c->n_tiles) : kDrainChunk;
        if (max_rescans > kDrainChunk) max_rescans = kDrainChunk;
        const int32_t *d_stop = c->d_drain_stats + 8;
        CU(cudaEventRecord(c->ev_fork, c->st));
        CU(cudaStreamWaitEvent(ps, c->ev_fork, 0));
        const int32_t n_chunks = (n_pods + chunk - 1) / chunk;
        auto pass_of = [&](int32_t k) {
            DrainPass dp;
            dp.walk = c->d_walk, dp.n_tiles = c->n_tiles, dp.row_base = c->cfg.row_base;
            dp.reqs = dreq, dp.pods = dall + (size_t)k * chunk;
            dp.n_pods = n_pods - k * chunk < chunk ? n_pods - k * chunk : chunk;
            dp.any_pair = dp.any_walk = false;
            for (int32_t p = k * chunk; p < k * chunk + dp.n_pods; p++) {
                const GenPod &gp = reinterpret_cast<const GenPod *>(c->h_stage + c->off_genall)[p];
                const bool pr = pair_pod(gp.n_reqs, hreq_all[gp.req_begin]);
                dp.any_pair |= pr, dp.any_walk |= !pr;
            }
            dp.ob = c->d_drain_ob[k & 1], dp.topk = c->d_topk[k & 1], dp.counts = c->d_dcounts[k & 1];
            dp.seglist = c->d_seglist, dp.segcount = c->d_segcount, dp.done = c->d_seldone; // (passes run one after another)
            dp.stop = c->d_drain_stats + 8;
            return dp;
        };
        CU(launch_drain_pass(c->dm, c->sms, pass_of(0), ps));
        CU(cudaEventRecord(c->ev_pass[0], ps));
        for (int32_t k = 0; k < n_chunks; k++) {
            if (k + 1 < n_chunks) { // side: chunk k-1 into the walk copy (pass k has read it), then pass k+1
                if (k > 0) {
                    CU(cudaStreamWaitEvent(ps, c->ev_commit[(k - 1) & 1], 0));
                    CU(launch_drain_apply(c->dm, c->d_carry[(k - 1) & 1], c->d_carry_words[(k - 1) & 1], c->d_walk,
                                          (k - 1) * chunk, d_stop, ps));
                }
                CU(launch_drain_pass(c->dm, c->sms, pass_of(k + 1), ps));
                CU(cudaEventRecord(c->ev_pass[(k + 1) & 1], ps));
            }
            CU(cudaStreamWaitEvent(c->st, c->ev_pass[k & 1], 0));
            const DrainPass dp = pass_of(k);
            CU(launch_drain_commit(c->dm, c->cfg.device, c->sms, raw_const(c), c->n_nodes, c->dmax, dp, dpick, k == 0,
                                   c->d_carry[(k + 1) & 1], c->d_carry_words[(k + 1) & 1], c->d_carry[k & 1],
                                   c->d_carry_words[k & 1], k * chunk, max_rescans, c->d_results + (size_t)k * chunk, c->d_picks, c->d_status,
                                   c->d_drain_stats, c->st));
            CU(cudaEventRecord(c->ev_commit[k & 1], c->st));
            c->launches += 4;
            c->plan.drain_chunks++;
        }
        // the last two chunks' rows (the side stream has applied the others)
        if (n_chunks > 1)
            CU(launch_drain_apply(c->dm, c->d_carry[(n_chunks - 2) & 1], c->d_carry_words[(n_chunks - 2) & 1], c->d_walk,
                                  (n_chunks - 2) * chunk, d_stop, c->st));
        CU(launch_drain_apply(c->dm, c->d_carry[(n_chunks - 1) & 1], c->d_carry_words[(n_chunks - 1) & 1], c->d_walk,
                              (n_chunks - 1) * chunk, d_stop, c->st));
        CU(cudaMemcpyAsync(c->h_drain_stats, c->d_drain_stats, 12 * sizeof(int32_t), cudaMemcpyDeviceToHost, c->st));
        // results come down with the stop word: did the chunks give up (too many re-scans)?  Then the pods from there
        // on take the per-pod chain below and the results are read again.
        rec(c, EV_SC1);
        c->scored = true;
        c->resolved = true;
        c->dirty_pods = 0;
        c->direct_out = false;
        rc = download_locked(c, results, picks);
        c->plan.drain_rescans = c->h_drain_stats[0];
#ifdef VGS_DRAIN_TIMING // (a build for kernel work only: per-phase clock counts of the commit kernel, in units of 64 cycles)
        fprintf(stderr, "drain cycles/64: A %d B1 %d B2 %d B %d C %d D %d E %d\n", c->h_drain_stats[1], c->h_drain_stats[6],
                c->h_drain_stats[7], c->h_drain_stats[2], c->h_drain_stats[3], c->h_drain_stats[4], c->h_drain_stats[5]);
#endif
        if (rc != VGS_OK) return rc;
        if (c->h_drain_stats[8] == 0) {
            c->version++;
            return VGS_OK;
        }
        p_first = c->h_drain_stats[8];
        c->plan.drain_chain_pods = n_pods - p_first;
    }
    for (int32_t p = p_first; p < n_pods; p++) {
        const GenPod &gp = reinterpret_cast<const GenPod *>(c->h_stage + c->off_genall)[p];
        if (gp.n_reqs <= kFilterMaxReqs && !(gp.flags & G_LEGACY) && c->n_tiles > 0 && c->tune.filter_one_launch) {
            // one launch per pod: score, last CTA resolves, commits into the raw planes and re-prepares the row
            FilterOne fo;
            memset(&fo, 0, sizeof fo);
            for (int32_t r = 0; r < gp.n_reqs; r++) {
                fo.reqs[r] = hreq_all[gp.req_begin + r];
                fo.pick_begin[r] = hpick_all[gp.req_begin + r];
            }
            fo.pick_begin[gp.n_reqs] = hpick_all[gp.req_begin + gp.n_reqs];
            fo.pod.req_begin = 0, fo.pod.n_reqs = gp.n_reqs, fo.pod.orig = 0, fo.pod.flags = gp.flags;
            fo.raw = raw_const(c);
            fo.dmax = c->dmax;
            // (pick_begin holds batch-wide offsets: the picks arrays are passed whole)
            fo.results = c->d_results + p, fo.host_results = c->h_results + p;
            fo.picks = c->d_picks, fo.host_picks = c->h_picks;
            fo.status = c->d_status, fo.host_status = c->h_status;
            fo.done = c->d_done;
            fo.commit_walk = c->d_walk;
            CU(launch_filter_one(c->dm, c->cfg.device, a, fo, c->st));
            c->launches++;
            continue;
        }
        if (gp.flags & G_LEGACY)
            CU(launch_score_general(c->dm, c->cfg.device, a, dreq, dall + p, 1, c->st));
        else
            CU(launch_score_walk(c->dm, c->cfg.device, a, dreq, dall + p, 1, c->st));
        CU(launch_resolve(raw_const(c), c->n_nodes, c->dmax, c->cfg.row_base, dpod + p, dreq, dpick, 1,
                          c->d_keys + p, c->d_counts + p, c->d_results + p, c->d_picks, c->d_status, c->st,
                          c->d_rowlist, nullptr, nullptr, false, nullptr, c->d_walk,
                          c->dm)); // resolve + commit into the raw planes + re-prepare the touched row
        c->launches += c->n_tiles ? 2 : 1;
    }
    rec(c, EV_SC1);
    c->scored = true;
    c->resolved = true;
    c->dirty_pods = 0; // every pod's K3 cleared its slot
    c->direct_out = false;
    c->version++;
    return download_locked(c, results, picks);
}

int vgs_sync(vgs_ctx *c) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    int frc = finish_snapshot(c);
    if (frc != VGS_OK) return frc;
    if (!c->have_snapshot) {
        CU(cudaStreamSynchronize(c->st));
        c->stage_in_flight = false;
        return VGS_OK;
    }
    int rc = pull_status(c); // synchronises the stream; reports what asynchronous deltas left in the status word
    if (rc != VGS_OK && (*c->h_status & (S_DELTA | S_RANGE | S_TYPE | S_NDEV))) c->have_snapshot = false;
    return rc;
}

int vgs_keys_device(vgs_ctx *c, uint64_t **dev_keys, int32_t **dev_counts) {
    if (!c) return VGS_EINVAL;
    if (dev_keys) *dev_keys = reinterpret_cast<uint64_t *>(c->d_keys);
    if (dev_counts) *dev_counts = c->d_counts;
    return VGS_OK;
}

int vgs_results_device(vgs_ctx *c, vgs_result **dev_results, vgs_pick **dev_picks) {
    if (!c) return VGS_EINVAL;
    if (dev_results) *dev_results = c->d_results;
    if (dev_picks) *dev_picks = c->d_picks;
    return VGS_OK;
}

int vgs_exchange_block(vgs_ctx *c, void **dev_block, uint64_t *bytes) {
    if (!c) return VGS_EINVAL;
    if (dev_block) *dev_block = c->d_keys;
    if (bytes) *bytes = (uint64_t)c->xblock_bytes;
    return VGS_OK;
}

int vgs_resolve_local(vgs_ctx *c) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (!c->scored) return fail(c, VGS_ESTATE, "resolve before score");
    rec(c, EV_RS0);
    c->direct_out = false;
    CU(launch_resolve(raw_const(c), c->n_nodes, c->dmax, c->cfg.row_base,
                      reinterpret_cast<const vgs_pod *>(c->d_stage + c->off_pods),
                      reinterpret_cast<const DevReq *>(c->d_stage + c->off_reqs),
                      reinterpret_cast<const int32_t *>(c->d_stage + c->off_pick), c->n_pods, c->d_keys,
                      c->d_counts, c->d_results, c->d_picks, c->d_status, c->st, nullptr, nullptr, nullptr, true));
    if (c->n_pods) c->launches++;
    rec(c, EV_RS1);
    return VGS_OK; // not "resolved" yet: the results are this shard's own until vgs_exchange_merge
}

int vgs_exchange_merge(vgs_ctx *c, const void *gathered, int32_t n_shards, uint64_t stride_bytes) {
    if (!c || !gathered || n_shards < 1) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (!c->scored) return fail(c, VGS_ESTATE, "merge before score");
    if (stride_bytes < c->xblock_bytes) return fail(c, VGS_EINVAL, "stride smaller than the exchange block");
    CU(launch_exchange_merge(gathered, n_shards, (size_t)stride_bytes, c->cfg.max_pods, c->picks_off,
                             reinterpret_cast<const vgs_pod *>(c->d_stage + c->off_pods),
                             reinterpret_cast<const int32_t *>(c->d_stage + c->off_pick), c->n_pods, c->d_keys,
                             c->d_counts, c->d_results, c->d_picks, c->st));
    if (c->n_pods) c->launches++;
    c->dirty_pods = 0;
    c->direct_out = false;
    c->resolved = true;
    return VGS_OK;
}

int vgs_keys_merge(vgs_ctx *c, const uint64_t *gk, const int32_t *gc, int32_t n_shards, int32_t stride,
                   int32_t count_stride) {
    if (!c || !gk || n_shards < 1 || stride < 0) return VGS_EINVAL;
    if (count_stride <= 0) count_stride = stride;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (!c->scored) return fail(c, VGS_ESTATE, "merge before score");
    if (stride < c->n_pods) return fail(c, VGS_EINVAL, "stride < n_pods");
    if (gc && count_stride < c->n_pods) return fail(c, VGS_EINVAL, "count_stride < n_pods");
    CU(launch_merge(reinterpret_cast<const unsigned long long *>(gk), gc, n_shards, stride, count_stride,
                    c->n_pods, c->d_keys, c->d_counts, c->st));
    if (c->n_pods) c->launches++;
    return VGS_OK;
}

int vgs_verify_download(vgs_ctx *c, uint32_t *fit_bitmap, float *scores) {
    if (!c) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    if (!c->cfg.verify) return fail(c, VGS_ESTATE, "context was not created with verify");
    if (!c->scored) return fail(c, VGS_ESTATE, "verify download before score");
    const size_t words = (size_t)(c->n_nodes + 31) / 32;
    if (fit_bitmap && words && c->n_pods)
        CU(cudaMemcpy2DAsync(fit_bitmap, words * sizeof(uint32_t), c->d_vfit,
                             (size_t)c->vwords * sizeof(uint32_t), words * sizeof(uint32_t), c->n_pods,
                             cudaMemcpyDeviceToHost, c->st));
    if (scores && c->n_nodes && c->n_pods)
        CU(cudaMemcpyAsync(scores, c->d_vscore, (size_t)c->n_pods * c->n_nodes * sizeof(float),
                           cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return VGS_OK;
}

void vgs_tuning_defaults(vgs_tuning *t) {
    if (!t) return;
    memset(t, 0, sizeof *t);
    t->slice_min = 512;
    t->slice_small = 64;
    t->slice_small_pairs = 16000000;
    t->node_words = -1;
    t->tables_in_smem = -1;
    t->tables_smem_kb = 48;
    t->scout_stride = 0;
    t->scout_split = 2;
    t->class_order = 1;
    t->zero_copy_upload = 1;
    t->general_legacy = 0;
    t->max_ctas = 0;
    t->filter_one_launch = 1;
    t->drain_chunk = 0;
    t->scouts_node_sliced = 11;
    t->slice_min_nodes = 6144;
}

int vgs_last_plan(vgs_ctx *c, vgs_plan *out) {
    if (!c || !out) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    *out = c->plan;
    return VGS_OK;
}

int vgs_last_timing(vgs_ctx *c, vgs_timing *out) {
    if (!c || !out) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->cfg.device);
    CU(cudaStreamSynchronize(c->st));
    memset(out, 0, sizeof *out);
    auto span = [&](int a, int b) -> float {
        float ms = 0.f;
        if (c->ev_set[a] && c->ev_set[b] &&
            cudaEventElapsedTime(&ms, c->ev[c->ev_alias[a]], c->ev[c->ev_alias[b]]) == cudaSuccess)
            return ms;
        return 0.f;
    };
    out->snapshot_h2d_ms = span(EV_UP0, EV_UP1);
    out->pods_h2d_ms = span(EV_PH0, EV_PH1);
    out->derive_ms = span(EV_UP1, EV_DR1);
    out->score_ms = span(EV_SC0, EV_SC1);
    out->resolve_ms = span(EV_RS0, EV_RS1);
    out->d2h_ms = span(EV_DH0, EV_DH1);
    out->kernel_launches = (int32_t)(c->launches & 0x7fffffff);
    return VGS_OK;
}

} // extern "C"
