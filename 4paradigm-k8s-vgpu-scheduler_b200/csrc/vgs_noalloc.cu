// vgs_noalloc.cu -- vgs_score_batch_noalloc: the scoring pass for callers that must not
// allocate GPU memory.
//
// Who that is.  The extender may run inside a container whose GPU memory is capped by the
// very product it schedules for (libvgpu.so intercepts cuMemAlloc and fails it once the
// pod's quota is used), or inside a process whose device memory belongs entirely to
// another allocator (a serving framework's pool).  In both cases a cudaMalloc made by this
// library at call time can fail or steal memory at an unpredictable moment, and cudaMalloc
// / cudaFree also synchronise the whole device, which a latency-sensitive caller cannot
// accept in the middle of its own GPU work.  vgs_init / vgs_snapshot_upload allocate; this
// entry point never does: the caller brings the snapshot planes and the output arrays in
// device memory it already owns, and everything the kernels need in between lives in one
// fixed array that is part of the library image.
//
// The price is a hard size limit (VGS_NOALLOC_MAX_*) and about 4 MiB of device memory held
// by every GPU that loads the library until the process exits.  With CUDA's default lazy
// module loading that memory (and the kernels' code) is committed by the FIRST call on a GPU,
// not when the library is loaded: a caller that runs under a memory quota should make one
// small warm-up call at start-up, while memory is still free, or run with
// CUDA_MODULE_LOADING=EAGER; after that first call no later call can fail for lack of memory.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <mutex>

#include "../../include/vgpuscore.h"
#include "vgs_kernels.h"
#include "vgs_layout.h"

using namespace vgs;

namespace {

constexpr size_t kA = 256; // alignment of every region
constexpr size_t up(size_t x) { return (x + kA - 1) / kA * kA; }

// regions of the scratch array, in bytes
constexpr size_t kOffStatus = 0;                                              // uint32
constexpr size_t kOffKeys = up(kOffStatus + sizeof(uint32_t));                // u64[MAX_PODS]
constexpr size_t kOffCounts = up(kOffKeys + 8 * VGS_NOALLOC_MAX_PODS);        // int32[MAX_PODS]
constexpr size_t kOffReqs = up(kOffCounts + 4 * VGS_NOALLOC_MAX_PODS);        // DevReq[MAX_REQS]
constexpr size_t kOffPods = up(kOffReqs + sizeof(DevReq) * VGS_NOALLOC_MAX_REQS);
constexpr size_t kOffGen = up(kOffPods + sizeof(vgs_pod) * VGS_NOALLOC_MAX_PODS);
constexpr size_t kOffPick = up(kOffGen + sizeof(GenPod) * VGS_NOALLOC_MAX_PODS);
constexpr size_t kOffWalk = up(kOffPick + 4 * (VGS_NOALLOC_MAX_REQS + 1));
constexpr size_t kWalkBytes = (size_t)(VGS_NOALLOC_MAX_NODES / kWarp) * kWalkWords * (8 / 4) * kWarp * sizeof(int4);
constexpr size_t kScratchBytes = kOffWalk + kWalkBytes;
static_assert(VGS_NOALLOC_MAX_NODES % kWarp == 0, "whole warp tiles");
static_assert(kScratchBytes < (size_t)5 << 20, "the fixed scratch is meant to stay modest");

// The scratch itself: part of the library image, so no allocation call ever creates it.
// In host code this name is only a handle for cudaGetSymbolAddress -- it is NOT a device
// address and must never be handed to a kernel or to cudaMemcpy as a pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

std::mutex g_lock;             // one call at a time per process: every call shares the array
unsigned char *g_base[64] = {}; // device address of g_noalloc_scratch, per GPU, filled at set-up

struct CurrentDevice {
    int prev = -1;
    bool ok;
    explicit CurrentDevice(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~CurrentDevice() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// same translation of a vgs_req as the allocating path performs when it uploads pods
DevReq to_dev_req(const vgs_req &q) {
    DevReq d;
    d.nums = q.nums;
    d.memfix = q.memreq > 0 ? q.memreq : 0;
    d.pct = q.mempct;
    d.coresreq = q.coresreq;
    d.flags = 0;
    if (q.mempct != VGS_MEMPCT_UNSET && q.memreq == 0) d.flags |= R_USEPCT;
    if (q.flags & VGS_REQ_NUMA_BIND) d.flags |= R_NUMA;
    d.blockmask = M_NOFREE | M_PAD;
    if (q.coresreq == 100) d.blockmask |= M_EXCL;
    if (q.coresreq == 0) d.blockmask |= M_FULL;
    if (q.coresreq > 100) d.blockmask |= M_ALWAYS;
    d.mask_lo = (uint32_t)q.typemask;
    d.mask_hi = (uint32_t)(q.typemask >> 32);
    return d;
}

} // namespace

extern "C" int vgs_score_batch_noalloc(int32_t device, void *stream, int32_t n_nodes, int32_t dmax,
                                       int32_t row_base, const int32_t *const dev_planes[VGS_NUM_PLANES],
                                       const int32_t *dev_ndev, int32_t n_pods, const vgs_pod *pods,
                                       int32_t n_reqs, const vgs_req *reqs, vgs_result *dev_results,
                                       vgs_pick *dev_picks) {
    // ---- arguments and the fixed limits, before any CUDA call ---------------------------
    if (device < 0 || device >= 64 || n_nodes < 0 || dmax < 1 || n_pods < 0 || n_reqs < 0) return VGS_EINVAL;
    if (!dev_planes || (n_nodes > 0 && !dev_ndev)) return VGS_EINVAL;
    if (n_pods > 0 && (!pods || !reqs || !dev_results)) return VGS_EINVAL;
    if (n_nodes > VGS_NOALLOC_MAX_NODES || n_pods > VGS_NOALLOC_MAX_PODS || n_reqs > VGS_NOALLOC_MAX_REQS ||
        dmax > VGS_NOALLOC_MAX_DEVICES)
        return VGS_ELIMIT;
    for (int f = 0; f < VGS_NUM_PLANES; f++)
        if (n_nodes > 0 && !dev_planes[f]) return VGS_EINVAL;

    // ---- request records, packed on the stack -------------------------------------------
    DevReq hreq[VGS_NOALLOC_MAX_REQS];
    vgs_pod hpod[VGS_NOALLOC_MAX_PODS];
    GenPod hgen[VGS_NOALLOC_MAX_PODS];
    int32_t hpick[VGS_NOALLOC_MAX_REQS + 1];
    int64_t total_picks = 0;
    for (int32_t r = 0; r < n_reqs; r++) {
        if (reqs[r].nums < 1) return VGS_EINVAL;
        if (reqs[r].flags & VGS_REQ_IDLE_SHARE) return VGS_EINVAL; // no type-flag input on this entry
        hreq[r] = to_dev_req(reqs[r]);
        hpick[r] = (int32_t)total_picks;
        total_picks += reqs[r].nums;
        if (total_picks > (int64_t)VGS_NOALLOC_MAX_REQS * VGS_MAX_DEVICES) return VGS_ELIMIT;
    }
    hpick[n_reqs] = (int32_t)total_picks;
    if (total_picks > 0 && !dev_picks) return VGS_EINVAL;
    for (int32_t p = 0; p < n_pods; p++) {
        const vgs_pod &pd = pods[p];
        if (pd.n_reqs < 1 || pd.n_reqs > VGS_MAX_REQS_PER_POD || pd.req_begin < 0 ||
            (int64_t)pd.req_begin + pd.n_reqs > n_reqs)
            return VGS_EINVAL;
        int32_t prev = -1;
        for (int32_t r = pd.req_begin; r < pd.req_begin + pd.n_reqs; r++) {
            if (reqs[r].ctr < prev || reqs[r].ctr < 0) return VGS_EINVAL;
            if (reqs[r].ctr != prev) hreq[r].flags |= R_NEWCTR;
            prev = reqs[r].ctr;
        }
        hpod[p] = pd;
        hgen[p].req_begin = pd.req_begin;
        hgen[p].n_reqs = pd.n_reqs;
        hgen[p].orig = p;
        hgen[p].flags = gen_pod_flags(&hreq[pd.req_begin], pd.n_reqs);
    }

    std::lock_guard<std::mutex> lk(g_lock);
    CurrentDevice cur(device);
    if (!cur.ok) return VGS_ECUDA;

    // ---- set-up, once per GPU, with that GPU current ------------------------------------
    // Ask the runtime where this GPU keeps its copy of the array.  (This first call may make
    // the runtime load the module; do it before any stream capture begins.)
    if (!g_base[device]) {
        unsigned char *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError(); // do not leave the error behind for the caller's next CUDA call
            return VGS_ECUDA;
        }
        g_base[device] = ptr;
    }
    unsigned char *base = g_base[device];
    cudaStream_t st = (cudaStream_t)stream;

    uint32_t *d_status = reinterpret_cast<uint32_t *>(base + kOffStatus);
    unsigned long long *d_keys = reinterpret_cast<unsigned long long *>(base + kOffKeys);
    int32_t *d_counts = reinterpret_cast<int32_t *>(base + kOffCounts);
    DevReq *d_reqs = reinterpret_cast<DevReq *>(base + kOffReqs);
    vgs_pod *d_pods = reinterpret_cast<vgs_pod *>(base + kOffPods);
    GenPod *d_gen = reinterpret_cast<GenPod *>(base + kOffGen);
    int32_t *d_pick = reinterpret_cast<int32_t *>(base + kOffPick);
    int4 *d_walk = reinterpret_cast<int4 *>(base + kOffWalk);

#define NA(call)                                                                                   \
    do {                                                                                           \
        if ((call) != cudaSuccess) {                                                               \
            cudaGetLastError();                                                                    \
            return VGS_ECUDA;                                                                      \
        }                                                                                          \
    } while (0)

    // The array still holds whatever the previous call left.  Status, candidate keys and fit
    // counts are accumulated with atomics, so they must start at zero; the walk tiles and the
    // request records are fully rewritten below for the sizes of THIS call.
    NA(cudaMemsetAsync(base, 0, kOffReqs, st));
    if (total_picks) NA(cudaMemsetAsync(dev_picks, 0, (size_t)total_picks * sizeof(vgs_pick), st));
    if (n_reqs) NA(cudaMemcpyAsync(d_reqs, hreq, (size_t)n_reqs * sizeof(DevReq), cudaMemcpyHostToDevice, st));
    NA(cudaMemcpyAsync(d_pick, hpick, (size_t)(n_reqs + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (n_pods) {
        NA(cudaMemcpyAsync(d_pods, hpod, (size_t)n_pods * sizeof(vgs_pod), cudaMemcpyHostToDevice, st));
        NA(cudaMemcpyAsync(d_gen, hgen, (size_t)n_pods * sizeof(GenPod), cudaMemcpyHostToDevice, st));
    }

    RawPlanes raw;
    for (int f = 0; f < VGS_NUM_PLANES; f++) raw.p[f] = dev_planes[f];
    raw.ndev = dev_ndev;
    raw.idle_types = 0;
    const int32_t n_tiles = (n_nodes + kWarp - 1) / kWarp;
    NA(launch_prepare(8, raw, n_nodes, dmax, d_walk, n_tiles, d_status, nullptr, 0, st));
    ScoreArgs a;
    a.walk = d_walk;
    a.n_tiles = n_tiles;
    a.n_nodes = n_nodes;
    a.row_base = row_base;
    a.keys = d_keys;
    a.counts = d_counts;
    a.vfit = nullptr;
    a.vscore = nullptr;
    a.vwords = 0;
    // (pods with more picks than the walk kernel's stamps hold cannot occur here: <= 1024 requests, but a
    // single pod may still carry > 100 picks -- those take the all-pairs kernel)
    bool any_legacy = false;
    for (int32_t p = 0; p < n_pods; p++) any_legacy = any_legacy || (hgen[p].flags & G_LEGACY);
    if (any_legacy) NA(launch_score_general(8, device, a, d_reqs, d_gen, n_pods, st));
    else NA(launch_score_walk(8, device, a, d_reqs, d_gen, n_pods, st));
    NA(launch_resolve(raw, n_nodes, dmax, row_base, d_pods, d_reqs, d_pick, n_pods, d_keys, d_counts, dev_results,
                      dev_picks, d_status, st));
    uint32_t hstatus = 0;
    NA(cudaMemcpyAsync(&hstatus, d_status, sizeof hstatus, cudaMemcpyDeviceToHost, st));
    NA(cudaStreamSynchronize(st));
#undef NA
    if (hstatus & S_NDEV) return VGS_EDMAX;
    if (hstatus & (S_TYPE | S_RANGE)) return VGS_ERANGE;
    if (hstatus & S_MISMATCH) return VGS_EINTERNAL;
    return VGS_OK;
}
