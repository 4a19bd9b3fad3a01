// vgs_group.cu -- vgs_group_*: the node set sharded over several GPUs of one box behind one handle
// (include/vgpuscore.h).  One call drives every GPU (the per-shard host work runs on one
// helper thread per further shard): one vgs_ctx per shard, each on a
// stream of its own, built only from the public single-shard entry points (pods upload / score / resolve_local
// / exchange block / exchange merge / download) plus the exchange itself:
//
//   NCCL  ncclAllGather of the shards' [keys | counts | picks] blocks inside one ncclGroupStart/End, one
//         communicator per GPU from ncclCommInitAll.  libnccl.so.2 is opened with dlopen at vgs_group_init, so
//         libvgpuscore.so itself does not depend on it.
//   PEER  every shard copies its block into the first GPU's gather buffer with a small kernel doing ordinary
//         16-byte stores through the peer mapping (cudaDeviceEnablePeerAccess); the first GPU's stream waits for
//         one event per shard and merges.  No collective launch; only the first GPU needs the gathered blocks,
//         because only it hands results to the host.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <thread>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/vgpuscore.h"

namespace {

// the few NCCL entry points used, resolved at run time (types as in nccl.h: opaque comm pointer, int enums)
typedef struct ncclComm *ncclComm_t;
struct Nccl {
    void *lib = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool load(std::string *err) {
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!lib) {
            *err = std::string("dlopen libnccl.so.2: ") + dlerror();
            return false;
        }
        auto sym = [&](const char *n) { return dlsym(lib, n); };
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        AllGather = reinterpret_cast<decltype(AllGather)>(sym("ncclAllGather"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        if (!CommInitAll || !CommDestroy || !AllGather || !GroupStart || !GroupEnd || !GetErrorString) {
            *err = "libnccl.so.2 lacks an expected symbol";
            return false;
        }
        return true;
    }
};
constexpr int kNcclInt8 = 0; // ncclInt8 / ncclChar

// block copy into (possibly another GPU's) memory: plain 16-byte loads and stores
__global__ void __launch_bounds__(256) k_push_block(const int4 *__restrict__ src, int4 *__restrict__ dst, int n16) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

double now_us() {
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Dev {
    int prev = -1;
    explicit Dev(int d) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == d) prev = -1;
        else cudaSetDevice(d);
    }
    ~Dev() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

} // namespace

// one helper thread per shard beyond the first: the per-shard host work of a pass (packing the request records
// into that context's staging block, queueing its copies and launches) runs side by side instead of G times in
// a row on the caller's thread
struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, stop = false, done = true;
    int rc = 0;
    void loop() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] { return has_job || stop; });
            if (stop) return;
            std::function<int()> j = std::move(job);
            has_job = false;
            lk.unlock();
            const int r = j();
            lk.lock();
            rc = r;
            done = true;
            cv.notify_all();
        }
    }
    void start() { th = std::thread([this] { loop(); }); }
    void submit(std::function<int()> j) {
        std::lock_guard<std::mutex> lk(mu);
        job = std::move(j);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return done; });
        return rc;
    }
    void shutdown() {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
            cv.notify_all();
        }
        if (th.joinable()) th.join();
    }
};

struct vgs_group {
    vgs_group_config cfg{};
    Worker worker[VGS_MAX_GROUP];
    bool workers_on = false;
    int G = 0;
    std::mutex mu;
    std::string err;
    vgs_ctx *ctx[VGS_MAX_GROUP] = {};
    cudaStream_t st[VGS_MAX_GROUP] = {};
    cudaEvent_t ev_done[VGS_MAX_GROUP] = {}; // shard g's block has arrived in the first GPU's buffer
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr; // first GPU: own resolve done / merge done
    int32_t lo[VGS_MAX_GROUP] = {}, hi[VGS_MAX_GROUP] = {};
    int32_t per = 0;             // rows per shard (capacity)
    void *block[VGS_MAX_GROUP] = {};
    uint64_t block_bytes = 0, stride = 0;
    unsigned char *gathered[VGS_MAX_GROUP] = {}; // NCCL: on every GPU; PEER: on the first GPU only
    bool peer_ok[VGS_MAX_GROUP] = {};            // shard g can store into the first GPU's memory directly
    Nccl nccl;
    ncclComm_t comm[VGS_MAX_GROUP] = {};
    bool have_comm = false;
    int32_t n_nodes = 0;
    double total_us = 0, exchange_us = 0;
};

namespace {

int gfail(vgs_group *g, int code, const std::string &what) {
    g->err = what;
    return code;
}
int gfail_ctx(vgs_group *g, int rc, int shard, const char *what) {
    char buf[600];
    snprintf(buf, sizeof buf, "%s (shard %d): %s: %s", what, shard, vgs_strerror(rc), vgs_last_error(g->ctx[shard]));
    g->err = buf;
    return rc;
}

void group_free(vgs_group *g) {
    if (g->workers_on)
        for (int i = 1; i < g->G; i++) g->worker[i].shutdown();
    for (int i = 0; i < g->G; i++) {
        Dev d(g->cfg.devices[i]);
        if (g->have_comm && g->comm[i]) g->nccl.CommDestroy(g->comm[i]);
        if (g->ctx[i]) vgs_destroy(g->ctx[i]);
        if (g->gathered[i]) cudaFree(g->gathered[i]);
        if (g->ev_done[i]) cudaEventDestroy(g->ev_done[i]);
        if (g->st[i]) cudaStreamDestroy(g->st[i]);
    }
    if (g->ev_t0) cudaEventDestroy(g->ev_t0);
    if (g->ev_t1) cudaEventDestroy(g->ev_t1);
    if (g->nccl.lib) dlclose(g->nccl.lib);
}

} // namespace

extern "C" {

int vgs_group_init(const vgs_group_config *cfg, vgs_group **out) {
    if (!cfg || !out) return VGS_EINVAL;
    *out = nullptr;
    if (cfg->n_devices < 1 || cfg->n_devices > VGS_MAX_GROUP || cfg->max_nodes < 1) return VGS_EINVAL;
    if (cfg->exchange != VGS_EXCHANGE_NCCL && cfg->exchange != VGS_EXCHANGE_PEER) return VGS_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return VGS_ECUDA;
    for (int i = 0; i < cfg->n_devices; i++)
        if (cfg->devices[i] < 0 || cfg->devices[i] >= ndev) return VGS_EINVAL;
    vgs_group *g = new (std::nothrow) vgs_group();
    if (!g) return VGS_EINVAL;
    g->cfg = *cfg;
    g->cfg.tuning = nullptr;
    g->G = cfg->n_devices;
    g->per = (cfg->max_nodes + g->G - 1) / g->G;
    auto bail = [&](int rc) {
        group_free(g);
        delete g;
        return rc;
    };
    int32_t max_pods = cfg->max_pods + (cfg->max_pods & 1); // the 12-byte-per-pod block travels in 8-byte units
    for (int i = 0; i < g->G; i++) {
        Dev d(cfg->devices[i]);
        if (cudaStreamCreateWithFlags(&g->st[i], cudaStreamNonBlocking) != cudaSuccess) return bail(VGS_ECUDA);
        if (cudaEventCreateWithFlags(&g->ev_done[i], cudaEventDisableTiming) != cudaSuccess) return bail(VGS_ECUDA);
        vgs_config c;
        memset(&c, 0, sizeof c);
        c.device = cfg->devices[i];
        c.max_nodes = g->per;
        c.max_devices = cfg->max_devices;
        c.max_pods = max_pods;
        c.max_reqs = cfg->max_reqs > max_pods ? cfg->max_reqs : max_pods;
        c.max_picks = cfg->max_picks > c.max_reqs ? cfg->max_picks : c.max_reqs;
        c.row_base = i * g->per;
        c.stream = g->st[i];
        c.tuning = cfg->tuning;
        int rc = vgs_init(&c, &g->ctx[i]);
        if (rc != VGS_OK) return bail(rc);
        uint64_t nb = 0;
        vgs_exchange_block(g->ctx[i], &g->block[i], &nb);
        g->block_bytes = nb;
    }
    g->stride = g->block_bytes; // a multiple of 16: blocks sit back to back in the gather buffer
    {
        Dev d(cfg->devices[0]);
        if (cudaEventCreate(&g->ev_t0) != cudaSuccess || cudaEventCreate(&g->ev_t1) != cudaSuccess) return bail(VGS_ECUDA);
    }
    if (cfg->exchange == VGS_EXCHANGE_NCCL) {
        for (int i = 0; i < g->G; i++)
            for (int j = 0; j < i; j++)
                if (cfg->devices[i] == cfg->devices[j]) return bail(VGS_EINVAL); // one NCCL rank per GPU
        if (g->G > 1) {
            std::string e;
            if (!g->nccl.load(&e)) return bail(VGS_ENCCL);
            int rc = g->nccl.CommInitAll(g->comm, g->G, cfg->devices);
            if (rc != 0) return bail(VGS_ENCCL);
            g->have_comm = true;
        }
        for (int i = 0; i < g->G; i++) {
            Dev d(cfg->devices[i]);
            if (cudaMalloc(&g->gathered[i], g->stride * g->G) != cudaSuccess) return bail(VGS_ECUDA);
        }
    } else {
        const int root = cfg->devices[0];
        {
            Dev d(root);
            if (cudaMalloc(&g->gathered[0], g->stride * g->G) != cudaSuccess) return bail(VGS_ECUDA);
        }
        for (int i = 0; i < g->G; i++) {
            const int dv = cfg->devices[i];
            if (dv == root) {
                g->peer_ok[i] = true;
                continue;
            }
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, dv, root) == cudaSuccess && can) {
                Dev d(dv);
                cudaError_t e = cudaDeviceEnablePeerAccess(root, 0);
                if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) g->peer_ok[i] = true;
                cudaGetLastError();
            } // else: cudaMemcpyPeerAsync (the driver stages it)
        }
    }
    for (int i = 1; i < g->G; i++) g->worker[i].start();
    g->workers_on = true;
    *out = g;
    return VGS_OK;
}

void vgs_group_destroy(vgs_group *g) {
    if (!g) return;
    for (int i = 0; i < g->G; i++) {
        Dev d(g->cfg.devices[i]);
        cudaStreamSynchronize(g->st[i]);
    }
    group_free(g);
    delete g;
}

const char *vgs_group_last_error(vgs_group *g) { return g ? g->err.c_str() : ""; }

int vgs_group_snapshot_type_flags(vgs_group *g, int32_t n_types, const uint32_t *flags) {
    if (!g) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(g->mu);
    for (int i = 0; i < g->G; i++) {
        int rc = vgs_snapshot_type_flags(g->ctx[i], n_types, flags);
        if (rc != VGS_OK) return gfail_ctx(g, rc, i, "type flags");
    }
    return VGS_OK;
}

int vgs_group_snapshot_upload(vgs_group *g, int32_t n_nodes, int32_t dmax, const int32_t *const planes[VGS_NUM_PLANES],
                              const int32_t *ndev) {
    if (!g || !planes || !ndev) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_nodes < 0 || n_nodes > g->cfg.max_nodes) return gfail(g, VGS_ETOOBIG, "snapshot larger than vgs_group_init capacity");
    if (dmax < 1) return gfail(g, VGS_EINVAL, "bad snapshot shape");
    // every copy is queued before any is waited for: the shards' uploads cross PCIe side by side
    for (int i = 0; i < g->G; i++) {
        const int32_t lo = std::min<int64_t>((int64_t)i * g->per, n_nodes), hi = std::min<int64_t>((int64_t)lo + g->per, n_nodes);
        g->lo[i] = lo, g->hi[i] = hi;
        const int32_t *p[VGS_NUM_PLANES];
        for (int f = 0; f < VGS_NUM_PLANES; f++) {
            if (!planes[f]) return gfail(g, VGS_EINVAL, "null plane");
            p[f] = planes[f] + (size_t)lo * dmax;
        }
        int rc = vgs_snapshot_upload_async(g->ctx[i], hi - lo, dmax, p, ndev + lo, nullptr);
        if (rc != VGS_OK) return gfail_ctx(g, rc, i, "snapshot upload");
    }
    for (int i = 0; i < g->G; i++) {
        int rc = vgs_sync(g->ctx[i]); // returns the shard's range errors
        if (rc != VGS_OK) return gfail_ctx(g, rc, i, "snapshot upload");
    }
    g->n_nodes = n_nodes;
    return VGS_OK;
}

int vgs_group_snapshot_delta(vgs_group *g, int32_t n, const vgs_delta *deltas) {
    if (!g) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(g->mu);
    for (int i = 0; i < g->G; i++) { // a shard ignores rows outside its range
        int rc = vgs_snapshot_delta(g->ctx[i], n, deltas);
        if (rc != VGS_OK) return gfail_ctx(g, rc, i, "snapshot delta");
    }
    return VGS_OK;
}

int vgs_group_score_batch(vgs_group *g, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs,
                          vgs_result *results, vgs_pick *picks) {
    if (!g) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(g->mu);
    const double t_begin = now_us();
    int rc;
    // 1. every shard: request records up, score kernels, own best node per pod -- queued by one helper thread per
    // shard (the caller's thread takes shard 0), nothing waited for on the GPU side
    auto shard_pass = [g, n_pods, pods, n_reqs, reqs](int i) -> int {
        int r;
        if ((r = vgs_pods_upload(g->ctx[i], n_pods, pods, n_reqs, reqs)) != VGS_OK) return r;
        if ((r = vgs_score_resident(g->ctx[i])) != VGS_OK) return r;
        return vgs_resolve_local(g->ctx[i]);
    };
    for (int i = 1; i < g->G; i++) g->worker[i].submit([shard_pass, i] { return shard_pass(i); });
    rc = shard_pass(0);
    int bad = rc != VGS_OK ? 0 : -1;
    for (int i = 1; i < g->G; i++) {
        const int r = g->worker[i].wait();
        if (r != VGS_OK && bad < 0) bad = i, rc = r;
    }
    if (bad >= 0) return gfail_ctx(g, rc, bad, "shard pass");
    {
        Dev d(g->cfg.devices[0]);
        cudaEventRecord(g->ev_t0, g->st[0]);
    }
    // 2. the one exchange of the pass
    if (g->cfg.exchange == VGS_EXCHANGE_NCCL && g->G > 1) {
        int nrc = g->nccl.GroupStart();
        for (int i = 0; i < g->G && nrc == 0; i++)
            nrc = g->nccl.AllGather(g->block[i], g->gathered[i], (size_t)g->stride, kNcclInt8, g->comm[i], g->st[i]);
        if (nrc == 0) nrc = g->nccl.GroupEnd();
        if (nrc != 0) return gfail(g, VGS_ENCCL, std::string("ncclAllGather: ") + g->nccl.GetErrorString(nrc));
        for (int i = 0; i < g->G; i++) // every rank merges: that also puts its own block back to rest
            if ((rc = vgs_exchange_merge(g->ctx[i], g->gathered[i], g->G, g->stride)) != VGS_OK)
                return gfail_ctx(g, rc, i, "merge");
    } else {
        // NCCL with one shard degenerates to the same thing: the block goes into the (local) gather buffer
        const int n16 = (int)(g->block_bytes / 16);
        for (int i = 0; i < g->G; i++) {
            Dev d(g->cfg.devices[i]);
            unsigned char *dst = g->gathered[0] + (size_t)i * g->stride;
            cudaError_t e = cudaSuccess;
            if (g->peer_ok[i] || g->cfg.exchange == VGS_EXCHANGE_NCCL) {
                k_push_block<<<(n16 + 255) / 256 < 64 ? (n16 + 255) / 256 : 64, 256, 0, g->st[i]>>>(
                    reinterpret_cast<const int4 *>(g->block[i]), reinterpret_cast<int4 *>(dst), n16);
                e = cudaGetLastError();
            } else {
                e = cudaMemcpyPeerAsync(dst, g->cfg.devices[0], g->block[i], g->cfg.devices[i], g->block_bytes, g->st[i]);
            }
            if (e == cudaSuccess) e = cudaEventRecord(g->ev_done[i], g->st[i]);
            if (e != cudaSuccess) return gfail(g, VGS_ECUDA, std::string("block push: ") + cudaGetErrorString(e));
        }
        {
            Dev d(g->cfg.devices[0]);
            for (int i = 1; i < g->G; i++) cudaStreamWaitEvent(g->st[0], g->ev_done[i], 0);
        }
        if ((rc = vgs_exchange_merge(g->ctx[0], g->gathered[0], g->G, g->stride)) != VGS_OK) return gfail_ctx(g, rc, 0, "merge");
    }
    {
        Dev d(g->cfg.devices[0]);
        cudaEventRecord(g->ev_t1, g->st[0]);
    }
    // 3. results from the first GPU
    if ((rc = vgs_results_download(g->ctx[0], results, picks)) != VGS_OK) return gfail_ctx(g, rc, 0, "download");
    // the other shards' status words (snapshot range / self-check) are looked at as well
    for (int i = 1; i < g->G; i++)
        if ((rc = vgs_sync(g->ctx[i])) != VGS_OK) return gfail_ctx(g, rc, i, "sync");
    g->total_us = now_us() - t_begin;
    float ms = 0.f;
    {
        Dev d(g->cfg.devices[0]);
        if (cudaEventElapsedTime(&ms, g->ev_t0, g->ev_t1) != cudaSuccess) ms = 0.f;
    }
    g->exchange_us = 1e3 * ms;
    return VGS_OK;
}

int vgs_group_last_timing(vgs_group *g, double *total_us, double *exchange_us) {
    if (!g) return VGS_EINVAL;
    std::lock_guard<std::mutex> lk(g->mu);
    if (total_us) *total_us = g->total_us;
    if (exchange_us) *exchange_us = g->exchange_us;
    return VGS_OK;
}

} // extern "C"
