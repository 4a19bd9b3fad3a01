/*
 * vgpuscore.h -- C ABI of libvgpuscore.so
 *
 * B200 (sm_100a) implementation of the one data-parallel hot path of the
 * 4paradigm k8s vGPU scheduler extender: the per-pod Filter scoring pass
 *
 *     Filter -> calcScore -> fitInDevices -> fitInCertainDevice + best-node pick
 *     (reference: pkg/scheduler/scheduler.go:377,386-387 and
 *      pkg/scheduler/score.go:45-50,71-214)
 *
 * The reference has no plugin interface at this seam (calcScore is package
 * private, sole caller scheduler.go:377); this header IS the seam.  Everything
 * is C99: plain pointers and sizes, caller-owned buffers, no callbacks, and the
 * library keeps no caller pointer after a call returns (cgo pointer rules).
 *
 * Conventions
 *   - every function returns 0 (VGS_OK) or a negative vgs_status;
 *     "no node fits" is NOT an error (best_row = -1), mirroring
 *     scheduler.go:381-385.  There is no CPU fallback: without a usable GPU
 *     vgs_init fails with VGS_ECUDA.
 *   - a context is bound to ONE GPU and holds ONE shard of the node set (rows
 *     [row_base, row_base + n_nodes)); multi-GPU = one context per GPU, the
 *     caller exchanges the per-pod candidate keys (see vgs_keys_* below).
 *   - calls on one context are serialised by an internal mutex.
 */
#ifndef VGPUSCORE_H
#define VGPUSCORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VGS_ABI_VERSION 2

/* Go 1.19 sort.Sort is a stable insertion sort only for n <= 12
 * (score.go:165); more devices per node would need pdqsort semantics. */
#define VGS_MAX_DEVICES 12
/* distinct device type strings (util.DeviceUsage.Type) per snapshot */
#define VGS_MAX_TYPES 64
/* ContainerDeviceRequest.MemPercentagereq value meaning "unset"
 * (pkg/device/nvidia/device.go:136) */
#define VGS_MEMPCT_UNSET 101
/* requests (containers x vendors) one pod may carry */
#define VGS_MAX_REQS_PER_POD 32
/* |Count| and |Used| must stay below this (ordering keys are packed) */
#define VGS_MAX_SHARE 1048576

/* plane order of vgs_snapshot_upload: fields of util.DeviceUsage
 * (pkg/util/types.go:105-117) that scoring reads, plus the type id. */
enum {
    VGS_PLANE_USED = 0,
    VGS_PLANE_COUNT = 1,
    VGS_PLANE_USEDMEM = 2,
    VGS_PLANE_TOTALMEM = 3,
    VGS_PLANE_USEDCORES = 4,
    VGS_PLANE_TOTALCORE = 5,
    VGS_PLANE_NUMA = 6,
    VGS_PLANE_TYPEID = 7, /* index into the caller's table of distinct Type strings */
    VGS_NUM_PLANES = 8
};

typedef enum {
    VGS_OK = 0,
    VGS_EINVAL = -1,  /* bad argument / malformed batch */
    VGS_ECUDA = -2,   /* CUDA runtime error (vgs_last_error has the text) */
    VGS_ETOOBIG = -3, /* exceeds the capacities given to vgs_init */
    VGS_EDMAX = -4,   /* more than VGS_MAX_DEVICES devices per node */
    VGS_ERANGE = -5,  /* snapshot value outside the supported range */
    VGS_ESTATE = -6,  /* call sequence error (e.g. score before upload) */
    VGS_EINTERNAL = -7, /* self-check failed (score kernel vs resolve kernel) */
    VGS_ELIMIT = -8 /* beyond the fixed VGS_NOALLOC_MAX_* sizes of vgs_score_batch_noalloc */
} vgs_status;

typedef struct vgs_ctx vgs_ctx;

/* Tuning of the score path (no counterpart in the reference).  Every field has a default that
 * vgs_tuning_defaults() writes; a context is configured once, at vgs_init, through
 * vgs_config.tuning (NULL = all defaults).  The library reads no environment variable.
 * None of these changes a result, only which kernel computes it. */
typedef struct {
    int32_t slice_min;        /* a single-request class (Nums 1 / Nums 2) runs on the pod-sliced kernels from this many pods (512) */
    int32_t slice_small;      /* ... or from this many when pods x nodes >= slice_small_pairs (64) */
    int64_t slice_small_pairs;/* (16000000) */
    int32_t node_words;       /* pod words (32 pods) of a sliced class handed to the node-sliced kernel:
                               * -1 = measured break-even (default), 0 = none, n = up to n words beyond a multiple of 32 */
    int32_t tables_in_smem;   /* bit tables of the pod-sliced kernel in shared memory: -1 = when they fit
                               * tables_smem_kb (default), 0 = never, 1 = always */
    int32_t tables_smem_kb;   /* (48) */
    int32_t scout_stride;     /* every scout_stride-th node tile is walked by the scout launch: 0 = auto (n_tiles / 96 in [4, 32]) */
    int32_t scout_split;      /* CTAs sharing one scout tile: 1, 2, 4 or 8 (2) */
    int32_t class_order;      /* 1 = issue the Nums 2 class chain first (default), 0 = the Nums 1 chain */
    int32_t zero_copy_upload; /* 1 = pinned host planes are pulled over PCIe by a copy kernel (default), 0 = cudaMemcpyAsync */
    int32_t general_legacy;   /* 1 = multi-request pods run on the round-1 all-pairs kernel (A/B and fallback testing), 0 = default */
    int32_t max_ctas;         /* upper bound on the CTAs of one pod-sliced score launch (sharing the GPU, tests): 0 = one wave */
    int32_t filter_one_launch;/* 1 = vgs_score_batch of ONE pod with <= 8 requests runs as a single kernel launch (the pod travels
                               * as launch parameters, the last CTA resolves the winner into the pinned result buffers);
                               * 0 = the phased path (copy, score kernel, resolve kernel) */
    int32_t drain_chunk;      /* vgs_drain_batch: pods per chunk of the batch-scale drain, 1..64 (0 = default 64); -1 = the
                               * per-pod chain (one pass over the whole snapshot per pod) */
    int32_t alloc_pad_kb;     /* layout experiments: this many KB between the request staging block and the bit tables inside the
                               * allocation that holds the score path's small arrays (0) */
    int32_t scouts_node_sliced; /* who walks the scout tiles of a pod-sliced class (the bests the main launch starts from):
                                 * bit 0 / bit 1 = the node-sliced kernel for the Nums 1 / Nums 2 class (else a cold
                                 * pod-sliced launch after sort / classify); bit 3 = on a stream of the class's own beside
                                 * sort / classify, bit 2 = in order on the class's stream (neither: on the stream of the
                                 * node-sliced words); a batch with a sliced class of more than 160 pod words keeps the
                                 * pod-sliced launches unless bit 4 is set.  Default 11 */
    int32_t slice_min_nodes;  /* ... and only on snapshots of at least this many nodes (6144): below, the classes' set-up (bit
                               * tables, preparation kernels) costs more than the per-pair kernel's whole pass */
} vgs_tuning;
void vgs_tuning_defaults(vgs_tuning *out);

typedef struct {
    int32_t device;      /* CUDA ordinal */
    int32_t max_nodes;   /* rows this context can hold */
    int32_t max_devices; /* device slots per node, 1..VGS_MAX_DEVICES */
    int32_t max_pods;    /* pods per batch */
    int32_t max_reqs;    /* requests per batch */
    int32_t max_picks;   /* sum of nums per batch */
    int32_t row_base;    /* global row index of local row 0 (node sharding) */
    int32_t verify;      /* !=0: also keep the full pod x node fit bitmap and scores */
    void *stream;        /* cudaStream_t to run on; NULL = library creates one */
    const vgs_tuning *tuning; /* NULL = defaults; copied by vgs_init */
} vgs_config;

/* One container x vendor request = util.ContainerDeviceRequest
 * (pkg/util/types.go:93-99) with the string policy pre-evaluated by the host:
 * bit t of typemask = checkType(annos, DeviceUsage{Type: types[t]}, req).pass
 * (score.go:71-84, nvidia/device.go:62-94,107-112). */
#define VGS_REQ_NUMA_BIND 1u /* checkType's numaAssert (nvidia/device.go:96-105) */
/* Cambricon request without memory (cambricon/device.go:97-99: Type "MLU", Memreq == 0): a
 * device whose type carries VGS_TYPE_IDLE_SHARE fails the type check while Used > 0 -- Used as
 * it stands when the request is examined, earlier requests of the same pod included.  The rule
 * of :94-96 (Memreq != 0 on a non-370 card) does not depend on usage and lives in typemask. */
#define VGS_REQ_IDLE_SHARE 2u
/* per device type (vgs_snapshot_type_flags): strings.Contains(d.Type, "370"), cambricon/device.go:94,97 */
#define VGS_TYPE_IDLE_SHARE 1u
typedef struct {
    int32_t ctr;      /* container index inside the pod, non-decreasing */
    int32_t nums;     /* Nums  (>= 1: k8sutil/pod.go:33 drops the rest) */
    int32_t memreq;   /* Memreq */
    int32_t mempct;   /* MemPercentagereq, VGS_MEMPCT_UNSET when unset */
    int32_t coresreq; /* Coresreq */
    uint32_t flags;   /* VGS_REQ_* */
    uint64_t typemask;
} vgs_req;

/* One pending pod = the [][]ContainerDeviceRequest of k8sutil.Resourcereqs
 * (pkg/k8sutil/pod.go:26-40), flattened.  Containers without a request simply
 * do not appear (score.go:195-198 gives them an empty device list). */
typedef struct {
    int32_t req_begin; /* first request in the reqs array */
    int32_t n_reqs;    /* 1..VGS_MAX_REQS_PER_POD */
} vgs_pod;

/* winner of one pod: NodeScore of the best node (score.go:29-33,
 * scheduler.go:386-387).  Ties: highest global row. */
typedef struct {
    int32_t best_row;  /* global row, -1 = no node fits */
    float score;       /* NodeScore.score */
    int32_t fit_count; /* len(*nodeScores): number of candidate nodes */
    int32_t n_picks;   /* sum of nums over the pod's requests, 0 when no fit */
} vgs_result;

/* One util.ContainerDevice (pkg/util/types.go:85-91) of the winner: slot is the
 * device's registration slot on that node (the caller maps it to the UUID). */
typedef struct {
    int32_t slot;
    int32_t usedmem;
    int32_t usedcores;
} vgs_pick;

/* usage change of one device: what addPod/delPod do to the next snapshot
 * (scheduler.go:280-297). */
typedef struct {
    int32_t row; /* global row */
    int32_t slot;
    int32_t d_used;
    int32_t d_usedmem;
    int32_t d_usedcores;
} vgs_delta;

/* CUDA-event time of the most recent run of each phase on the context's
 * stream (0 if the phase never ran). */
typedef struct {
    float snapshot_h2d_ms; /* vgs_snapshot_upload / _set_device copies */
    float derive_ms;       /* snapshot preparation kernel (K0) */
    float pods_h2d_ms;     /* request records host -> device */
    float score_ms;        /* score kernels (K1) incl. key reset */
    float resolve_ms;      /* resolve kernel (K3) incl. pick reset */
    float d2h_ms;          /* results device -> host */
    int32_t kernel_launches; /* kernels launched since vgs_init (cumulative) */
    int32_t reserved;
} vgs_timing;

/* ---- lifetime ----------------------------------------------------------- */
int vgs_abi_version(void);
const char *vgs_strerror(int status);
int vgs_init(const vgs_config *cfg, vgs_ctx **out);
void vgs_destroy(vgs_ctx *ctx);
const char *vgs_last_error(vgs_ctx *ctx);

/* ---- snapshot: replaces the per-Filter rebuild in getNodesUsage
 *      (scheduler.go:249-310) with a device-resident copy -------------------- */
/* planes[f] is node-major: planes[f][row * dmax + slot]; ndev[row] =
 * len(node.Devices).  Copies, prepares the walk-ordered planes on the GPU,
 * validates ranges.  Blocks until done. */
int vgs_snapshot_upload(vgs_ctx *ctx, int32_t n_nodes, int32_t dmax,
                        const int32_t *const planes[VGS_NUM_PLANES], const int32_t *ndev,
                        uint64_t *out_version);
/* Same, but returns as soon as the copy and the preparation are queued, so that the caller can go on
 * building the pod batch while the planes cross PCIe.  The planes must stay untouched until the next
 * call on this context that needs the snapshot (vgs_score_*, vgs_drain_batch, vgs_snapshot_delta /
 * _refresh / _download / _type_flags, vgs_sync): that call waits for the upload and returns its
 * range errors (VGS_EDMAX / VGS_ERANGE) in place of its own result.  vgs_pods_upload does NOT wait: the
 * request records and the bit tables depend on the pods alone, so a batch uploaded while the copy is
 * pending goes up on a second stream beside it.  No counterpart in the reference. */
int vgs_snapshot_upload_async(vgs_ctx *ctx, int32_t n_nodes, int32_t dmax,
                              const int32_t *const planes[VGS_NUM_PLANES], const int32_t *ndev,
                              uint64_t *out_version);
/* The same upload from planes in a NARROW wire format: plane f holds n_nodes * dmax elements of widths[f] bytes,
 * ndev holds n_nodes elements of widths[VGS_NUM_PLANES] bytes; 1, 2 or 4 bytes each, signed two's complement,
 * widened to int32 on the GPU by the kernel that reads them.  What the upload costs is the bytes that cross PCIe:
 * the packer that fills the planes from []*util.DeviceUsage (pkg/util/types.go:105-117) picks, per plane, the
 * narrowest type its values fit -- Used / Count / Usedcores / Totalcore / Numa / type id are bytes on any real
 * cluster (split counts, core percentages <= 100, < 64 types), memory stays int32: 14 instead of 32 bytes per
 * device.  asynchronous != 0: as vgs_snapshot_upload_async.  Pinned, 16-byte aligned planes are read straight
 * over PCIe; anything else is copied to a device staging block first (allocated at the first such call).
 * planes[f] == NULL (ndev == NULL) leaves that plane as it is on the GPU; it needs a resident snapshot of the same
 * n_nodes and dmax (VGS_ESTATE otherwise).  That is the upload for a Filter whose node registry has not changed:
 * Count / Totalmem / Totalcore / Numa / type come from the registration annotations (nodes.go:59-101), only Used /
 * Usedmem / Usedcores are recomputed from the pod list (scheduler.go:280-297) -- 6 bytes per device on the wire. */
int vgs_snapshot_upload_narrow(vgs_ctx *ctx, int32_t n_nodes, int32_t dmax, const void *const planes[VGS_NUM_PLANES],
                               const void *ndev, const uint8_t widths[VGS_NUM_PLANES + 1], int32_t asynchronous,
                               uint64_t *out_version);
/* same, from planes that already live in this GPU's memory */
int vgs_snapshot_set_device(vgs_ctx *ctx, int32_t n_nodes, int32_t dmax,
                            const int32_t *const dev_planes[VGS_NUM_PLANES],
                            const int32_t *dev_ndev, uint64_t *out_version);
/* flags[t] = VGS_TYPE_* of type id t (all zero until called; types >= n_types get 0).  The
 * preparation kernel folds them into the walk copy; with a snapshot resident it re-runs. */
int vgs_snapshot_type_flags(vgs_ctx *ctx, int32_t n_types, const uint32_t *flags);
/* re-run the preparation kernel over the resident planes (asynchronous; a range
 * error surfaces at the next vgs_results_download) */
int vgs_snapshot_refresh(vgs_ctx *ctx);
/* Queues the usage changes (scatter into the resident planes + re-preparation of the touched rows) and returns;
 * the deltas are copied before it does.  A delta naming a slot the node does not have (VGS_EINVAL) or driving
 * Used out of range (VGS_ERANGE) is reported by the next vgs_results_download / vgs_score_batch / vgs_sync on
 * this context, after which the snapshot has to be uploaded again.  Rows outside the shard are ignored. */
int vgs_snapshot_delta(vgs_ctx *ctx, int32_t n, const vgs_delta *deltas);
/* read the resident planes back (testing the delta path) */
int vgs_snapshot_download(vgs_ctx *ctx, int32_t *const planes[VGS_NUM_PLANES], int32_t *ndev);

/* ---- scoring: replaces calcScore (score.go:183-214) + the pick
 *      (scheduler.go:386-387) for a batch of pods against one snapshot -------- */
/* results[n_pods]; picks[sum of nums], request-major: request r owns
 * picks[B(r) .. B(r)+nums), B = exclusive prefix sum of nums over reqs[]. */
int vgs_score_batch(vgs_ctx *ctx, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs,
                    const vgs_req *reqs, vgs_result *results, vgs_pick *picks);

/* Sequential placement of a batch ("what-if drain"): pod p is scored against
 * the snapshot as the winners of pods 0..p-1 left it, exactly as P consecutive Filter calls
 * with addPod in between would (scheduler.go:372-394 per pod; usage folding :280-297).  The
 * whole chain runs on the GPU without host round trips; the resident snapshot keeps all the
 * commits afterwards.  Batches of 32 pods and more run in chunks (vgs_tuning.drain_chunk): the
 * chunk's pods are scored against the whole snapshot once, then placed in order by one CTA that
 * re-evaluates only the rows touched inside the chunk (a commit changes one row); smaller
 * batches, and pods the walk kernel does not take (more than 100 devices asked for), score the
 * whole snapshot per pod.  Single shard only.  Same outputs as vgs_score_batch.  The batch does not
 * stay resident for vgs_score_resident (upload it with vgs_pods_upload first). */
int vgs_drain_batch(vgs_ctx *ctx, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs,
                    const vgs_req *reqs, vgs_result *results, vgs_pick *picks);

/* the same pass in separate phases (resident benchmarking, node sharding) */
int vgs_pods_upload(vgs_ctx *ctx, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs,
                    const vgs_req *reqs);
int vgs_score_resident(vgs_ctx *ctx);   /* score kernels -> per-pod candidate keys */
int vgs_resolve_resident(vgs_ctx *ctx); /* winners owned by this shard -> results, picks */
int vgs_results_download(vgs_ctx *ctx, vgs_result *results, vgs_pick *picks);
int vgs_sync(vgs_ctx *ctx);

/* node sharding: per-pod candidate key = (ordered score bits << 32) | global
 * row, 0 = none; max over shards = the global winner under the canonical tie
 * rule.  keys_device returns this GPU's candidate block: max_pods uint64 keys
 * immediately followed by max_pods int32 fit counts (one allocation, so one
 * all-gather of max_pods * 12 bytes moves both).  keys_merge takes the
 * gathered blocks (device memory) and reduces them back into the context:
 * shard g's keys start at dev_gathered_keys + g * key_stride (uint64 units),
 * its counts at dev_gathered_counts + g * count_stride (int32 units;
 * <= 0 means "same as key_stride"). */
int vgs_keys_device(vgs_ctx *ctx, uint64_t **dev_keys, int32_t **dev_counts);
int vgs_keys_merge(vgs_ctx *ctx, const uint64_t *dev_gathered_keys,
                   const int32_t *dev_gathered_counts, int32_t n_shards, int32_t key_stride,
                   int32_t count_stride);
/* The same exchange in ONE collective.  The candidate block is followed, in the same allocation, by
 * the picks array ("exchange block": max_pods keys, max_pods counts, padding, max_picks picks).  Each
 * shard resolves its OWN best node per pod (vgs_resolve_local, which keeps the candidates), the blocks
 * are all-gathered once, and vgs_exchange_merge picks the winning shard per pod out of the gathered
 * blocks (shard g at dev_gathered + g * stride_bytes), sums the counts, writes results and copies the
 * winner's picks -- after which vgs_results_download returns the global answer on every shard. */
int vgs_exchange_block(vgs_ctx *ctx, void **dev_block, uint64_t *bytes);
int vgs_resolve_local(vgs_ctx *ctx);
int vgs_exchange_merge(vgs_ctx *ctx, const void *dev_gathered, int32_t n_shards, uint64_t stride_bytes);
/* device arrays behind vgs_results_download (shard combine of results/picks) */
int vgs_results_device(vgs_ctx *ctx, vgs_result **dev_results, vgs_pick **dev_picks);

/* ---- node sharding inside the library: several GPUs of one box behind one handle --------------------
 * (No counterpart in the reference.)  One process, one call, one context per GPU (the library keeps one helper
 * thread per GPU beyond the first for the per-shard host work of a pass); GPU g owns the
 * contiguous rows [g * ceil(N/G), (g+1) * ceil(N/G)) of the snapshot.  A pass scores
 * every shard with no collective, every shard resolves its own best node per pod, and the shards'
 * [keys | counts | picks] blocks (vgs_exchange_block) are combined by ONE exchange:
 *   VGS_EXCHANGE_NCCL  ncclAllGather over NVLink (libnccl.so.2 is opened at vgs_group_init; absent
 *                      library -> VGS_ENCCL).  One communicator per GPU (ncclCommInitAll).
 *   VGS_EXCHANGE_PEER  every shard copies its block into the first GPU's gather buffer with ordinary
 *                      peer-mapped stores (cudaDeviceEnablePeerAccess), ordered by events; no collective
 *                      launch.  Also the mode that works with several shards on ONE GPU (tests).
 * The first GPU merges (max key = max score, then highest global row: the single-GPU canonical winner) and
 * holds the results.  Same inputs and outputs as vgs_snapshot_upload / _delta / vgs_score_batch. */
#define VGS_ENCCL (-9) /* NCCL unavailable or a collective failed (vgs_group_last_error has the text) */
#define VGS_MAX_GROUP 8
enum { VGS_EXCHANGE_NCCL = 0, VGS_EXCHANGE_PEER = 1 };
typedef struct vgs_group vgs_group;
typedef struct {
    int32_t n_devices;                  /* 1..VGS_MAX_GROUP shards */
    int32_t devices[VGS_MAX_GROUP];     /* CUDA ordinal of each shard (repeats allowed with VGS_EXCHANGE_PEER) */
    int32_t max_nodes;                  /* rows of the WHOLE snapshot */
    int32_t max_devices, max_pods, max_reqs, max_picks; /* as in vgs_config */
    int32_t exchange;                   /* VGS_EXCHANGE_* */
    const vgs_tuning *tuning;           /* NULL = defaults */
} vgs_group_config;
int vgs_group_init(const vgs_group_config *cfg, vgs_group **out);
void vgs_group_destroy(vgs_group *g);
const char *vgs_group_last_error(vgs_group *g);
/* planes of the whole node set, as for vgs_snapshot_upload; the library hands every GPU its row range */
int vgs_group_snapshot_upload(vgs_group *g, int32_t n_nodes, int32_t dmax, const int32_t *const planes[VGS_NUM_PLANES],
                              const int32_t *ndev);
int vgs_group_snapshot_type_flags(vgs_group *g, int32_t n_types, const uint32_t *flags);
/* deltas name GLOBAL rows; every shard applies the ones it owns */
int vgs_group_snapshot_delta(vgs_group *g, int32_t n, const vgs_delta *deltas);
int vgs_group_score_batch(vgs_group *g, int32_t n_pods, const vgs_pod *pods, int32_t n_reqs, const vgs_req *reqs,
                          vgs_result *results, vgs_pick *picks);
/* wall-clock microseconds of the last vgs_group_score_batch and, of that, the part after the last shard's
 * score kernels were queued (exchange + merge + download) */
int vgs_group_last_timing(vgs_group *g, double *total_us, double *exchange_us);

/* ---- scoring without any GPU allocation --------------------------------------
 * (No counterpart in the reference: the Go extender never touches a GPU.)
 *
 * For callers that cannot let this library allocate device memory: an extender
 * running in a container whose GPU memory is capped by the vGPU product itself
 * (an allocation made at call time can be refused once the quota is spent), a
 * process whose device memory is owned by another allocator's pool, or a caller
 * that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply.
 * vgs_init and the context API allocate; this call never does (no cudaMalloc,
 * cudaFree, or allocating library routine).
 *
 * The caller supplies, in memory of GPU `device` that it already owns, the
 * snapshot planes (same layout as vgs_snapshot_upload) and the output arrays
 * dev_results[n_pods], dev_picks[sum of nums]; pods/reqs are host arrays as in
 * vgs_score_batch.  All intermediate data (prepared snapshot, request records,
 * candidate keys) lives in one fixed array inside the library image, ~4 MiB,
 * which every GPU that loads the library holds until the process exits.  Hence
 * the hard limits below; anything larger returns VGS_ELIMIT and touches nothing.
 *
 * The first call on each GPU looks up that array's device address (one-time
 * set-up); make that first call outside of any stream capture.  With CUDA's
 * default lazy module loading it is also the call that commits the array and the
 * kernels' code to device memory: make it at start-up (a one-node, one-pod call
 * will do) while memory is free, or set CUDA_MODULE_LOADING=EAGER.  Every call on a
 * GPU shares the array and finds the previous call's contents in it (the call
 * clears what has to start at zero).  Two calls executing at once on the same GPU
 * would corrupt each other, so the function takes a process-wide lock and does
 * not return until the GPU has finished and the outputs are complete.
 * stream: cudaStream_t to run on (NULL = the default stream).
 * Results are identical to vgs_score_batch on the same inputs. */
#define VGS_NOALLOC_MAX_NODES 16384
#define VGS_NOALLOC_MAX_DEVICES 8
#define VGS_NOALLOC_MAX_PODS 256
#define VGS_NOALLOC_MAX_REQS 1024
int vgs_score_batch_noalloc(int32_t device, void *stream, int32_t n_nodes, int32_t dmax,
                            int32_t row_base, const int32_t *const dev_planes[VGS_NUM_PLANES],
                            const int32_t *dev_ndev, int32_t n_pods, const vgs_pod *pods,
                            int32_t n_reqs, const vgs_req *reqs, vgs_result *dev_results,
                            vgs_pick *dev_picks);

/* verify mode (cfg.verify != 0): fit_bitmap[n_pods][ceil32(n_nodes)] words,
 * bit (row & 31) of word row/32; scores[n_pods][n_nodes], 0 when not fit. */
int vgs_verify_download(vgs_ctx *ctx, uint32_t *fit_bitmap, float *scores);

int vgs_last_timing(vgs_ctx *ctx, vgs_timing *out);

/* How the most recent vgs_score_resident / vgs_score_batch ran: which kernel family took each pod class
 * and with what launch geometry (what vgs_tuning steers).  Classes: 0 = one request, Nums 1; 1 = one
 * request, Nums 2; 2 = one request, Nums >= 3; 3 = everything else (several requests, numa-bind). */
enum { VGS_KERNEL_NONE = 0, VGS_KERNEL_PAIR = 1, VGS_KERNEL_SLICED = 2, VGS_KERNEL_GENERAL = 3 };
typedef struct {
    int32_t pods;             /* pods of the class in the batch */
    int32_t kernel;           /* VGS_KERNEL_* */
    int32_t words_pod_sliced; /* pod words (32 pods) on the pod-sliced kernel */
    int32_t words_node_sliced;/* pod words on the node-sliced kernel */
    int32_t tables_in_smem;
    int32_t scout_stride, scout_split;
    int32_t tiles_per_cta_scout, tiles_per_cta_main;
    int32_t launches;         /* kernels launched for the class */
} vgs_class_plan;
typedef struct {
    vgs_class_plan cls[4];
    int32_t dict_full;        /* a sliced class met more distinct candidate scores than its hash set holds since
                               * vgs_init: results stay exact (such scores are compared one by one), pruning is lost */
    int32_t drain_chunks;     /* the most recent vgs_drain_batch: chunks it ran as (0 = the per-pod chain) */
    int32_t drain_rescans;    /* ... and pods whose candidate list ran dry (a whole-snapshot scan inside the chunk) */
    int32_t drain_chain_pods; /* ... and pods placed by the per-pod chain after the chunks gave up (re-scans on a big snapshot
                               * cost more than that chain: near-identical pods on nodes that take one or two of them) */
} vgs_plan;
int vgs_last_plan(vgs_ctx *ctx, vgs_plan *out);

/* ---- host-side policy helpers (what the Go shim keeps in Go) ----------------
 * bit t = checkType pass for device type string types[t] under the pod's
 * annotations; NULL annotation = key absent.  Mirrors score.go:71-84 for the
 * NVIDIA vendor (nvidia/device.go:62-94,107-112; pass nvidia.com/use-gputype /
 * nouse-gputype) and the Hygon vendor (hygon/device.go:45-84; req_type "DCU",
 * pass hygon.com/use-dcutype / nouse-dcutype).  Other request types: mask 0. */
uint64_t vgs_host_typemask(const char *req_type, const char *use_gputype,
                           const char *nouse_gputype, const char *const *types, int32_t n_types);
/* strings.ToUpper as Go 1.19 computes it (what checkGPUtype / checkDCUtype / checkMLUtype fold both sides of their
 * substring test with, nvidia/device.go:66-86): ASCII byte by byte, anything else rune by rune through the simple
 * uppercase mapping of Unicode 13.0.0, bytes outside a valid UTF-8 encoding replaced by U+FFFD.  Writes at most
 * cap - 1 bytes + NUL into out (out may be NULL); returns the length of the whole result, -1 for a NULL string. */
int32_t vgs_host_to_upper(const char *s, char *out, int32_t cap);
/* the Cambricon vendor (cambricon/device.go:60-104; req_type "MLU", pass
 * cambricon.com/use-mlutype / nouse-mlutype): bit t = the card contains "MLU", survives
 * checkMLUtype, and is not ruled out by ":94-96 Memreq != 0 on a card that is not a 370".
 * *req_flags receives VGS_REQ_IDLE_SHARE when memreq == 0 (:97-99), else 0. */
uint64_t vgs_host_typemask_mlu(int32_t memreq, const char *use_mlutype, const char *nouse_mlutype,
                               const char *const *types, int32_t n_types, uint32_t *req_flags);
/* VGS_TYPE_* of one device type string */
uint32_t vgs_host_type_flags(const char *type_name);
/* nvidia/device.go:96-105 assertNuma */
int32_t vgs_host_numa_bind(const char *numa_bind_anno);
/* util.DecodeNodeDevices (pkg/util/util.go:68-99): the node registration
 * annotation "id,count,devmem,devcore,type,numa,health:" per device, decoded
 * straight into the columns the snapshot planes are filled from.  Id and Type
 * come back as (offset, length) into str.  Returns the device count,
 * VGS_EINVAL for what the reference rejects (no ':' at all, or an entry that
 * does not have exactly 7 fields), VGS_ETOOBIG beyond max_dev. */
int32_t vgs_host_decode_node_devices(const char *str, int32_t max_dev, int32_t *count, int32_t *devmem,
                                     int32_t *devcore, int32_t *numa, int32_t *health, int32_t *id_off,
                                     int32_t *id_len, int32_t *type_off, int32_t *type_len);
/* util.DecodePodDevices (pkg/util/util.go:128-172): the 4pd.io/vgpu-ids-new annotation back into
 * per-container device lists -- what the pod informer feeds addPod with (scheduler.go:72-92).
 * Containers are the ';' pieces; ctr_n[c] = devices of container c, the devices themselves
 * container-major in the flat columns; UUID and Type as (offset, length) into str; Usedmem /
 * Usedcores through strconv.ParseInt(s, 10, 32) with the error dropped (0 on a syntax error, the
 * int32 limit on overflow).  Returns the container count; 0 for the empty string AND for a
 * malformed entry (fewer than 4 fields), which the reference turns into an empty PodDevices with a
 * nil error (:167-169); VGS_ETOOBIG beyond max_ctrs / max_dev. */
int32_t vgs_host_decode_pod_devices(const char *str, int32_t max_ctrs, int32_t max_dev, int32_t *ctr_n,
                                    int32_t *uuid_off, int32_t *uuid_len, int32_t *type_off,
                                    int32_t *type_len, int32_t *usedmem, int32_t *usedcores);
/* util.EncodePodDevices (pkg/util/util.go:110-126): "UUID,Type,mem,cores:" per
 * pick, ';' between containers.  Returns length or VGS_ETOOBIG. */
int32_t vgs_host_encode_pod_devices(int32_t n_ctrs, const int32_t *picks_per_ctr,
                                    const char *const *uuids, const char *const *types,
                                    const vgs_pick *picks, char *out, int32_t cap);

#ifdef __cplusplus
}
#endif
#endif /* VGPUSCORE_H */
