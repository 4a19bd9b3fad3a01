/*
 * oracle/score_ref.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference scheduler extender's Filter scoring pass
 * (pkg/scheduler/score.go, the three vendors' CheckType in pkg/device/{nvidia,
 * hygon,cambricon}/device.go, scheduler.go:386-387).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (libvgpuscore.so)
 * never links or calls it.
 *
 * PARITY UNPINNED: the reference holds no test, golden vector or fixture for
 * pkg/scheduler, and no Go toolchain exists on this image, so this restatement
 * could not be checked against the reference's own outputs.  It is pinned only
 * by the hand-derived known-answer vectors of SURVEY.md section 8(c), plus
 * seven more derived the same way for the Cambricon vendor
 * (tests/golden/kat.json, tests/golden/make_kat.py).
 */
#ifndef VGPU_ORACLE_SCORE_REF_H
#define VGPU_ORACLE_SCORE_REF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* pkg/util/types.go:105-117 (DeviceUsage).  Id is replaced by the device's
 * registration slot, which is what the extender maps back to a UUID. */
typedef struct {
    int32_t Slot; /* stands in for Id / Index */
    int32_t Used;
    int32_t Count;
    int32_t Usedmem;
    int32_t Totalmem;
    int32_t Totalcore;
    int32_t Usedcores;
    int32_t Numa;
    const char *Type;
} ref_DeviceUsage;

/* pkg/scheduler/nodes.go:44-48 (DeviceUsageList / NodeUsage): a slice of
 * POINTERS; sort.Sort permutes the pointers in place. */
typedef struct {
    ref_DeviceUsage **Devices;
    int32_t len;
} ref_NodeUsage;

/* pkg/util/types.go:93-99 */
typedef struct {
    int32_t Nums;
    const char *Type;
    int32_t Memreq;
    int32_t MemPercentagereq;
    int32_t Coresreq;
} ref_ContainerDeviceRequest;

/* pkg/util/types.go:85-91; UUID replaced by Slot. */
typedef struct {
    int32_t Idx;
    int32_t Slot;
    int32_t Usedmem;
    int32_t Usedcores;
} ref_ContainerDevice;

/* The three pod annotations the NVIDIA policy reads
 * (pkg/device/nvidia/device.go:20-22); NULL = key absent. */
typedef struct {
    const char *use_gputype;
    const char *nouse_gputype;
    const char *numa_bind;
    /* Hygon vendor (pkg/device/hygon/device.go:20-21) */
    const char *use_dcutype;
    const char *nouse_dcutype;
    /* Cambricon vendor (pkg/device/cambricon/device.go:19-20) */
    const char *use_mlutype;
    const char *nouse_mlutype;
} ref_Annos;

/* One result record per pod / per pick: same shape the product's C ABI uses
 * (declared independently here so the oracle does not include product code). */
typedef struct {
    int32_t best_row; /* -1: no node fits */
    float score;
    int32_t fit_count;
    int32_t n_picks;
} ref_Result;

typedef struct {
    int32_t slot;
    int32_t usedmem;
    int32_t usedcores;
} ref_Pick;

/* ---- string policy (pkg/device/nvidia/device.go:62-112, score.go:71-84) -- */
int ref_strings_contains(const char *s, const char *sub);
/* strings.ToUpper (Go 1.19, Unicode 13.0.0) of s[0..n) into out (cap bytes, NUL-terminated); returns the whole length */
size_t ref_strings_to_upper(const char *s, size_t n, char *out, size_t cap);
int ref_parse_bool(const char *s, int *ok);
int ref_checkGPUtype(const ref_Annos *annos, const char *cardtype);
int ref_assertNuma(const ref_Annos *annos);
/* returns pass; *numa receives numaAssert */
int ref_checkType(const ref_Annos *annos, const ref_DeviceUsage *d,
                  const ref_ContainerDeviceRequest *n, int *numa);

uint64_t ref_typemask(const ref_Annos *annos, const char *req_type, int32_t memreq,
                      const char *const *type_names, int n_types, int *numa, int *idle);

/* ---- scoring (score.go:45-50, 86-214) ---------------------------------- */
void ref_sort_devices(ref_NodeUsage *node);
int ref_fitInCertainDevice(ref_NodeUsage *node, ref_ContainerDeviceRequest request,
                           const ref_Annos *annos, ref_ContainerDevice *tmpDevs,
                           int *n_tmp);
int ref_fitInDevices(ref_NodeUsage *node, const ref_ContainerDeviceRequest *requests,
                     int n_requests, const ref_Annos *annos, float *score,
                     ref_ContainerDevice *devs, int *n_devs);

/*
 * Faithful batch driver: for every pod, rebuild a fresh pointer-of-structs
 * snapshot from the packed planes (scheduler.go:259-278) and run calcScore
 * (score.go:183-214) over nodes in row order, then the canonical best-node
 * pick (highest score, ties -> highest row).
 *
 * Snapshot: node-major planes [n_nodes*dmax]; ndev[n_nodes]; type_id indexes
 * type_names.  Pods: pod p owns containers ctr_begin[p]..ctr_begin[p+1]-1,
 * container c owns requests req_begin[c]..req_begin[c+1]-1 (possibly none).
 *
 * Outputs (any may be NULL): results[n_pods]; picks = request-major, request r
 * owns entries [pick_begin[r], pick_begin[r]+Nums) with pick_begin the
 * exclusive prefix sum of max(Nums,0); fit[n_pods*n_nodes] (uint8) and
 * scores[n_pods*n_nodes] (float, 0 when not fit).
 * log_io != 0 additionally formats the reference's per-device log lines
 * (score.go:65-68,90,93,136,147,202,205) into log_fd (R1-io baseline).
 */
typedef struct {
    int32_t n_nodes, dmax;
    const int32_t *used, *count, *usedmem, *totalmem, *usedcores, *totalcore, *numa,
        *type_id;
    const int32_t *ndev;
    int32_t n_types;
    const char *const *type_names;
} ref_Snapshot;

typedef struct {
    int32_t n_pods;
    const int32_t *ctr_begin; /* n_pods+1 */
    const int32_t *req_begin; /* n_ctrs+1 */
    const int32_t *nums, *memreq, *mempct, *coresreq; /* n_reqs */
    const char *const *req_type;                     /* n_reqs */
    const ref_Annos *annos;                          /* n_pods */
} ref_PodBatch;

int ref_score_batch(const ref_Snapshot *snap, const ref_PodBatch *pods, ref_Result *results,
                    ref_Pick *picks, uint8_t *fit, float *scores, int log_io, int log_fd,
                    double *seconds_scoring, double *seconds_rebuild);

/*
 * "Best-effort CPU" (BASELINE.md R2): same arithmetic on the packed int32
 * planes with a per-request 64-bit type mask instead of strings, pods spread
 * over n_threads host threads.  Must agree with ref_score_batch bit for bit.
 */
typedef struct {
    int32_t n_pods;
    const int32_t *ctr_begin;
    const int32_t *req_begin;
    const int32_t *nums, *memreq, *mempct, *coresreq;
    const uint64_t *typemask; /* n_reqs */
    const uint8_t *numa_bind; /* n_reqs: bit 0 numaAssert, bit 1 "370 cards must be unused" */
} ref_PodBatchPacked;

int ref_score_batch_packed(const ref_Snapshot *snap, const ref_PodBatchPacked *pods,
                           ref_Result *results, ref_Pick *picks, uint8_t *fit, float *scores,
                           int n_threads, double *seconds);

/* result wire format: pkg/util/util.go:110-126 (EncodeContainerDevices /
 * EncodePodDevices).  ctr_n[c] picks per container; returns bytes written
 * (excluding NUL) or -1 if cap is too small. */
int ref_encode_pod_devices(int n_ctrs, const int32_t *ctr_n, const char *const *uuids,
                           const char *const *types, const int32_t *usedmem,
                           const int32_t *usedcores, char *out, int cap);

#ifdef __cplusplus
}
#endif
#endif
