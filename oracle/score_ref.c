/*
 * oracle/score_ref.c -- TEST INFRASTRUCTURE ONLY (see score_ref.h).
 *
 * Line-by-line CPU restatement of the reference's Filter scoring pass.  Every
 * function cites the reference lines it follows (paths relative to the
 * reference checkout).  PARITY UNPINNED: no reference test or golden vector
 * exists for pkg/scheduler and the reference (Go) cannot be built on this
 * image; pinned only by tests/golden/kat.json.
 *
 * Third-party behaviour restated here (not in the reference tree):
 *   Go 1.19 standard library (go.mod:3, docker/Dockerfile:1):
 *     sort.Sort for n <= 12  -> insertionSort (stable)        [score.go:165]
 *     strings.Contains / ToUpper (ASCII subset) / Split       [nvidia/device.go:65-89]
 *     strconv.ParseBool                                       [nvidia/device.go:99]
 *   Go int32 arithmetic wraps; float32 is IEEE binary32 RNE, no fusion in a/b+c.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (oracle/Makefile).
 */
#define _GNU_SOURCE
#include "score_ref.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#define REF_MAX_DEV 12 /* sort.Sort is insertion sort only up to 12 elements */

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* Go int32 arithmetic wraps on overflow. */
static inline int32_t i32_sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static inline int32_t i32_add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t i32_mul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }

/* ---- logging stand-in for fmt.Println / klog (R1-io baseline) ----------- */
static int g_log_on = 0;
static int g_log_fd = -1;
#define LOGF(...)                                                                                  \
    do {                                                                                           \
        if (g_log_on) dprintf(g_log_fd, __VA_ARGS__);                                              \
    } while (0)

/* ---- Go strings / strconv ------------------------------------------------ */

/* strings.Contains: true for the empty substring. */
int ref_strings_contains(const char *s, const char *sub) { return strstr(s, sub) != NULL; }

/* strings.ToUpper restricted to ASCII (real GPU model names are ASCII). */
static char *to_upper_dup(const char *s, size_t n) {
    char *r = (char *)malloc(n + 1);
    for (size_t i = 0; i < n; i++) {
        char c = s[i];
        r[i] = (c >= 'a' && c <= 'z') ? (char)(c - 32) : c;
    }
    r[n] = 0;
    return r;
}

/* strconv.ParseBool: 1,t,T,TRUE,true,True / 0,f,F,FALSE,false,False. */
int ref_parse_bool(const char *s, int *ok) {
    static const char *T[] = {"1", "t", "T", "TRUE", "true", "True"};
    static const char *F[] = {"0", "f", "F", "FALSE", "false", "False"};
    for (int i = 0; i < 6; i++)
        if (strcmp(s, T[i]) == 0) {
            *ok = 1;
            return 1;
        }
    for (int i = 0; i < 6; i++)
        if (strcmp(s, F[i]) == 0) {
            *ok = 1;
            return 0;
        }
    *ok = 0;
    return 0;
}

/* does ToUpper(cardtype) contain ToUpper(token[0..n))? */
static int upper_contains(const char *cardtype, const char *tok, size_t n) {
    char *a = to_upper_dup(cardtype, strlen(cardtype));
    char *b = to_upper_dup(tok, n);
    int r = strstr(a, b) != NULL;
    free(a);
    free(b);
    return r;
}

/* any comma token of list is a case-insensitive substring of cardtype.
 * Follows the two branches of nvidia/device.go:65-75 (no comma -> whole
 * string; comma -> strings.Split, empty tokens included). */
static int any_token_matches(const char *list, const char *cardtype) {
    if (strchr(list, ',') == NULL) return upper_contains(cardtype, list, strlen(list));
    const char *p = list;
    for (;;) {
        const char *q = strchr(p, ',');
        size_t n = q ? (size_t)(q - p) : strlen(p);
        if (upper_contains(cardtype, p, n)) return 1;
        if (!q) break;
        p = q + 1;
    }
    return 0;
}

/* pkg/device/nvidia/device.go:62-94 checkGPUtype */
int ref_checkGPUtype(const ref_Annos *annos, const char *cardtype) {
    if (annos->use_gputype) { /* :63-77  use wins; nouse is not consulted */
        return any_token_matches(annos->use_gputype, cardtype) ? 1 : 0;
    }
    if (annos->nouse_gputype) { /* :78-92 */
        return any_token_matches(annos->nouse_gputype, cardtype) ? 0 : 1;
    }
    return 1; /* :93 */
}

/* pkg/device/nvidia/device.go:96-105 assertNuma */
int ref_assertNuma(const ref_Annos *annos) {
    if (annos->numa_bind) {
        int ok;
        int enforce = ref_parse_bool(annos->numa_bind, &ok);
        if (ok && enforce) return 1;
    }
    return 0;
}

/* pkg/device/hygon/device.go:45-77 checkDCUtype: the same list logic on the
 * hygon.com/{use,nouse}-dcutype annotations */
static int ref_checkDCUtype(const ref_Annos *annos, const char *cardtype) {
    if (annos->use_dcutype) return any_token_matches(annos->use_dcutype, cardtype) ? 1 : 0;
    if (annos->nouse_dcutype) return any_token_matches(annos->nouse_dcutype, cardtype) ? 0 : 1;
    return 1;
}

/* pkg/device/cambricon/device.go:60-91 checkMLUtype: the same list logic on the
 * cambricon.com/{use,nouse}-mlutype annotations */
static int ref_checkMLUtype(const ref_Annos *annos, const char *cardtype) {
    if (annos->use_mlutype) return any_token_matches(annos->use_mlutype, cardtype) ? 1 : 0;
    if (annos->nouse_mlutype) return any_token_matches(annos->nouse_mlutype, cardtype) ? 0 : 1;
    return 1;
}

/* pkg/device/cambricon/device.go:93-104 CheckType (found iff n.Type contains "MLU";
 * numaAssert always false) */
static int ref_checkTypeMLU(const ref_Annos *annos, const ref_DeviceUsage *d,
                            const ref_ContainerDeviceRequest *n) {
    if (!ref_strings_contains(d->Type, "370") && n->Memreq != 0) return 0;               /* :94-96 */
    if (ref_strings_contains(d->Type, "370") && n->Memreq == 0 && d->Used > 0) return 0; /* :97-99 */
    return ref_checkMLUtype(annos, d->Type);                                             /* :100 */
}

/* pkg/scheduler/score.go:71-84 checkType over device.GetDevices(): NVIDIA
 * (nvidia/device.go:107-112), Hygon (hygon/device.go:79-84), Cambricon
 * (cambricon/device.go:93-104).  The three "found" conditions are mutually
 * exclusive, so the map order of GetDevices() does not matter.  A request type
 * no vendor claims is "Unrecognized device". */
int ref_checkType(const ref_Annos *annos, const ref_DeviceUsage *d,
                  const ref_ContainerDeviceRequest *n, int *numa) {
    *numa = 0;
    if (!ref_strings_contains(d->Type, n->Type)) return 0; /* :73-75 */
    if (strcmp(n->Type, "NVIDIA") == 0) {                  /* device.go:108 */
        *numa = ref_assertNuma(annos);
        return ref_checkGPUtype(annos, d->Type);
    }
    if (strcmp(n->Type, "DCU") == 0) return ref_checkDCUtype(annos, d->Type); /* hygon: numaAssert false */
    if (ref_strings_contains(n->Type, "MLU")) return ref_checkTypeMLU(annos, d, n);
    return 0; /* :82-83 */
}

/* type mask for the packed (R2) arm: bit t = checkType pass for an UNUSED device of
 * type_names[t] and a request of this type and Memreq; *numa = numaAssert; *idle = the
 * request additionally loses "370" cards that are in use (cambricon/device.go:97-99,
 * applied per device by the packed arm).  Built from the oracle's own string policy above. */
uint64_t ref_typemask(const ref_Annos *annos, const char *req_type, int32_t memreq,
                      const char *const *type_names, int n_types, int *numa, int *idle) {
    uint64_t m = 0;
    ref_ContainerDeviceRequest rq;
    memset(&rq, 0, sizeof rq);
    rq.Type = req_type;
    rq.Memreq = memreq;
    *numa = 0;
    if (idle) *idle = ref_strings_contains(req_type, "MLU") && memreq == 0;
    for (int t = 0; t < n_types && t < 64; t++) {
        ref_DeviceUsage d;
        memset(&d, 0, sizeof d);
        d.Type = type_names[t];
        int nm = 0;
        if (ref_checkType(annos, &d, &rq, &nm)) m |= (uint64_t)1 << t;
        if (nm) *numa = 1;
    }
    return m;
}

/* ---- ordering ------------------------------------------------------------ */

/* score.go:45-50 DeviceUsageList.Less */
static int dev_less(const ref_DeviceUsage *a, const ref_DeviceUsage *b) {
    if (a->Numa == b->Numa) return i32_sub(a->Count, a->Used) < i32_sub(b->Count, b->Used);
    return a->Numa < b->Numa;
}

/* sort.Sort(node.Devices) score.go:165 -- Go 1.19 sort.insertionSort:
 *   for i := a+1; i < b; i++ { for j := i; j > a && Less(j, j-1); j-- { Swap(j, j-1) } } */
void ref_sort_devices(ref_NodeUsage *node) {
    for (int i = 1; i < node->len; i++)
        for (int j = i; j > 0 && dev_less(node->Devices[j], node->Devices[j - 1]); j--) {
            ref_DeviceUsage *t = node->Devices[j];
            node->Devices[j] = node->Devices[j - 1];
            node->Devices[j - 1] = t;
        }
}

/* ---- score.go:86-152 fitInCertainDevice ---------------------------------- */
int ref_fitInCertainDevice(ref_NodeUsage *node, ref_ContainerDeviceRequest request,
                           const ref_Annos *annos, ref_ContainerDevice *tmpDevs, int *n_tmp) {
    ref_ContainerDeviceRequest k = request;
    int32_t originReq = k.Nums;
    int prevnuma = -1;
    int nt = 0;
    LOGF("Allocating device for container request {%d %s %d %d %d}\n", k.Nums, k.Type, k.Memreq,
         k.MemPercentagereq, k.Coresreq);
    for (int i = node->len - 1; i >= 0; i--) {
        ref_DeviceUsage *dev = node->Devices[i];
        LOGF("Scoring pod %d:%d:%d:%di%ddevice:%d\n", k.Memreq, k.MemPercentagereq, k.Coresreq,
             k.Nums, i, dev->Slot);
        int numa;
        int found = ref_checkType(annos, dev, &k, &numa); /* :94 */
        if (!found) {
            LOGF("card type mismatch,continueing... %s %s\n", dev->Type, k.Type);
            continue;
        }
        if (numa && prevnuma != dev->Numa) { /* :99-104 */
            LOGF("Numa not fit, resotoreing....k.nums= %d numa= %d : %d : %d\n", k.Nums, numa,
                 prevnuma, dev->Numa);
            k.Nums = originReq;
            prevnuma = dev->Numa;
            nt = 0;
        }
        int32_t memreq = 0;
        if (dev->Count <= dev->Used) continue; /* :107 */
        if (k.Coresreq > 100) {                /* :110-113 */
            *n_tmp = nt;
            return 0;
        }
        if (k.Memreq > 0) memreq = k.Memreq;              /* :114-116 */
        if (k.MemPercentagereq != 101 && k.Memreq == 0) { /* :117-120 int32 mul wraps */
            memreq = i32_mul(dev->Totalmem, k.MemPercentagereq) / 100;
        }
        if (i32_sub(dev->Totalmem, dev->Usedmem) < memreq) continue;        /* :121 */
        if (i32_sub(dev->Totalcore, dev->Usedcores) < k.Coresreq) continue; /* :124 */
        if (dev->Totalcore == 100 && k.Coresreq == 100 && dev->Used > 0) continue; /* :128 */
        if (dev->Totalcore != 0 && dev->Usedcores == dev->Totalcore && k.Coresreq == 0)
            continue;     /* :132 */
        if (k.Nums > 0) { /* :135-145 */
            LOGF("device %d first fitted\n", dev->Slot);
            k.Nums--;
            tmpDevs[nt].Idx = i;
            tmpDevs[nt].Slot = dev->Slot;
            tmpDevs[nt].Usedmem = memreq;
            tmpDevs[nt].Usedcores = k.Coresreq;
            nt++;
        }
        if (k.Nums == 0) { /* :146-149 */
            LOGF("device allocate success\n");
            *n_tmp = nt;
            return 1;
        }
    }
    *n_tmp = nt;
    return 0; /* :151 */
}

/* ---- score.go:154-181 fitInDevices --------------------------------------- */
int ref_fitInDevices(ref_NodeUsage *node, const ref_ContainerDeviceRequest *requests,
                     int n_requests, const ref_Annos *annos, float *score,
                     ref_ContainerDevice *devs, int *n_devs) {
    int32_t total = 0, free_ = 0;
    int sums = 0;
    int nd = 0;
    *score = 0.0f;
    for (int r = 0; r < n_requests; r++) {
        const ref_ContainerDeviceRequest *k = &requests[r];
        sums += (int)k->Nums;            /* :161 */
        if ((int)k->Nums > node->len) { /* :162-164 */
            *n_devs = nd;
            return 0;
        }
        ref_sort_devices(node); /* :165 */
        ref_ContainerDevice tmp[REF_MAX_DEV];
        int nt = 0;
        int fit = ref_fitInCertainDevice(node, *k, annos, tmp, &nt); /* :166 */
        if (fit) {
            for (int t = 0; t < nt; t++) { /* :168-174 */
                ref_DeviceUsage *d = node->Devices[tmp[t].Idx];
                total = i32_add(total, d->Count);
                free_ = i32_add(free_, i32_sub(d->Count, d->Used));
                d->Used = i32_add(d->Used, 1);
                d->Usedcores = i32_add(d->Usedcores, tmp[t].Usedcores);
                d->Usedmem = i32_add(d->Usedmem, tmp[t].Usedmem);
            }
            for (int t = 0; t < nt; t++) devs[nd++] = tmp[t]; /* :175 */
        } else {
            *n_devs = nd;
            return 0; /* :177 */
        }
    }
    /* :180  float32(total)/float32(free) + float32(len(node.Devices)-sums) */
    volatile float q = (float)total / (float)free_;
    *score = q + (float)(node->len - sums);
    *n_devs = nd;
    return 1;
}

/* ---- score.go:183-214 calcScore, one node -------------------------------- */
/* picks_out: request-major (see header); may be NULL. Returns 1 if the node is
 * a candidate (score.go:209), with its accumulated float32 score. */
static int calc_score_node(ref_NodeUsage *node, int row, const ref_PodBatch *pods, int p,
                           const int32_t *pick_begin, float *score_out, ref_Pick *picks_out) {
    float score = 0.0f; /* :187 */
    if (g_log_on) {     /* viewStatus :64-69,186 */
        LOGF("viewing status\n");
        for (int i = 0; i < node->len; i++) {
            ref_DeviceUsage *d = node->Devices[i];
            LOGF("&{%d %d %d %d %d %d %d %d %d %s true}\n", d->Slot, d->Slot, d->Used, d->Count,
                 d->Usedmem, d->Totalmem, d->Totalcore, d->Usedcores, d->Numa, d->Type);
        }
    }
    int n_ctrs = pods->ctr_begin[p + 1] - pods->ctr_begin[p];
    int appended = 0;
    for (int ci = 0; ci < n_ctrs; ci++) { /* :190 */
        int c = pods->ctr_begin[p] + ci;
        int r0 = pods->req_begin[c], r1 = pods->req_begin[c + 1];
        int sums = 0;
        for (int r = r0; r < r1; r++) sums += (int)pods->nums[r]; /* :191-194 */
        if (sums == 0) {                                          /* :195-198 */
            appended++;
            continue;
        }
        ref_ContainerDeviceRequest reqs[8];
        int nr = r1 - r0;
        if (nr > 8) return -1;
        for (int r = 0; r < nr; r++) {
            reqs[r].Nums = pods->nums[r0 + r];
            reqs[r].Type = pods->req_type[r0 + r];
            reqs[r].Memreq = pods->memreq[r0 + r];
            reqs[r].MemPercentagereq = pods->mempct[r0 + r];
            reqs[r].Coresreq = pods->coresreq[r0 + r];
        }
        ref_ContainerDevice devs[8 * REF_MAX_DEV];
        int nd = 0;
        float nodescore;
        int fit = ref_fitInDevices(node, reqs, nr, &pods->annos[p], &nodescore, devs, &nd); /* :199 */
        if (fit) {
            appended++; /* :201 */
            LOGF("node: %d  score: %g\n", row, (double)nodescore);
            score = score + nodescore; /* :203 float32 accumulate */
            if (picks_out) {
                /* fit => every request r took exactly Nums devices, in order */
                int t = 0;
                for (int r = 0; r < nr; r++)
                    for (int j = 0; j < reqs[r].Nums; j++, t++) {
                        ref_Pick *pk = &picks_out[pick_begin[r0 + r] + j];
                        pk->slot = devs[t].Slot;
                        pk->usedmem = devs[t].Usedmem;
                        pk->usedcores = devs[t].Usedcores;
                    }
            }
        } else {
            LOGF("node: %d  not fitted\n", row);
            break; /* :206 */
        }
    }
    *score_out = score;
    return appended == n_ctrs; /* :209 */
}

static void rebuild_node(const ref_Snapshot *s, int row, ref_DeviceUsage *pool,
                         ref_DeviceUsage **ptrs, ref_NodeUsage *node) {
    /* scheduler.go:259-278: one fresh DeviceUsage per registered device, in
     * registration order; usage already folded into the planes (:280-297). */
    int nd = s->ndev[row];
    for (int d = 0; d < nd; d++) {
        size_t i = (size_t)row * s->dmax + d;
        ref_DeviceUsage *u = &pool[d];
        u->Slot = d;
        u->Used = s->used[i];
        u->Count = s->count[i];
        u->Usedmem = s->usedmem[i];
        u->Totalmem = s->totalmem[i];
        u->Totalcore = s->totalcore[i];
        u->Usedcores = s->usedcores[i];
        u->Numa = s->numa[i];
        u->Type = s->type_names[s->type_id[i]];
        ptrs[d] = u;
    }
    node->Devices = ptrs;
    node->len = nd;
}

static int32_t *make_pick_begin(int n_reqs, const int32_t *nums) {
    int32_t *pb = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n_reqs + 1));
    int32_t acc = 0;
    for (int r = 0; r < n_reqs; r++) {
        pb[r] = acc;
        acc += nums[r] > 0 ? nums[r] : 0;
    }
    pb[n_reqs] = acc;
    return pb;
}

int ref_score_batch(const ref_Snapshot *snap, const ref_PodBatch *pods, ref_Result *results,
                    ref_Pick *picks, uint8_t *fit, float *scores, int log_io, int log_fd,
                    double *seconds_scoring, double *seconds_rebuild) {
    if (snap->dmax > REF_MAX_DEV || snap->dmax < 1) return -1;
    int N = snap->n_nodes, D = snap->dmax;
    int n_ctrs = pods->ctr_begin[pods->n_pods];
    int n_reqs = pods->req_begin[n_ctrs];
    int32_t *pick_begin = make_pick_begin(n_reqs, pods->nums);
    ref_DeviceUsage *pool = (ref_DeviceUsage *)malloc(sizeof(ref_DeviceUsage) * (size_t)N * D);
    ref_DeviceUsage **ptrs = (ref_DeviceUsage **)malloc(sizeof(void *) * (size_t)N * D);
    ref_NodeUsage *nodes = (ref_NodeUsage *)malloc(sizeof(ref_NodeUsage) * (size_t)N);
    ref_Pick *tmp_picks = (ref_Pick *)malloc(sizeof(ref_Pick) * (size_t)(pick_begin[n_reqs] + 1));
    double t_score = 0, t_build = 0;
    g_log_on = log_io;
    g_log_fd = log_fd;
    int rc = 0;
    for (int p = 0; p < pods->n_pods && rc == 0; p++) {
        double t0 = now_s();
        for (int row = 0; row < N; row++)
            rebuild_node(snap, row, pool + (size_t)row * D, ptrs + (size_t)row * D, &nodes[row]);
        double t1 = now_s();
        /* calcScore over nodes (map order in Go; row order here) + pick
         * scheduler.go:386-387 under the canonical tie rule: max score, then
         * highest row. */
        int best = -1, cnt = 0;
        float best_score = 0.0f;
        for (int row = 0; row < N; row++) {
            float sc;
            int ok = calc_score_node(&nodes[row], row, pods, p, pick_begin, &sc, tmp_picks);
            if (ok < 0) {
                rc = -1;
                break;
            }
            if (fit) fit[(size_t)p * N + row] = (uint8_t)ok;
            if (scores) scores[(size_t)p * N + row] = ok ? sc : 0.0f;
            if (ok) {
                cnt++;
                if (best < 0 || sc >= best_score) {
                    best = row;
                    best_score = sc;
                    if (picks) {
                        int c0 = pods->ctr_begin[p], c1 = pods->ctr_begin[p + 1];
                        int a = pick_begin[pods->req_begin[c0]], b = pick_begin[pods->req_begin[c1]];
                        memcpy(picks + a, tmp_picks + a, sizeof(ref_Pick) * (size_t)(b - a));
                    }
                }
            }
        }
        double t2 = now_s();
        t_build += t1 - t0;
        t_score += t2 - t1;
        if (results) {
            results[p].best_row = best;
            results[p].score = best >= 0 ? best_score : 0.0f;
            results[p].fit_count = cnt;
            int c0 = pods->ctr_begin[p], c1 = pods->ctr_begin[p + 1];
            results[p].n_picks =
                best >= 0 ? pick_begin[pods->req_begin[c1]] - pick_begin[pods->req_begin[c0]] : 0;
        }
    }
    g_log_on = 0;
    if (seconds_scoring) *seconds_scoring = t_score;
    if (seconds_rebuild) *seconds_rebuild = t_build;
    free(pick_begin);
    free(pool);
    free(ptrs);
    free(nodes);
    free(tmp_picks);
    return rc;
}

/* ======================================================================== */
/* R2: packed planes + type masks, multi-threaded.  Same decisions as above,  */
/* written against the int32 structure-of-arrays instead of pointer structs.  */
/* ======================================================================== */

typedef struct {
    int32_t used[REF_MAX_DEV], usedmem[REF_MAX_DEV], usedcores[REF_MAX_DEV];
    int8_t ord[REF_MAX_DEV]; /* ord[i] = slot at sorted position i */
} packed_node_state;

static inline int packed_less(const ref_Snapshot *s, size_t base, const packed_node_state *st,
                              int a, int b) {
    int32_t na = s->numa[base + a], nb = s->numa[base + b];
    if (na == nb)
        return i32_sub(s->count[base + a], st->used[a]) < i32_sub(s->count[base + b], st->used[b]);
    return na < nb;
}

/* returns 1 when node `row` fits pod p; *score_out = accumulated float32 */
static int packed_eval_node(const ref_Snapshot *s, int row, const ref_PodBatchPacked *pods, int p,
                            const int32_t *pick_begin, float *score_out, ref_Pick *picks_out) {
    int nd = s->ndev[row];
    size_t base = (size_t)row * s->dmax;
    packed_node_state st;
    for (int d = 0; d < nd; d++) {
        st.used[d] = s->used[base + d];
        st.usedmem[d] = s->usedmem[base + d];
        st.usedcores[d] = s->usedcores[base + d];
        st.ord[d] = (int8_t)d;
    }
    float score = 0.0f;
    for (int c = pods->ctr_begin[p]; c < pods->ctr_begin[p + 1]; c++) {
        int r0 = pods->req_begin[c], r1 = pods->req_begin[c + 1];
        int sums = 0;
        for (int r = r0; r < r1; r++) sums += pods->nums[r];
        if (sums == 0) continue;
        int32_t total = 0, free_ = 0;
        sums = 0;
        for (int r = r0; r < r1; r++) {
            int32_t nums = pods->nums[r];
            sums += nums;
            if (nums > nd) return 0;
            for (int i = 1; i < nd; i++) /* insertion sort on the permutation */
                for (int j = i; j > 0 && packed_less(s, base, &st, st.ord[j], st.ord[j - 1]); j--) {
                    int8_t t = st.ord[j];
                    st.ord[j] = st.ord[j - 1];
                    st.ord[j - 1] = t;
                }
            int32_t left = nums, prevnuma = -1;
            int taken[REF_MAX_DEV], nt = 0;
            int32_t takenmem[REF_MAX_DEV];
            int ok = 0;
            int32_t coresreq = pods->coresreq[r], memfix = pods->memreq[r], pct = pods->mempct[r];
            for (int i = nd - 1; i >= 0; i--) {
                int d = st.ord[i];
                size_t x = base + d;
                if (!((pods->typemask[r] >> s->type_id[x]) & 1)) continue;
                if ((pods->numa_bind[r] & 2) && st.used[d] > 0 &&
                    ref_strings_contains(s->type_names[s->type_id[x]], "370"))
                    continue; /* cambricon/device.go:97-99, part of checkType */
                if ((pods->numa_bind[r] & 1) && prevnuma != s->numa[x]) {
                    left = nums;
                    prevnuma = s->numa[x];
                    nt = 0;
                }
                int32_t memreq = 0;
                if (s->count[x] <= st.used[d]) continue;
                if (coresreq > 100) break;
                if (memfix > 0) memreq = memfix;
                if (pct != 101 && memfix == 0) memreq = i32_mul(s->totalmem[x], pct) / 100;
                if (i32_sub(s->totalmem[x], st.usedmem[d]) < memreq) continue;
                if (i32_sub(s->totalcore[x], st.usedcores[d]) < coresreq) continue;
                if (s->totalcore[x] == 100 && coresreq == 100 && st.used[d] > 0) continue;
                if (s->totalcore[x] != 0 && st.usedcores[d] == s->totalcore[x] && coresreq == 0)
                    continue;
                if (left > 0) {
                    left--;
                    taken[nt] = d;
                    takenmem[nt] = memreq;
                    nt++;
                }
                if (left == 0) {
                    ok = 1;
                    break;
                }
            }
            if (!ok) return 0;
            for (int t = 0; t < nt; t++) {
                int d = taken[t];
                total = i32_add(total, s->count[base + d]);
                free_ = i32_add(free_, i32_sub(s->count[base + d], st.used[d]));
                st.used[d] = i32_add(st.used[d], 1);
                st.usedcores[d] = i32_add(st.usedcores[d], coresreq);
                st.usedmem[d] = i32_add(st.usedmem[d], takenmem[t]);
                if (picks_out) {
                    ref_Pick *pk = &picks_out[pick_begin[r] + t];
                    pk->slot = d;
                    pk->usedmem = takenmem[t];
                    pk->usedcores = coresreq;
                }
            }
        }
        volatile float q = (float)total / (float)free_;
        float nodescore = q + (float)(nd - sums);
        score = score + nodescore;
    }
    *score_out = score;
    return 1;
}

typedef struct {
    const ref_Snapshot *snap;
    const ref_PodBatchPacked *pods;
    const int32_t *pick_begin;
    uint8_t *fit;
    float *scores;
    uint64_t *best; /* per thread [n_pods] : (score_bits<<32)|row, 0 = none */
    int32_t *cnt;   /* per thread [n_pods] */
    int64_t *next;  /* shared work counter */
    int64_t n_items;
    int blocks_per_pod, block_nodes;
} packed_job;

/* monotone map float32 bits -> uint32 (negative scores exist when one container
 * carries several requests: len(node.Devices)-sums < 0), never 0 for finite input */
static inline uint32_t f32_order(uint32_t b) { return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
static inline uint32_t f32_unorder(uint32_t o) { return (o & 0x80000000u) ? (o ^ 0x80000000u) : ~o; }

static void *packed_worker(void *arg) {
    packed_job *j = (packed_job *)arg;
    int N = j->snap->n_nodes;
    for (;;) {
        int64_t it = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED);
        if (it >= j->n_items) break;
        int p = (int)(it / j->blocks_per_pod);
        int b = (int)(it % j->blocks_per_pod);
        int lo = b * j->block_nodes, hi = lo + j->block_nodes;
        if (hi > N) hi = N;
        uint64_t best = j->best[p];
        int32_t cnt = 0;
        for (int row = lo; row < hi; row++) {
            float sc;
            int ok = packed_eval_node(j->snap, row, j->pods, p, j->pick_begin, &sc, NULL);
            if (j->fit) j->fit[(size_t)p * N + row] = (uint8_t)ok;
            if (j->scores) j->scores[(size_t)p * N + row] = ok ? sc : 0.0f;
            if (ok) {
                uint32_t bits;
                memcpy(&bits, &sc, 4);
                uint64_t key = ((uint64_t)f32_order(bits) << 32) | (uint32_t)row;
                if (key > best) best = key;
                cnt++;
            }
        }
        j->best[p] = best;
        j->cnt[p] += cnt;
    }
    return NULL;
}

int ref_score_batch_packed(const ref_Snapshot *snap, const ref_PodBatchPacked *pods,
                           ref_Result *results, ref_Pick *picks, uint8_t *fit, float *scores,
                           int n_threads, double *seconds) {
    if (snap->dmax > REF_MAX_DEV || snap->dmax < 1) return -1;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    int P = pods->n_pods, N = snap->n_nodes;
    int n_ctrs = pods->ctr_begin[P];
    int n_reqs = pods->req_begin[n_ctrs];
    int32_t *pick_begin = make_pick_begin(n_reqs, pods->nums);
    uint64_t *best = (uint64_t *)calloc((size_t)n_threads * (size_t)(P > 0 ? P : 1), 8);
    int32_t *cnt = (int32_t *)calloc((size_t)n_threads * (size_t)(P > 0 ? P : 1), 4);
    int64_t next = 0;
    packed_job jobs[256];
    pthread_t th[256];
    int block_nodes = 2048;
    int bpp = (N + block_nodes - 1) / block_nodes;
    if (bpp < 1) bpp = 1;
    double t0 = now_s();
    for (int t = 0; t < n_threads; t++) {
        packed_job *j = &jobs[t];
        j->snap = snap;
        j->pods = pods;
        j->pick_begin = pick_begin;
        j->fit = fit;
        j->scores = scores;
        j->best = best + (size_t)t * P;
        j->cnt = cnt + (size_t)t * P;
        j->next = &next;
        j->n_items = (int64_t)P * bpp;
        j->blocks_per_pod = bpp;
        j->block_nodes = block_nodes;
    }
    for (int t = 1; t < n_threads; t++) pthread_create(&th[t], NULL, packed_worker, &jobs[t]);
    packed_worker(&jobs[0]);
    for (int t = 1; t < n_threads; t++) pthread_join(th[t], NULL);
    for (int p = 0; p < P; p++) {
        uint64_t b = 0;
        int32_t c = 0;
        for (int t = 0; t < n_threads; t++) {
            if (best[(size_t)t * P + p] > b) b = best[(size_t)t * P + p];
            c += cnt[(size_t)t * P + p];
        }
        int c0 = pods->ctr_begin[p], c1 = pods->ctr_begin[p + 1];
        ref_Result res;
        res.fit_count = c;
        if (c > 0) {
            uint32_t bits = f32_unorder((uint32_t)(b >> 32));
            res.best_row = (int32_t)(uint32_t)b;
            memcpy(&res.score, &bits, 4);
            res.n_picks = pick_begin[pods->req_begin[c1]] - pick_begin[pods->req_begin[c0]];
            if (picks) {
                float sc;
                packed_eval_node(snap, res.best_row, pods, p, pick_begin, &sc, picks);
            }
        } else {
            res.best_row = -1;
            res.score = 0.0f;
            res.n_picks = 0;
        }
        if (results) results[p] = res;
    }
    if (seconds) *seconds = now_s() - t0;
    free(pick_begin);
    free(best);
    free(cnt);
    return 0;
}

/* ---- pkg/util/util.go:110-126 EncodeContainerDevices / EncodePodDevices --- */
int ref_encode_pod_devices(int n_ctrs, const int32_t *ctr_n, const char *const *uuids,
                           const char *const *types, const int32_t *usedmem,
                           const int32_t *usedcores, char *out, int cap) {
    int w = 0, t = 0;
    for (int c = 0; c < n_ctrs; c++) {
        if (c > 0) { /* strings.Join(ss, ";") */
            if (w + 1 >= cap) return -1;
            out[w++] = ';';
        }
        for (int j = 0; j < ctr_n[c]; j++, t++) {
            int n = snprintf(out + w, (size_t)(cap - w), "%s,%s,%d,%d:", uuids[t], types[t],
                             usedmem[t], usedcores[t]);
            if (n < 0 || n >= cap - w) return -1;
            w += n;
        }
    }
    if (w >= cap) return -1;
    out[w] = 0;
    return w;
}
