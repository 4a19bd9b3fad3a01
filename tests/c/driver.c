/* driver.c -- the call sequence the Go shim makes from Scheduler.Filter (scheduler.go:372-394), in plain C99:
 *
 *   vgs_init -> vgs_snapshot_type_flags -> vgs_snapshot_upload      (once / when the node set changes)
 *   per Filter:  vgs_host_typemask + vgs_host_numa_bind (string policy, stays on the host)
 *                -> vgs_score_batch(P = 1) -> vgs_host_encode_pod_devices -> vgs_snapshot_delta (addPod)
 *
 * against three hand-derived known-answer vectors of SURVEY.md 8(c) (KAT 1, 4, 8; literals below, never
 * computed).  Exit code 0 and "driver: ok" when every value matches.  Built and run by tests/test_c_driver.py. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "vgpuscore.h"

#define D 8
static const char *kTypes[2] = {"NVIDIA-Tesla V100-PCIE-32GB", "NVIDIA-NVIDIA A100-SXM4-40GB"};
static int failures = 0;

#define CHECK(cond, ...)                                                                           \
    do {                                                                                           \
        if (!(cond)) {                                                                             \
            printf("driver: FAIL %s:%d ", __FILE__, __LINE__);                                     \
            printf(__VA_ARGS__);                                                                   \
            printf("\n");                                                                          \
            failures++;                                                                            \
        }                                                                                          \
    } while (0)

typedef struct {
    int32_t *planes[VGS_NUM_PLANES];
    int32_t *ndev;
    int32_t n;
} snapshot;

static snapshot make_snapshot(int32_t n_nodes, const int32_t numa[D], const int32_t used[D]) {
    snapshot s;
    int f, r, j;
    s.n = n_nodes;
    for (f = 0; f < VGS_NUM_PLANES; f++) s.planes[f] = (int32_t *)calloc((size_t)n_nodes * D, sizeof(int32_t));
    s.ndev = (int32_t *)calloc((size_t)n_nodes, sizeof(int32_t));
    for (r = 0; r < n_nodes; r++) {
        s.ndev[r] = D;
        for (j = 0; j < D; j++) {
            s.planes[VGS_PLANE_USED][r * D + j] = used ? used[j] : 0;
            s.planes[VGS_PLANE_COUNT][r * D + j] = 10;
            s.planes[VGS_PLANE_TOTALMEM][r * D + j] = 32768;
            s.planes[VGS_PLANE_TOTALCORE][r * D + j] = 100;
            s.planes[VGS_PLANE_NUMA][r * D + j] = numa ? numa[j] : 0;
            s.planes[VGS_PLANE_TYPEID][r * D + j] = 0;
        }
    }
    return s;
}
static void free_snapshot(snapshot *s) {
    int f;
    for (f = 0; f < VGS_NUM_PLANES; f++) free(s->planes[f]);
    free(s->ndev);
}

static vgs_ctx *open_ctx(int32_t max_nodes) {
    vgs_config cfg;
    vgs_ctx *ctx = NULL;
    uint32_t flags[2];
    int rc;
    memset(&cfg, 0, sizeof cfg);
    cfg.device = 0;
    cfg.max_nodes = max_nodes;
    cfg.max_devices = D;
    cfg.max_pods = 1; /* Filter scores one pod per call */
    cfg.max_reqs = VGS_MAX_REQS_PER_POD;
    cfg.max_picks = VGS_MAX_REQS_PER_POD * D;
    rc = vgs_init(&cfg, &ctx);
    if (rc != VGS_OK) {
        printf("driver: vgs_init -> %d (%s)\n", rc, vgs_strerror(rc));
        exit(2);
    }
    flags[0] = vgs_host_type_flags(kTypes[0]);
    flags[1] = vgs_host_type_flags(kTypes[1]);
    rc = vgs_snapshot_type_flags(ctx, 2, flags);
    CHECK(rc == VGS_OK, "type flags -> %d", rc);
    return ctx;
}

/* one NVIDIA request per container, as k8sutil.Resourcereqs would produce it */
static int filter(vgs_ctx *ctx, int n_ctrs, const int32_t (*ctr)[4], const char *numa_anno, vgs_result *res,
                  vgs_pick *picks) {
    vgs_req reqs[4];
    vgs_pod pod;
    int c;
    const uint64_t mask = vgs_host_typemask("NVIDIA", NULL, NULL, kTypes, 2);
    const uint32_t numa = vgs_host_numa_bind(numa_anno) ? VGS_REQ_NUMA_BIND : 0u;
    for (c = 0; c < n_ctrs; c++) {
        reqs[c].ctr = c;
        reqs[c].nums = ctr[c][0];
        reqs[c].memreq = ctr[c][1];
        reqs[c].mempct = ctr[c][2];
        reqs[c].coresreq = ctr[c][3];
        reqs[c].flags = numa;
        reqs[c].typemask = mask;
    }
    pod.req_begin = 0;
    pod.n_reqs = n_ctrs;
    return vgs_score_batch(ctx, 1, &pod, n_ctrs, reqs, res, picks);
}

int main(void) {
    vgs_result res;
    vgs_pick picks[16];
    uint64_t ver = 0;
    int rc;
    CHECK(vgs_abi_version() == VGS_ABI_VERSION, "abi %d", vgs_abi_version());
    CHECK(sizeof(vgs_config) == 48 && sizeof(vgs_tuning) == 72 && sizeof(vgs_req) == 32 && sizeof(vgs_plan) == 176,
          "struct sizes %zu %zu %zu %zu", sizeof(vgs_config), sizeof(vgs_tuning), sizeof(vgs_req), sizeof(vgs_plan));

    { /* KAT 1: 64 idle nodes, 1 vGPU 3000M 30% (examples/nvidia/default_use.yaml) */
        const int32_t ctr[1][4] = {{1, 3000, VGS_MEMPCT_UNSET, 30}};
        snapshot s = make_snapshot(64, NULL, NULL);
        vgs_ctx *ctx = open_ctx(64);
        vgs_delta d;
        char anno[256];
        const char *uuids[1] = {"GPU-7"}, *types[1] = {"NVIDIA"};
        int32_t per_ctr[1] = {1};
        rc = vgs_snapshot_upload(ctx, s.n, D, (const int32_t *const *)s.planes, s.ndev, &ver);
        CHECK(rc == VGS_OK && ver == 1, "upload -> %d ver %llu: %s", rc, (unsigned long long)ver, vgs_last_error(ctx));
        rc = filter(ctx, 1, ctr, NULL, &res, picks);
        CHECK(rc == VGS_OK, "score -> %d: %s", rc, vgs_last_error(ctx));
        /* all 64 nodes tie at 10/10 + 7 = 8.0 -> canonical winner = highest row */
        CHECK(res.best_row == 63 && res.fit_count == 64 && res.n_picks == 1, "kat1 row %d fit %d picks %d", res.best_row,
              res.fit_count, res.n_picks);
        CHECK(memcmp(&res.score, &(uint32_t){0x41000000u}, 4) == 0, "kat1 score %f", res.score);
        CHECK(picks[0].slot == 7 && picks[0].usedmem == 3000 && picks[0].usedcores == 30, "kat1 pick %d %d %d",
              picks[0].slot, picks[0].usedmem, picks[0].usedcores);
        /* the annotation the shim writes (util.EncodePodDevices, util.go:110-126) */
        rc = vgs_host_encode_pod_devices(1, per_ctr, uuids, types, picks, anno, (int32_t)sizeof anno);
        CHECK(rc > 0 && strcmp(anno, "GPU-7,NVIDIA,3000,30:") == 0, "kat1 annotation '%s'", anno);
        /* addPod (scheduler.go:394 -> :280-297 on the next Filter): the placement becomes usage */
        d.row = res.best_row, d.slot = picks[0].slot, d.d_used = 1, d.d_usedmem = picks[0].usedmem,
        d.d_usedcores = picks[0].usedcores;
        rc = vgs_snapshot_delta(ctx, 1, &d);
        CHECK(rc == VGS_OK, "delta -> %d: %s", rc, vgs_last_error(ctx));
        /* same pod again: on row 63 slot 7 now sorts first (9 free), the walk starts at slot 6; still
         * 10/10 + 7 = 8.0 on every node -> row 63 again, slot 6 */
        rc = filter(ctx, 1, ctr, NULL, &res, picks);
        CHECK(rc == VGS_OK, "score 2 -> %d", rc);
        CHECK(res.best_row == 63 && res.fit_count == 64, "kat1b row %d fit %d", res.best_row, res.fit_count);
        CHECK(memcmp(&res.score, &(uint32_t){0x41000000u}, 4) == 0, "kat1b score %f", res.score);
        CHECK(picks[0].slot == 6 && picks[0].usedmem == 3000 && picks[0].usedcores == 30, "kat1b pick %d", picks[0].slot);
        vgs_destroy(ctx);
        free_snapshot(&s);
    }
    { /* KAT 4: two containers, the order persists: slot 7 then slot 6, 8 + 8 */
        const int32_t ctr[2][4] = {{1, 1000, VGS_MEMPCT_UNSET, 0}, {1, 1000, VGS_MEMPCT_UNSET, 0}};
        snapshot s = make_snapshot(1, NULL, NULL);
        vgs_ctx *ctx = open_ctx(1);
        rc = vgs_snapshot_upload(ctx, s.n, D, (const int32_t *const *)s.planes, s.ndev, &ver);
        CHECK(rc == VGS_OK, "upload -> %d", rc);
        rc = filter(ctx, 2, ctr, NULL, &res, picks);
        CHECK(rc == VGS_OK, "score -> %d: %s", rc, vgs_last_error(ctx));
        CHECK(res.best_row == 0 && res.fit_count == 1 && res.n_picks == 2, "kat4 row %d", res.best_row);
        CHECK(memcmp(&res.score, &(uint32_t){0x41800000u}, 4) == 0, "kat4 score %f", res.score);
        CHECK(picks[0].slot == 7 && picks[1].slot == 6 && picks[0].usedmem == 1000 && picks[1].usedcores == 0,
              "kat4 picks %d %d", picks[0].slot, picks[1].slot);
        vgs_destroy(ctx);
        free_snapshot(&s);
    }
    { /* KAT 8: numa-bind -- slot 7 (numa 1) is taken, then dropped at the numa change; 20/20 + 6 */
        const int32_t numa[D] = {0, 0, 0, 0, 1, 1, 1, 1}, used[D] = {0, 0, 0, 0, 10, 10, 10, 0};
        const int32_t ctr[1][4] = {{2, 1000, VGS_MEMPCT_UNSET, 0}};
        snapshot s = make_snapshot(1, numa, used);
        vgs_ctx *ctx = open_ctx(1);
        rc = vgs_snapshot_upload(ctx, s.n, D, (const int32_t *const *)s.planes, s.ndev, &ver);
        CHECK(rc == VGS_OK, "upload -> %d", rc);
        rc = filter(ctx, 1, ctr, "true", &res, picks);
        CHECK(rc == VGS_OK, "score -> %d: %s", rc, vgs_last_error(ctx));
        CHECK(res.best_row == 0 && res.n_picks == 2, "kat8 row %d", res.best_row);
        CHECK(memcmp(&res.score, &(uint32_t){0x40e00000u}, 4) == 0, "kat8 score %f", res.score);
        CHECK(picks[0].slot == 3 && picks[1].slot == 2, "kat8 picks %d %d", picks[0].slot, picks[1].slot);
        /* 8b: without the annotation the same request takes slots 7 and 3 */
        rc = filter(ctx, 1, ctr, NULL, &res, picks);
        CHECK(rc == VGS_OK && picks[0].slot == 7 && picks[1].slot == 3, "kat8b picks %d %d", picks[0].slot, picks[1].slot);
        /* a request nothing can satisfy is not an error (scheduler.go:381-385) */
        {
            const int32_t big[1][4] = {{9, 1000, VGS_MEMPCT_UNSET, 0}};
            rc = filter(ctx, 1, big, NULL, &res, picks);
            CHECK(rc == VGS_OK && res.best_row == -1 && res.fit_count == 0 && res.n_picks == 0, "kat11 rc %d row %d", rc,
                  res.best_row);
        }
        vgs_destroy(ctx);
        free_snapshot(&s);
    }
    if (failures) {
        printf("driver: %d check(s) failed\n", failures);
        return 1;
    }
    printf("driver: ok\n");
    return 0;
}
