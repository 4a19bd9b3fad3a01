"""Parity tests proper: the CUDA path through the C ABI against the CPU oracle, bit-exact
(winner row, score bits, fit counts, device picks, and in verify mode the whole
pod x node fit bitmap and score matrix)."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v
from helpers import assert_same, cambricon_cluster, check_kat, kat_inputs, load_kats

pytestmark = pytest.mark.gpu
K = load_kats()
FORCE_SLICED = {"slice_min": 1, "slice_small": 1, "slice_min_nodes": 0}  # vgs_tuning: every Nums 1 / Nums 2 class on the sliced kernels


def gpu_score(snap, batch, verify=True, **kw):
    # (the small snapshots of these tests keep reaching the sliced kernels: production routes snapshots under 6144
    # nodes to the per-pair kernel, test_small_snapshots_take_the_per_pair_kernel)
    kw["tuning"] = dict({"slice_min_nodes": 0}, **(kw.get("tuning") or {}))
    with v.Context(max_nodes=max(snap.n_nodes, 1), max_devices=snap.dmax, max_pods=max(batch.n_pods, 1),
                   max_reqs=max(batch.n_reqs, batch.n_pods, 1), max_picks=max(int(batch.nums.sum()), batch.n_reqs, 1),
                   verify=verify, **kw) as ctx:
        ctx.upload(snap)
        return ctx.score(batch)


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_kat_gpu(built, k):
    snap, batch = kat_inputs(k)
    check_kat(k, gpu_score(snap, batch), batch)


@pytest.mark.parametrize("cfg,n,p", [("cfg1", 64, 1), ("cfg2", 10_000, 256), ("cfg3", 5000, 400),
                                     ("cfg3", 33, 31), ("cfg4", 4000, 300), ("cfg4", 1000, 33),
                                     # <= 32 pods: candidates are kept per CTA across its node tiles
                                     ("cfg3", 70_000, 1), ("cfg3", 9000, 32), ("cfg4", 40_000, 7)])
def test_differential_full_matrix(built, cfg, n, p):
    snap, batch = v.workload.make_config(cfg, n, p)
    got = gpu_score(snap, batch)
    want = oracle.score_packed(snap, batch, full=True)
    assert_same(got, want, f"{cfg} {n}x{p}")
    assert want.fit.any()


def test_faithful_oracle_agrees_with_gpu(built):
    """string/pointer restatement (R1) vs the GPU path incl. host-built type masks"""
    snap, batch = v.workload.make_config("cfg3", 1500, 200)
    assert_same(gpu_score(snap, batch), oracle.score_faithful(snap, batch, full=True), "R1")
    snap, batch = v.workload.make_config("cfg4", 800, 100)
    assert_same(gpu_score(snap, batch), oracle.score_faithful(snap, batch, full=True), "R1 multi")


def _random_snapshot(rng, n, dmax, ragged=True, numa_vals=3, types=5):
    planes = np.zeros((8, n, dmax), np.int32)
    count = rng.integers(1, 12, (n, dmax))
    planes[1] = count
    planes[0] = np.minimum(rng.integers(0, 14, (n, dmax)), count + rng.integers(0, 2, (n, dmax)))
    totalmem = rng.choice([8000, 16000, 32768, 81920], (n, dmax))
    planes[3] = totalmem
    planes[2] = (totalmem * rng.random((n, dmax)) * (planes[0] > 0)).astype(np.int32)
    totalcore = rng.choice([100, 100, 100, 200, 0], (n, dmax))
    planes[5] = totalcore
    planes[4] = np.minimum(totalcore, rng.choice([0, 0, 10, 50, 100, 200], (n, dmax))) * (planes[0] > 0)
    planes[6] = rng.integers(0, numa_vals, (n, dmax)) * 3 - 1  # includes numa -1
    planes[7] = rng.integers(0, types, (n, dmax))
    ndev = rng.integers(0, dmax + 1, n).astype(np.int32) if ragged else np.full(n, dmax, np.int32)
    names = ["NVIDIA-Tesla V100", "NVIDIA-NVIDIA A100-SXM4-40GB", "NVIDIA-Tesla T4", "MLU370-X8", "NVIDIA-NVIDIA H100"][:types]
    return v.Snapshot(planes, ndev, names)


def _random_pods(rng, p, dmax, max_ctrs=4):
    pods = []
    for _ in range(p):
        ctrs = []
        for _ in range(int(rng.integers(1, max_ctrs + 1))):
            if rng.random() < 0.15:
                ctrs.append([])
                continue
            nums = int(rng.choice([1, 1, 2, 3, dmax, dmax + 1]))
            u = rng.random()
            if u < 0.4:
                mem, pct = int(rng.choice([0, 500, 4000, 20000])), 101
            elif u < 0.8:
                mem, pct = 0, int(rng.choice([0, 10, 50, 100, 101]))
            else:  # gpumem AND gpumem-percentage set (nvidia/device.go:125-145), negative values: score.go:114-120
                mem, pct = int(rng.choice([500, 4000, -5, -4000])), int(rng.choice([10, 50, 100, 0]))
            cores = int(rng.choice([0, 0, 10, 50, 100, 101, -10]))
            ctrs.append([(nums, mem, pct, cores)])
        if not any(ctrs):
            ctrs.append([(1, 0, 101, 0)])
        annos = {}
        r = rng.random()
        if r < 0.2:
            annos["nvidia.com/use-gputype"] = str(rng.choice(["a100", "V100,t4", "h100,", "mlu"]))
        elif r < 0.4:
            annos["nvidia.com/nouse-gputype"] = str(rng.choice(["a100", "V100,t4", "T4"]))
        if rng.random() < 0.35:
            annos["nvidia.com/numa-bind"] = str(rng.choice(["true", "1", "false", "T"]))
        pods.append({"containers": ctrs, "annos": annos})
    return v.PodBatch.from_pods(pods)


@pytest.mark.parametrize("dmax,seed", [(8, 1), (8, 2), (4, 3), (12, 4), (12, 5), (1, 6), (7, 7)])
def test_random_ragged_edge_cases(built, dmax, seed):
    """ragged nodes (0..dmax devices), hetero Count/Totalcore, negative numa ids, every
    request shape incl. Nums > devices, Coresreq > 100, empty containers, numa-bind"""
    rng = np.random.default_rng(seed)
    snap = _random_snapshot(rng, 777, dmax)
    batch = _random_pods(rng, 150, dmax)
    got = gpu_score(snap, batch)
    assert_same(got, oracle.score_packed(snap, batch, full=True), f"dmax={dmax}")
    assert_same(got, oracle.score_faithful(snap, batch, full=True), f"dmax={dmax} R1")


def test_multi_request_container(built):
    """two requests inside ONE container (the second sees the first's usage; sums spans both)"""
    rng = np.random.default_rng(11)
    snap = _random_snapshot(rng, 500, 8, ragged=False, types=3)
    pods = [{"containers": [[(2, 1000, 101, 10), (1, 0, 50, 0)], [(1, 100, 101, 0)]]},
            {"containers": [[(5, 100, 101, 0), (5, 100, 101, 0)]]},  # sums > len(devices): negative tail
            {"containers": [[(1, 100, 101, 0), (1, 100, 101, 0), (1, 100, 101, 0)]], "annos": {"nvidia.com/numa-bind": "true"}}]
    batch = v.PodBatch.from_pods(pods)
    assert_same(gpu_score(snap, batch), oracle.score_packed(snap, batch, full=True), "multi-req")


def test_shared_first_request_runs(built):
    """the walk kernel parks the node state after a first request that consecutive pods share: groups of pods
    with equal first requests and different continuations -- numa-bind, exclusive (usage-flag tracking, never
    shared), mixed vendors, a follower with a single request, percentages -- against the whole fit / score matrix"""
    rng = np.random.default_rng(41)
    snap = _random_snapshot(rng, 1500, 8, ragged=True, types=5)
    firsts = [(1, 1000, 101, 0), (2, 0, 50, 10), (1, 4000, 101, 100), (3, 500, 101, 0), (1, 0, 25, 0)]
    pods = []
    for i in range(400):
        f = firsts[int(rng.integers(0, len(firsts)))]
        ctrs = [[f]]
        for _ in range(int(rng.integers(0, 4))):
            if rng.random() < 0.5:
                ctrs.append([(int(rng.choice([1, 1, 2, 4])), int(rng.choice([500, 4000, 20000])), 101, int(rng.choice([0, 10, 100])))])
            else:
                ctrs.append([(int(rng.choice([1, 2])), 0, int(rng.choice([10, 50, 100])), int(rng.choice([0, 30])))])
        if rng.random() < 0.15:  # a second request inside the FIRST container
            ctrs[0] = [f, (1, 100, 101, 0)]
        annos = {}
        if rng.random() < 0.2:
            annos["nvidia.com/numa-bind"] = "true"
        if rng.random() < 0.2:
            annos["nvidia.com/use-gputype"] = "a100,t4"
        pods.append({"containers": ctrs, "annos": annos})
    batch = v.PodBatch.from_pods(pods)
    want = oracle.score_packed(snap, batch, full=True)
    assert want.fit.any()
    assert_same(gpu_score(snap, batch), want, "shared first request")
    assert_same(gpu_score(snap, batch, verify=False), oracle.score_packed(snap, batch), "shared first request (fast)")
    # the round-1 all-pairs kernel (vgs_tuning.general_legacy) must agree as well
    assert_same(gpu_score(snap, batch, tuning={"general_legacy": 1}), want, "all-pairs kernel")


def test_walk_kernel_many_picks_fall_back(built):
    """a pod with more picks than the walk kernel's arrival stamps hold (> 100) runs on the all-pairs kernel;
    one just below the limit runs on the walk kernel: same oracle"""
    node = [dict(Used=0, Count=40, Usedmem=0, Totalmem=32768, Usedcores=0, Totalcore=0, Numa=s // 4) for s in range(8)]
    snap = v.Snapshot.from_nodes([node] * 40)
    pods = [{"containers": [[(8, 10, 101, 0)] for _ in range(13)]},   # 104 picks
            {"containers": [[(8, 10, 101, 0)] for _ in range(12)]},   # 96 picks
            {"containers": [[(4, 10, 101, 0)] for _ in range(25)]}]   # 100 picks
    batch = v.PodBatch.from_pods(pods)
    want = oracle.score_packed(snap, batch, full=True)
    assert want.fit.all()
    assert_same(gpu_score(snap, batch), want, "many picks")


def test_empty_and_tiny(built):
    _, batch = v.workload.make_config("cfg3", 8, 40)
    empty = v.Snapshot(np.zeros((8, 0, 8), np.int32), np.zeros(0, np.int32), [t for t, _ in v.workload.GPU_TYPES])
    got = gpu_score(empty, batch, verify=False)
    assert (got.best_row == -1).all() and (got.fit_count == 0).all()
    snap, one = v.workload.make_config("cfg3", 1, 1)
    assert_same(gpu_score(snap, one), oracle.score_packed(snap, one, full=True), "1x1")


def test_error_codes(built):
    snap, batch = v.workload.make_config("cfg2", 100, 4)
    names = snap.type_names
    with v.Context(max_nodes=100, max_pods=4) as ctx:
        pods, reqs = v.capi.pack_batch(batch, names)
        with pytest.raises(v.VgsError) as e:
            ctx.score_packed(pods, reqs)  # no snapshot yet
        assert e.value.code == v.capi.VGS_ESTATE
        ctx.upload(snap)
        bad = reqs.copy()
        bad["nums"][0] = 0
        with pytest.raises(v.VgsError) as e:
            ctx.score_packed(pods, bad)
        assert e.value.code == v.capi.VGS_EINVAL
        big, _ = v.workload.make_config("cfg2", 101, 1)
        with pytest.raises(v.VgsError) as e:
            ctx.upload(big)
        assert e.value.code == v.capi.VGS_ETOOBIG
        rng = snap.planes.copy()
        rng[1, 5, 2] = 1 << 21
        with pytest.raises(v.VgsError) as e:
            ctx.upload(v.Snapshot(rng, snap.ndev, names))
        assert e.value.code == v.capi.VGS_ERANGE
        nd = snap.ndev.copy()
        nd[3] = 9
        with pytest.raises(v.VgsError) as e:
            ctx.upload(v.Snapshot(snap.planes, nd, names))
        assert e.value.code == v.capi.VGS_EDMAX
        ctx.upload(snap)  # context still usable
        assert_same(ctx.score(batch), oracle.score_packed(snap, batch), "after errors")
    with pytest.raises(v.VgsError) as e:
        v.Context(max_nodes=10, max_devices=13)
    assert e.value.code == v.capi.VGS_EDMAX


def test_delta_equals_rebuild(built):
    """commit the winners as usage deltas (addPod) == re-uploading a rebuilt snapshot"""
    snap, batch = v.workload.make_config("cfg4", 3000, 64)
    with v.Context(max_nodes=3000, max_pods=64, max_reqs=512, max_picks=1024, verify=True) as ctx:
        ctx.upload(snap)
        first = ctx.score(batch)
        planes = snap.planes.copy()
        deltas = []
        pb = batch.pick_begin
        for p in range(batch.n_pods):
            if first.best_row[p] < 0:
                continue
            r0, r1 = batch.pod_req_range(p)
            for r in range(r0, r1):
                for j in range(int(batch.nums[r])):
                    s, m, c = (int(x) for x in first.picks[int(pb[r]) + j])
                    row = int(first.best_row[p])
                    deltas.append((row, s, 1, m, c))
                    planes[0, row, s] += 1
                    planes[2, row, s] += m
                    planes[4, row, s] += c
        ctx.delta(np.array(deltas, dtype=v.capi.DELTA_DTYPE))
        got_planes, got_ndev = ctx.download_snapshot()
        assert np.array_equal(got_planes, planes) and np.array_equal(got_ndev, snap.ndev)
        rebuilt = v.Snapshot(planes, snap.ndev, snap.type_names)
        after = ctx.score(batch)
        assert_same(after, oracle.score_packed(rebuilt, batch, full=True), "after delta")
        # undo (delPod) -> original answers again
        undo = np.array([(r, s, -1, -m, -c) for r, s, _, m, c in deltas], dtype=v.capi.DELTA_DTYPE)
        ctx.delta(undo)
        assert_same(ctx.score(batch), first, "after undo")


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_kat_single_launch_filter(built, k):
    """the Filter call proper -- one pod, snapshot resident, no verify buffers -- runs as ONE kernel launch (pod as
    launch parameters, last CTA resolves): every known-answer vector through it, and through the phased path"""
    snap, batch = kat_inputs(k)
    for one_launch in (1, 0):
        with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=1, max_reqs=32, max_picks=256,
                       tuning={"filter_one_launch": one_launch}) as ctx:
            ctx.upload(snap)
            l0 = ctx.timing().kernel_launches
            res = ctx.score(batch)
            launches = ctx.timing().kernel_launches - l0
            check_kat(k, res, batch)
            check_kat(k, ctx.score(batch), batch)  # and again: the candidate block and the CTA counter are back at rest
        if batch.n_reqs <= 8:
            assert launches == (1 if one_launch else 2), (one_launch, launches)


def test_single_launch_filter_random_pods(built):
    """random single pods (multi-container, numa-bind, percentages, exclusive, > devices) one call each against a
    ragged snapshot: winner, score bits, fit count and picks equal the oracle's; 12-slot nodes too"""
    for dmax, seed in ((8, 31), (12, 32), (3, 33)):
        rng = np.random.default_rng(seed)
        snap = _random_snapshot(rng, 3000 + seed, dmax)
        batch = _random_pods(rng, 60, dmax)
        want = oracle.score_packed(snap, batch)
        with v.Context(max_nodes=snap.n_nodes, max_devices=dmax, max_pods=2, max_reqs=64, max_picks=512) as ctx:
            ctx.upload(snap)
            pb = batch.pick_begin
            for p in range(batch.n_pods):
                got = ctx.score(batch.select([p]))
                r0, r1 = batch.pod_req_range(p)
                assert got.best_row[0] == want.best_row[p], (dmax, p)
                assert got.score.view(np.uint32)[0] == want.score.view(np.uint32)[p], (dmax, p)
                assert got.fit_count[0] == want.fit_count[p], (dmax, p)
                assert np.array_equal(got.picks, want.picks[int(pb[r0]):int(pb[r1])]), (dmax, p)
            # a batch of two goes the phased way and leaves the single-launch path usable afterwards
            two = ctx.score(batch.select([0, 1]))
            assert np.array_equal(two.best_row, want.best_row[:2])
            assert ctx.score(batch.select([5])).best_row[0] == want.best_row[5]


def test_delta_errors_surface_at_the_next_status_read(built):
    """vgs_snapshot_delta only queues work; a delta naming a slot the node does not have is reported by the next
    call that reads the status word, and the snapshot then has to be uploaded again"""
    snap, batch = v.workload.make_config("cfg3", 500, 8)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    with v.Context(max_nodes=500, max_pods=8) as ctx:
        ctx.upload(snap)
        good = ctx.score_packed(pods, reqs)
        ctx.delta(np.array([(3, 11, 1, 100, 10)], dtype=v.capi.DELTA_DTYPE))  # slot 11 of an 8-device node
        with pytest.raises(v.VgsError) as e:
            ctx.score_packed(pods, reqs)
        assert e.value.code == v.capi.VGS_EINVAL
        with pytest.raises(v.VgsError) as e:
            ctx.score_packed(pods, reqs)
        assert e.value.code == v.capi.VGS_ESTATE
        ctx.upload(snap)
        again = ctx.score_packed(pods, reqs)
        assert np.array_equal(good[0], again[0]) and np.array_equal(good[1], again[1])
        ctx.delta(np.array([(3, 0, 1 << 21, 0, 0)], dtype=v.capi.DELTA_DTYPE))  # Used out of range
        with pytest.raises(v.VgsError) as e:
            ctx.sync()
        assert e.value.code == v.capi.VGS_ERANGE
        # a delta for a row of another shard is not an error
        ctx.upload(snap)
        ctx.delta(np.array([(100_000, 0, 1, 1, 1)], dtype=v.capi.DELTA_DTYPE))
        ctx.sync()
        again = ctx.score_packed(pods, reqs)
        assert np.array_equal(good[0], again[0])


def test_shards_on_one_gpu_equal_single(built):
    """3 row shards (3 contexts on one GPU) + key merge == one context; 1-GPU result is the oracle"""
    import torch
    snap, batch = v.workload.make_config("cfg4", 5000, 96)
    single = gpu_score(snap, batch, verify=False)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    G, P = 3, batch.n_pods
    n_picks = int(reqs["nums"].sum())
    ctxs = []
    for g in range(G):
        lo, hi = v.sharding.shard_range(snap.n_nodes, G, g)
        c = v.Context(max_nodes=hi - lo, max_pods=P, max_reqs=len(reqs), max_picks=n_picks, row_base=lo)
        c.upload(snap.rows(lo, hi))
        c.pods_upload(pods, reqs)
        c.score_resident()
        c.sync()
        ctxs.append(c)
    dev = torch.device("cuda", 0)
    gk = torch.zeros((G, P), dtype=torch.int64, device=dev)
    gc = torch.zeros((G, P), dtype=torch.int32, device=dev)
    for g, c in enumerate(ctxs):
        kp, cp = c.keys_device()
        gk[g] = v.sharding.device_view(kp, P, torch.int64, dev)
        gc[g] = v.sharding.device_view(cp, P, torch.int32, dev)
    torch.cuda.synchronize()
    picks = np.zeros((n_picks, 3), np.int32)
    res = None
    for c in ctxs:
        c.keys_merge(gk.data_ptr(), gc.data_ptr(), G, P)
        c.resolve_resident()
        r, pk = c.results_download()
        picks += np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1)
        res = r if res is None else res
        assert np.array_equal(r, res)
        c.close()
    assert np.array_equal(res["best_row"], single.best_row)
    assert np.array_equal(res["score"].view(np.uint32), single.score.view(np.uint32))
    assert np.array_equal(res["fit_count"], single.fit_count)
    assert np.array_equal(picks, single.picks)


@pytest.mark.parametrize("cfg", ["cfg3", "cfg4"])
def test_full_size_properties(built, cfg):
    """BASELINE configs 3 and 4 at full size (100k nodes x 4096 pods): properties that do not
    need the oracle over all 4e8 pairs -- determinism, a sampled-pod oracle check over all
    nodes, and every winner re-checked by the oracle on its own node."""
    snap, batch = v.workload.make_config(cfg)
    with v.Context(max_nodes=snap.n_nodes, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        a = ctx.score(batch)
        b = ctx.score(batch)
    assert_same(a, b, "idempotence")
    assert (a.best_row >= 0).mean() > 0.8
    idx = list(range(0, batch.n_pods, 64))
    sub = batch.select(idx)
    want = oracle.score_packed(snap, sub)
    assert np.array_equal(a.best_row[idx], want.best_row)
    assert np.array_equal(a.score[idx].view(np.uint32), want.score.view(np.uint32))
    assert np.array_equal(a.fit_count[idx], want.fit_count)
    # every pod: oracle on the winner node alone must reproduce score and picks
    pb = batch.pick_begin
    placed = np.nonzero(a.best_row >= 0)[0]
    for p in placed[:: max(1, len(placed) // 512)]:
        row = int(a.best_row[p])
        one = oracle.score_packed(snap.rows(row, row + 1), batch.select([int(p)]))
        assert one.best_row[0] == 0
        assert one.score.view(np.uint32)[0] == a.score.view(np.uint32)[p]
        r0, r1 = batch.pod_req_range(int(p))
        assert np.array_equal(one.picks, a.picks[int(pb[r0]):int(pb[r1])])


# ---- pod-sliced K1 (classes of single-request pods with Nums 1 / 2) -------------------------

def test_sliced_default_threshold(built):
    """classes >= 512 pods take the pod-sliced kernels; whole fit/score matrix vs the oracle"""
    # 9000 nodes: enough tiles for the two-launch (scouts + main) path of the sliced kernel
    for cfg, n, p in (("cfg3", 3000, 2600), ("cfg2", 5000, 1024), ("cfg3", 40, 4096), ("cfg3", 9000, 1400)):
        snap, batch = v.workload.make_config(cfg, n, p)
        assert_same(gpu_score(snap, batch), oracle.score_packed(snap, batch, full=True), f"sliced {cfg}")


@pytest.mark.parametrize("node_words,tabsmem,scout", [(10, -1, 0), (0, 0, 5), (31, 1, 17), (-1, -1, 64)])
def test_sliced_word_split_between_kernels(built, node_words, tabsmem, scout):
    """a class's pod words beyond a multiple of 32 go to the node-sliced kernel (lane = node), the rest to
    pod-sliced warps (lane = pod word); vgs_tuning.node_words moves the split, tables_in_smem / scout_stride
    the pod-sliced kernel's variant and launch geometry.  Same answers wherever a word runs -- and
    vgs_last_plan shows that the knobs really moved the split."""
    # (the scout tiles: a cold pod-sliced launch in the two cases that steer its geometry, the node-sliced kernel --
    # the default -- in the others)
    tune = {"node_words": node_words, "slice_min": 64, "slice_small": 64, "slice_min_nodes": 0, "tables_in_smem": tabsmem,
            "scout_stride": scout, "scout_split": 4 if scout == 17 else 2}
    if scout in (5, 17):
        tune["scouts_node_sliced"] = 0
    rng = np.random.default_rng(77)
    snap, _ = v.workload.make_config("cfg3", 9000, 1)
    def pods_of(n1, n2):
        out = []
        for i in range(n1 + n2):
            nums = 1 if i < n1 else 2
            if rng.random() < 0.5:
                mem, pct = int(rng.choice([1000, 4000, 16000, 30000])), 101
            else:
                mem, pct = 0, int(rng.choice([10, 25, 50, 100]))
            an = {"nvidia.com/use-gputype": "a100,t4"} if rng.random() < 0.2 else {}
            out.append({"containers": [[(nums, mem, pct, int(rng.choice([0, 10, 30, 100])))]], "annos": an})
        return v.PodBatch.from_pods(out)
    # (Nums 1 pods, Nums 2 pods): 33 + 3 words / 40 + 35 words, a few-word class next to a big one, exact multiples
    for n1, n2 in ((1040, 70), (1270, 1100), (200, 2048), (1024, 1024 + 5)):
        batch = pods_of(n1, n2)
        want = oracle.score_packed(snap, batch, full=True)
        assert_same(gpu_score(snap, batch, tuning=tune), want, f"verify {n1}+{n2} words={node_words}")
        with v.Context(max_nodes=snap.n_nodes, max_pods=batch.n_pods, tuning=tune) as ctx:
            ctx.upload(snap)
            assert_same(ctx.score(batch), oracle.score_packed(snap, batch), f"{n1}+{n2} words={node_words}")
            plan = ctx.plan()
        for k, n in ((0, n1), (1, n2)):
            cp, W = plan.cls[k], (n + 31) // 32
            assert cp.pods == n and cp.kernel == v.capi.KERNEL_SLICED, (k, cp.pods, cp.kernel)
            assert cp.words_pod_sliced + cp.words_node_sliced == W
            extra = W % 32
            if node_words == 0:
                assert cp.words_node_sliced == 0
            elif node_words == 31:
                assert cp.words_node_sliced == extra  # every word beyond a multiple of 32
            elif node_words == 10:
                assert cp.words_node_sliced == (extra if extra <= 10 else 0)
            if cp.words_pod_sliced:
                if tabsmem >= 0:
                    assert cp.tables_in_smem == tabsmem
                if scout >= 2:  # 9000 nodes = 282 tiles: scouts + main launch
                    assert cp.scout_stride == scout and cp.tiles_per_cta_main >= 2
                    if tune.get("scouts_node_sliced") == 0:
                        assert cp.scout_split in (1, 2, 4) and cp.tiles_per_cta_scout >= 1
                    else:
                        assert cp.scout_split == 0 and cp.tiles_per_cta_scout == 0  # no pod-sliced scout launch


def test_small_snapshots_take_the_per_pair_kernel(built):
    """default routing: a big class on a snapshot under vgs_tuning.slice_min_nodes runs on the per-pair kernel, on a
    larger one on the sliced kernels -- same answers"""
    for n, kernel in ((3000, v.capi.KERNEL_PAIR), (7000, v.capi.KERNEL_SLICED)):
        snap, batch = v.workload.make_config("cfg3", n, 2600)  # both single-request classes above slice_min (512) pods
        with v.Context(max_nodes=n, max_pods=2600) as ctx:
            ctx.upload(snap)
            got = ctx.score(batch)
            plan = ctx.plan()
        assert plan.cls[0].kernel == kernel and plan.cls[1].kernel == kernel, (n, plan.cls[0].kernel, plan.cls[1].kernel)
        assert_same(got, oracle.score_packed(snap, batch), f"routing {n}")


def test_sliced_long_scout_span_keeps_rows(built):
    """a scout CTA whose tiles lie far apart (large stride x many tiles per CTA, forced by many pod-word
    columns and a big stride) still names the right winner row: the per-pod row inside a CTA is kept as
    (virtual tile, lane), not as a 16-bit node offset"""
    snap, batch = v.workload.make_config("cfg3", 140_000, 1)
    rng = np.random.default_rng(3)
    pods = [{"containers": [[(int(rng.choice([1, 2])), int(rng.choice([1000, 3000, 9000])), 101, int(rng.choice([0, 10, 30])))]]}
            for _ in range(2048)]
    batch = v.PodBatch.from_pods(pods)
    # 4375 tiles, stride 32 -> 137 scout tiles on 2 CTAs = 69 tiles each, 2177 tiles (69664 nodes) apart
    tune = {"scout_stride": 32, "scout_split": 1, "slice_min": 64, "max_ctas": 2, "scouts_node_sliced": 0}
    with v.Context(max_nodes=snap.n_nodes, max_pods=batch.n_pods, tuning=tune) as ctx:
        ctx.upload(snap)
        got = ctx.score(batch)
        plan = ctx.plan()
    idx = list(range(0, batch.n_pods, 97))
    want = oracle.score_packed(snap, batch.select(idx))
    assert np.array_equal(got.best_row[idx], want.best_row)
    assert np.array_equal(got.score[idx].view(np.uint32), want.score.view(np.uint32))
    assert np.array_equal(got.fit_count[idx], want.fit_count)
    # the geometry this test is about: ((tiles_per - 1) * stride + 1) * 32 nodes per scout CTA
    for cp in (plan.cls[0], plan.cls[1]):
        assert cp.kernel == v.capi.KERNEL_SLICED and cp.scout_stride == 32
        assert ((cp.tiles_per_cta_scout - 1) * cp.scout_stride + 1) * 32 > 65535


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_sliced_forced_kats(built, k):
    snap, batch = kat_inputs(k)
    check_kat(k, gpu_score(snap, batch, tuning=FORCE_SLICED), batch)


@pytest.mark.parametrize("dmax,seed", [(8, 21), (8, 22), (7, 23), (4, 24), (1, 25), (2, 26)])
def test_sliced_forced_random(built, dmax, seed):
    rng = np.random.default_rng(seed)
    snap = _random_snapshot(rng, 1000 + seed, dmax)
    batch = _random_pods(rng, 300, dmax, max_ctrs=1)  # single-container pods -> classes A/B/C
    got = gpu_score(snap, batch, tuning=FORCE_SLICED)
    assert_same(got, oracle.score_packed(snap, batch, full=True), f"sliced dmax={dmax}")
    assert_same(got, oracle.score_faithful(snap, batch, full=True), f"sliced dmax={dmax} R1")


@pytest.mark.parametrize("n,p,tune", [(3000, 700, None), (9500, 1300, None), (2000, 300, {"slice_min": 1, "slice_small": 1, "slice_min_nodes": 0})])
def test_stress_snapshot_many_score_values(built, n, p, tune):
    """the robustness workload: Count 1..32 per device, ragged nodes, 48 types mixed inside nodes, scores rising
    along the walk, thousands of distinct candidate scores (more than the 32 score classes and, at size, more than
    the sliced kernels' hash set holds): whole fit / score matrix against both oracle arms"""
    snap, batch = v.workload.make_config("stress", n, p)
    want = oracle.score_packed(snap, batch, full=True)
    assert len(np.unique(want.scores[want.fit])) > 500
    got = gpu_score(snap, batch, tuning=tune)
    assert_same(got, want, "stress")
    assert_same(got, oracle.score_faithful(snap, batch, full=True), "stress R1")
    assert_same(gpu_score(snap, batch, verify=False, tuning=tune), oracle.score_packed(snap, batch), "stress fast")


def test_sliced_many_distinct_thresholds(built):
    """every pod its own Memreq / Coresreq value: long cumulative tables, binary search cut-offs"""
    rng = np.random.default_rng(5)
    snap, _ = v.workload.make_config("cfg3", 2000, 1)
    pods = [{"containers": [[(int(rng.choice([1, 2])), int(rng.integers(0, 40000)), 101, int(rng.integers(0, 101)))]]}
            for _ in range(900)]
    pods += [{"containers": [[(int(rng.choice([1, 2])), 0, int(rng.integers(0, 101)), int(rng.integers(0, 101)))]]}
             for _ in range(300)]
    batch = v.PodBatch.from_pods(pods)
    assert_same(gpu_score(snap, batch, tuning=FORCE_SLICED), oracle.score_packed(snap, batch, full=True), "distinct")


def test_sliced_falls_back_when_preconditions_fail(built):
    """Totalmem * pct wrapping int32 (score.go:119) and pct outside [0,100] are only handled by
    the per-pair kernels: the library must route there and still match the oracle"""
    snap, batch = v.workload.make_config("cfg3", 1500, 600)
    big = snap.planes.copy()
    big[3] = 30_000_000       # * 100 wraps int32
    big[2] = big[2] * 900
    wrapped = v.Snapshot(big, snap.ndev, snap.type_names)
    assert_same(gpu_score(wrapped, batch, tuning=FORCE_SLICED), oracle.score_packed(wrapped, batch, full=True), "bigmem")
    odd = v.PodBatch.from_pods([{"containers": [[(1, 0, pct, 0)]]} for pct in (150, -5, 100, 0, 250)] * 20)
    assert_same(gpu_score(snap, odd, tuning=FORCE_SLICED), oracle.score_packed(snap, odd, full=True), "odd pct")
    # and the same context flips back to the sliced path when a normal snapshot returns
    with v.Context(max_nodes=1500, max_pods=600, verify=True, tuning=FORCE_SLICED) as ctx:
        ctx.upload(wrapped)
        a = ctx.score(batch)
        assert ctx.plan().cls[0].kernel == v.capi.KERNEL_PAIR
        # a usage delta re-prepares one row only: the snapshot must stay off the sliced kernels
        ctx.delta(np.array([(7, 0, 0, 0, 0)], dtype=v.capi.DELTA_DTYPE))
        a2 = ctx.score(batch)
        assert ctx.plan().cls[0].kernel == v.capi.KERNEL_PAIR
        assert_same(a2, a, "bigmem stays sticky across a delta")
        ctx.upload(snap)
        b = ctx.score(batch)
        assert ctx.plan().cls[0].kernel == v.capi.KERNEL_SLICED
    assert_same(a, oracle.score_packed(wrapped, batch, full=True), "ctx bigmem")
    assert_same(b, oracle.score_packed(snap, batch, full=True), "ctx normal")


def test_mixed_vendor_cluster_nvidia_and_hygon(built):
    """NVIDIA and Hygon DCU devices in one snapshot: a DCU request only sees
    DCU cards and honours hygon.com/use-dcutype / nouse-dcutype; numa-bind is an NVIDIA-only notion"""
    rng = np.random.default_rng(8)
    nodes = []
    for i in range(600):
        t = ["NVIDIA-Tesla V100", "DCU-Z100", "DCU-Z100L", "NVIDIA-NVIDIA A100"][i % 4]
        nodes.append([dict(Used=int(rng.integers(0, 4)), Count=4, Usedmem=int(rng.integers(0, 8000)), Totalmem=16384,
                           Usedcores=int(rng.choice([0, 30, 60])), Totalcore=100, Numa=s // 4, Type=t) for s in range(8)])
    snap = v.Snapshot.from_nodes(nodes)
    pods = []
    for i in range(120):
        typ = "DCU" if i % 2 else "NVIDIA"
        annos = {}
        if i % 3 == 0:
            annos["hygon.com/use-dcutype"] = "z100l"
        if i % 5 == 0:
            annos["hygon.com/nouse-dcutype"] = "Z100L"
        if i % 7 == 0:
            annos["nvidia.com/use-gputype"] = "a100"
        if i % 4 == 0:
            annos["nvidia.com/numa-bind"] = "true"
        pods.append({"containers": [[(int(rng.choice([1, 2])), int(rng.choice([0, 2000])), 101, 10, typ)]], "annos": annos})
    batch = v.PodBatch.from_pods(pods)
    got = gpu_score(snap, batch)
    assert_same(got, oracle.score_faithful(snap, batch, full=True), "hygon R1")
    assert_same(got, oracle.score_packed(snap, batch, full=True), "hygon R2")
    # a DCU pod never lands on an NVIDIA node
    for p in range(batch.n_pods):
        if got.best_row[p] >= 0:
            assert ("DCU" in nodes[got.best_row[p]][0]["Type"]) == (batch.req_type[p] == "DCU")


def test_cambricon_vendor(built):
    """Cambricon's CheckType (cambricon/device.go:93-104) depends on the request's
    Memreq and on the device's Used AS THE POD'S EARLIER REQUESTS LEFT IT; full fit/score matrices
    against both oracle arms, then the same batch sharded and drained"""
    rng = np.random.default_rng(370)
    snap, batch = cambricon_cluster(rng, 900, 260)
    got = gpu_score(snap, batch)
    assert_same(got, oracle.score_faithful(snap, batch, full=True), "mlu R1")
    assert_same(got, oracle.score_packed(snap, batch, full=True), "mlu R2")
    assert 0 < int((got.best_row >= 0).sum()) < batch.n_pods
    # the rule bites: with the type flags cleared the same pods see more nodes
    with v.Context(max_nodes=snap.n_nodes, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum()), verify=True) as ctx:
        ctx.upload(snap)
        ctx.set_type_flags([0] * len(snap.type_names))
        loose = ctx.score(batch)
        assert int(loose.fit_count.sum()) > int(got.fit_count.sum())
        ctx.set_type_flags([v.capi.host_type_flags(t) for t in snap.type_names])
        assert_same(ctx.score(batch), got, "flags restored")
    # single-request pods only, enough of them to reach the pod-sliced kernels' threshold:
    # idle-share requests must be kept out of those classes
    pods = [{"containers": [[(int(rng.choice([1, 2])), int(rng.choice([0, 0, 3000])), 0, 0, "MLU")]]} for _ in range(1300)]
    big = v.PodBatch.from_pods(pods)
    assert_same(gpu_score(snap, big, verify=False), oracle.score_packed(snap, big), "mlu classes")


def test_cambricon_drain_and_shards(built):
    rng = np.random.default_rng(371)
    snap, batch = cambricon_cluster(rng, 400, 90)
    want = oracle.score_packed(snap, batch)
    halves = []
    for lo, hi in ((0, 200), (200, 400)):
        with v.Context(max_nodes=200, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                       max_picks=int(batch.nums.sum()), row_base=lo) as ctx:
            ctx.upload(snap.rows(lo, hi))
            halves.append(ctx.score(batch))
    for p in range(batch.n_pods):
        cand = [(h.score[p], h.best_row[p]) for h in halves if h.best_row[p] >= 0]
        assert (max(cand)[1] if cand else -1) == want.best_row[p]
        assert sum(int(h.fit_count[p]) for h in halves) == want.fit_count[p]
    # sequential placement: pod p sees the cards pods 0..p-1 took (an MLU370 taken once is gone for
    # every later memory-less request)
    rows, scores, counts, picks, planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=snap.n_nodes, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
    assert np.array_equal(got.best_row, rows)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32))
    assert np.array_equal(got.picks, picks)
    assert np.array_equal(after, planes)


@pytest.mark.parametrize("cfg,n,p", [("cfg4", 5000, 96), ("cfg3", 9000, 1500)])
def test_shards_exchange_in_one_collective(built, cfg, n, p):
    """vgs_resolve_local + one gather of the [keys | counts | picks] blocks + vgs_exchange_merge: every
    shard ends up with the single-context answer (winner, score bits, summed counts, the winner's picks)"""
    import torch
    snap, batch = v.workload.make_config(cfg, n, p)
    single = gpu_score(snap, batch, verify=False)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    G, P = 3, batch.n_pods
    n_picks = int(reqs["nums"].sum())
    dev = torch.device("cuda", 0)
    ctxs, blocks = [], []
    for g in range(G):
        lo, hi = v.sharding.shard_range(snap.n_nodes, G, g)
        c = v.Context(max_nodes=hi - lo, max_pods=P, max_reqs=len(reqs), max_picks=n_picks, row_base=lo)
        c.upload(snap.rows(lo, hi))
        ctxs.append(c)
    for rep in range(2):  # twice: the merge must leave every block clean for the next pass
        blocks = []
        for c in ctxs:
            c.pods_upload(pods, reqs)
            c.score_resident()
            c.resolve_local()
            c.sync()
            bp, nb = c.exchange_block()
            blocks.append(v.sharding.device_view(bp, nb // 4, torch.int32, dev).clone())
        gathered = torch.stack(blocks).contiguous()
        torch.cuda.synchronize()
        for c in ctxs:
            c.exchange_merge(gathered.data_ptr(), G, gathered.stride(0) * 4)
            r, pk = c.results_download()
            assert np.array_equal(r["best_row"], single.best_row)
            assert np.array_equal(r["score"].view(np.uint32), single.score.view(np.uint32))
            assert np.array_equal(r["fit_count"], single.fit_count)
            assert np.array_equal(r["n_picks"], single.n_picks)
            assert np.array_equal(np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1), single.picks)
    for c in ctxs:
        c.close()


def test_cfg5_shape_eight_shards(built):
    """BASELINE config 5 at full size: 1M nodes x 8 devices x 16384 pods as 8 row shards of 125k nodes (here: 8
    contexts on one GPU), every shard resolving its own best node, ONE gather of the exchange blocks, merge.
    Checked against the oracle over all 1M nodes for a sample of pods, and every shard must hold the same answer."""
    import torch
    snap, batch = v.workload.make_config("cfg5")
    assert snap.n_nodes == 1_000_000 and batch.n_pods == 16384
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    G, P = 8, batch.n_pods
    n_picks = int(reqs["nums"].sum())
    dev = torch.device("cuda", 0)
    ctxs, blocks = [], []
    for g in range(G):
        lo, hi = v.sharding.shard_range(snap.n_nodes, G, g)
        c = v.Context(max_nodes=hi - lo, max_pods=P, max_reqs=len(reqs), max_picks=n_picks, row_base=lo)
        c.upload(snap.rows(lo, hi))
        c.pods_upload(pods, reqs)
        c.score_resident()
        c.resolve_local()
        c.sync()
        plan = c.plan()
        assert plan.cls[0].kernel == v.capi.KERNEL_SLICED and plan.cls[1].kernel == v.capi.KERNEL_SLICED
        bp, nb = c.exchange_block()
        blocks.append(v.sharding.device_view(bp, nb // 4, torch.int32, dev).clone())
        ctxs.append(c)
    gathered = torch.stack(blocks).contiguous()
    torch.cuda.synchronize()
    first = None
    for c in ctxs:
        c.exchange_merge(gathered.data_ptr(), G, gathered.stride(0) * 4)
        r, pk = c.results_download()
        if first is None:
            first = (r, pk)
        else:
            assert np.array_equal(r, first[0]) and np.array_equal(pk, first[1])
        c.close()
    r, pk = first
    idx = list(range(0, P, 256))
    want = oracle.score_packed(snap, batch.select(idx))
    assert np.array_equal(r["best_row"][idx], want.best_row)
    assert np.array_equal(r["score"][idx].view(np.uint32), want.score.view(np.uint32))
    assert np.array_equal(r["fit_count"][idx], want.fit_count)
    pb = batch.pick_begin
    got = np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1)
    got = np.concatenate([got[int(pb[batch.pod_req_range(p)[0]]):int(pb[batch.pod_req_range(p)[1]])] for p in idx])
    assert np.array_equal(got, want.picks)
    assert (r["best_row"] >= 0).mean() > 0.8 and r["best_row"].max() > 875_000  # winners come from the last shard too


def test_candidate_block_is_clean_between_passes(built):
    """K3 clears the candidate keys/counts it reads and writes every pick itself, so a pass needs no
    memset -- unless an earlier pass was scored and never resolved, possibly for MORE pods than the next one"""
    snap, batch = v.workload.make_config("cfg3", 3000, 200)
    want = oracle.score_packed(snap, batch)
    small = batch.select(list(range(10)))
    want_small = oracle.score_packed(snap, small)
    with v.Context(max_nodes=3000, max_pods=256, max_reqs=512, max_picks=1024) as ctx:
        ctx.upload(snap)
        pods, reqs = v.capi.pack_batch(batch, snap.type_names)
        sp, sr = v.capi.pack_batch(small, snap.type_names)
        for _ in range(3):  # back-to-back passes, no clearing in between
            assert_same(ctx.score(batch), want, "repeat")
        ctx.pods_upload(pods, reqs)
        ctx.score_resident()  # 200 dirty slots, never resolved
        ctx.score_resident()  # ... and scored again: counts must not double
        ctx.resolve_resident()
        res, picks = ctx.results_download()
        assert np.array_equal(res["fit_count"], want.fit_count) and np.array_equal(res["best_row"], want.best_row)
        ctx.pods_upload(pods, reqs)
        ctx.score_resident()  # dirty again, abandoned for a SMALLER batch
        assert_same(ctx.score(small), want_small, "small after abandoned pass")
        assert_same(ctx.score(batch), want, "big after small")  # slots 10..199 must have been cleared too
        # a pod that fits nowhere gets zero picks even when the slots held another pod's picks before
        nofit = v.PodBatch.from_pods([{"containers": [[(2, 10**9, 101, 0)]]}] * 3)
        got = ctx.score(nofit)
        assert (got.best_row == -1).all() and not got.picks.any()


def test_pinned_planes_take_the_zero_copy_upload(built):
    """pinned, aligned host planes are pulled over PCIe by a kernel; pageable ones by cudaMemcpy:
    same resident snapshot either way"""
    import torch
    snap, batch = v.workload.make_config("cfg3", 3001, 64)  # odd size: exercises the int32 tail
    pin = torch.from_numpy(snap.planes).pin_memory()
    pnd = torch.from_numpy(snap.ndev).pin_memory()
    with v.Context(max_nodes=3001, max_pods=64) as ctx:
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, [pin[f].data_ptr() for f in range(8)], pnd.data_ptr(), device=False)
        ctx.type_names = snap.type_names
        planes, ndev = ctx.download_snapshot()
        assert np.array_equal(planes, snap.planes) and np.array_equal(ndev, snap.ndev)
        assert_same(ctx.score(batch), oracle.score_packed(snap, batch), "pinned upload")
    with v.Context(max_nodes=3001, max_pods=64, tuning={"zero_copy_upload": 0}) as ctx:
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, [pin[f].data_ptr() for f in range(8)], pnd.data_ptr(), device=False)
        planes, _ = ctx.download_snapshot()
        assert np.array_equal(planes, snap.planes)


def test_async_uploads_of_snapshots_with_other_properties(built):
    """asynchronous uploads one after another of snapshots that allow different things -- more types (other table
    rows), a Totalmem the bit tables cannot take (per-pair kernels only) -- each scored right after it"""
    import torch
    n, p = 6001, 1400
    snap_a, batch = v.workload.make_config("cfg3", n, p)
    snap_b = v.workload.make_stress_snapshot(n)          # 48 types instead of a handful
    snap_c = v.Snapshot(snap_a.planes.copy(), snap_a.ndev.copy(), snap_a.type_names)
    snap_c.planes[3][11, 0] = 1 << 30                      # Totalmem beyond the sliced tables: per-pair kernels only
    with v.Context(max_nodes=n, max_pods=p) as ctx:
        for snap in (snap_a, snap_b, snap_a, snap_c, snap_a, snap_a):
            want = oracle.score_packed(snap, batch)
            pods, reqs = v.capi.pack_batch(batch, snap.type_names)
            pin = torch.from_numpy(snap.planes.copy()).pin_memory()
            pnd = torch.from_numpy(snap.ndev.copy()).pin_memory()
            ctx.set_type_flags([v.capi.host_type_flags(t) for t in snap.type_names])
            ctx.upload_ptrs(n, snap.dmax, [pin[f].data_ptr() for f in range(8)], pnd.data_ptr(), device=False, asynchronous=True)
            ctx.pods_upload(pods, reqs)
            ctx.score_resident()
            ctx.resolve_resident()
            r, pk = ctx.results_download()
            assert np.array_equal(r["best_row"], want.best_row) and np.array_equal(r["fit_count"], want.fit_count)
            assert np.array_equal(r["score"].view(np.uint32), want.score.view(np.uint32))
            assert np.array_equal(np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1), want.picks)


def test_async_snapshot_upload(built):
    """vgs_snapshot_upload_async returns with the copy queued; the next call that needs the snapshot waits
    for it and gets its range errors; results equal the blocking upload's, pass after pass"""
    import torch
    snap, batch = v.workload.make_config("cfg3", 6001, 700)
    want = oracle.score_packed(snap, batch)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    pin = torch.from_numpy(snap.planes.copy()).pin_memory()
    pnd = torch.from_numpy(snap.ndev.copy()).pin_memory()
    ptrs = [pin[f].data_ptr() for f in range(8)]
    with v.Context(max_nodes=6001, max_pods=700) as ctx:
        ctx.type_names = snap.type_names
        for _ in range(3):  # the bench's end-to-end step
            ctx.upload_ptrs(snap.n_nodes, snap.dmax, ptrs, pnd.data_ptr(), device=False, asynchronous=True)
            ctx.pods_upload(pods, reqs)
            ctx.score_resident()
            ctx.resolve_resident()
            r, pk = ctx.results_download()
            assert np.array_equal(r["best_row"], want.best_row) and np.array_equal(r["fit_count"], want.fit_count)
            assert np.array_equal(np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1), want.picks)
        # completed by other entry points too
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, ptrs, pnd.data_ptr(), device=False, asynchronous=True)
        planes, ndev = ctx.download_snapshot()
        assert np.array_equal(planes, snap.planes) and np.array_equal(ndev, snap.ndev)
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, ptrs, pnd.data_ptr(), device=False, asynchronous=True)
        ctx.sync()
        assert_same(ctx.score(batch), want, "after sync")
        # a bad snapshot: the error belongs to the upload and comes out of the next call
        pin[v.model.P_COUNT][5, 3] = 1 << 21
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, ptrs, pnd.data_ptr(), device=False, asynchronous=True)
        ctx.pods_upload(pods, reqs)  # depends on the pods alone: goes up beside the copy, does not wait for it
        with pytest.raises(v.VgsError) as e:
            ctx.score_resident()  # the first call that needs the snapshot
        assert e.value.code == v.capi.VGS_ERANGE
        with pytest.raises(v.VgsError) as e:
            ctx.score_resident()  # still no snapshot
        assert e.value.code == v.capi.VGS_ESTATE
        pin[v.model.P_COUNT][5, 3] = int(snap.planes[v.model.P_COUNT][5, 3])
        ctx.upload_ptrs(snap.n_nodes, snap.dmax, ptrs, pnd.data_ptr(), device=False, asynchronous=True)
        assert_same(ctx.score(batch), want, "recovered")


def test_refresh_and_phased_calls(built):
    """the phased entry points (pods_upload / score_resident / resolve_resident / results_download) and
    vgs_snapshot_refresh give the same answer as the one-shot vgs_score_batch"""
    snap, batch = v.workload.make_config("cfg3", 4000, 700)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    with v.Context(max_nodes=4000, max_pods=700) as ctx:
        ctx.upload(snap)
        want_r, want_p = ctx.score_packed(pods, reqs)
        ctx.pods_upload(pods, reqs)
        for _ in range(2):
            ctx.refresh()
            ctx.score_resident()
            ctx.resolve_resident()
            got_r, got_p = ctx.results_download()
            assert np.array_equal(got_r, want_r) and np.array_equal(got_p, want_p)
        t = ctx.timing()
        assert t.score_ms > 0 and t.derive_ms > 0 and t.kernel_launches > 0


def _oracle_drain(snap, batch):
    """P consecutive oracle Filters with the winners' usage folded in between (test-side restatement
    of scheduler.go:372-394 + :280-297 on the packed planes)"""
    planes = snap.planes.copy()
    pb = batch.pick_begin
    rows, scores, counts, picks = [], [], [], np.zeros((int(batch.nums.sum()), 3), np.int32)
    for p in range(batch.n_pods):
        one = batch.select([p])
        r = oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), one, threads=4)
        rows.append(int(r.best_row[0])), scores.append(r.score[0]), counts.append(int(r.fit_count[0]))
        if r.best_row[0] >= 0:
            r0, r1 = batch.pod_req_range(p)
            picks[int(pb[r0]):int(pb[r1])] = r.picks
            for s, m, c in r.picks:
                planes[0, rows[-1], s] += 1
                planes[2, rows[-1], s] += m
                planes[4, rows[-1], s] += c
    return np.array(rows, np.int32), np.array(scores, np.float32), np.array(counts, np.int32), picks, planes


@pytest.mark.parametrize("cfg,n,p", [("cfg3", 600, 150), ("cfg4", 400, 120), ("cfg2", 64, 200)])
def test_drain_equals_sequential_filters(built, cfg, n, p):
    """vgs_drain_batch == P consecutive Filters with addPod in between: same winners, scores, picks,
    and the resident snapshot ends up equal to the rebuilt one"""
    snap, batch = v.workload.make_config(cfg, n, p)
    rows, scores, counts, picks, planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
        assert np.array_equal(got.best_row, rows)
        assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32))
        assert np.array_equal(got.fit_count, counts)
        assert np.array_equal(got.picks, picks)
        assert np.array_equal(after, planes)
        # state is carried: the drained snapshot answers a fresh batch like the rebuilt one
        again = ctx.score(batch)
    assert_same(again, oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), batch), "after drain")
    assert (rows >= 0).sum() > 0 and len(set(rows[rows >= 0].tolist())) < (rows >= 0).sum()  # nodes get reused


@pytest.mark.parametrize("chunk", [0, 7, 33, -1])
@pytest.mark.parametrize("cfg,n,p", [("cfg3", 3000, 500), ("cfg4", 700, 150), ("cfg2", 64, 200), ("stress", 900, 130)])
def test_drain_chunked_equals_per_pod_chain(built, cfg, n, p, chunk):
    """the batch-scale drain (chunks scored once, placed in order from the touched rows) against the oracle's
    P sequential Filters, for several chunk lengths and for the per-pod chain (drain_chunk = -1)"""
    snap, batch = v.workload.make_config(cfg, n, p)
    rows, scores, counts, picks, planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=p, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum()), tuning={"drain_chunk": chunk}) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        plan = ctx.plan()
        after, _ = ctx.download_snapshot()
        assert np.array_equal(got.best_row, rows)
        assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32))
        assert np.array_equal(got.fit_count, counts)
        assert np.array_equal(got.picks, picks)
        assert np.array_equal(after, planes)
        want_chunks = 0 if chunk < 0 else -(-p // (chunk or 64))
        assert plan.drain_chunks == want_chunks
        # the drained snapshot answers a fresh batch like the rebuilt one, and a second drain carries on from it
        again = ctx.score(batch)
        assert_same(again, oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), batch), "after drain")
        more = ctx.drain(batch)
    rows2, scores2, counts2, picks2, _ = _oracle_drain(v.Snapshot(planes, snap.ndev, snap.type_names), batch)
    assert np.array_equal(more.best_row, rows2) and np.array_equal(more.fit_count, counts2)
    assert np.array_equal(more.score.view(np.uint32), scores2.view(np.uint32)) and np.array_equal(more.picks, picks2)


@pytest.mark.parametrize("pods_of", ["mixed", "multi"])
def test_drain_chunked_wide_nodes(built, pods_of):
    """nodes with 12 and with ragged device counts (the 12-wide tiles): chunked drain == sequential oracle Filters"""
    n, p = 1500, 260
    snap = v.workload.make_snapshot(n, seed=991, dmax=12, count=6)
    rng = np.random.default_rng(5)
    snap.ndev[:] = rng.integers(5, 13, n)
    for f in range(8):
        snap.planes[f][np.arange(12)[None, :] >= snap.ndev[:, None]] = 0
    batch = (v.workload.pods_mixed if pods_of == "mixed" else v.workload.pods_multi)(p, 77)
    rows, scores, counts, picks, planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_devices=12, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        plan = ctx.plan()
        after, _ = ctx.download_snapshot()
    assert plan.drain_chunks == -(-p // 64)
    assert np.array_equal(got.best_row, rows) and np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32)) and np.array_equal(got.picks, picks)
    assert np.array_equal(after, planes)


@pytest.mark.parametrize("cfg,p", [("cfg3", 4096), ("cfg4", 1024)])
def test_drain_full_size(built, cfg, p):
    """the drain at BASELINE's node count: the chunked path and the per-pod chain against each other for every pod,
    and against sequential oracle Filters for the first 160 pods (two and a half chunks: carried rows included)"""
    snap, batch = v.workload.make_config(cfg, 100_000, p)
    out = {}
    for chunk in (0, -1):
        with v.Context(max_nodes=snap.n_nodes, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum()),
                       tuning={"drain_chunk": chunk}) as ctx:
            ctx.upload(snap)
            out[chunk] = ctx.drain(batch)
            if chunk == 0:
                plan = ctx.plan()
                after, _ = ctx.download_snapshot()
            else:
                after_chain, _ = ctx.download_snapshot()
    a, b = out[0], out[-1]
    assert plan.drain_chunks == p // 64
    assert np.array_equal(a.best_row, b.best_row) and np.array_equal(a.fit_count, b.fit_count)
    assert np.array_equal(a.score.view(np.uint32), b.score.view(np.uint32)) and np.array_equal(a.picks, b.picks)
    assert np.array_equal(after, after_chain)
    head = batch.select(list(range(160)))
    rows, scores, counts, picks, _ = _oracle_drain(snap, head)
    assert np.array_equal(a.best_row[:160], rows) and np.array_equal(a.fit_count[:160], counts)
    assert np.array_equal(a.score[:160].view(np.uint32), scores.view(np.uint32))
    assert np.array_equal(a.picks[:len(picks)], picks)
    assert (a.best_row >= 0).sum() > 0.9 * p


@pytest.mark.parametrize("chunk", [0, 7, -1])
def test_drain_hand_derived_sequence(built, chunk):
    """A drain worked out by hand from score.go / scheduler.go, no oracle involved.  11 nodes of two cards (Count 2,
    1000 MB, 100 cores, one numa, one type); node 0 has card 0 in use once (300 MB, 20 cores); 40 identical pods
    ask for one card with 400 MB and 30 cores.  Per Filter the cards are sorted ascending by Count-Used
    (score.go:45-50, stable for <= 12) and walked from the END (:92), so the emptier card is met first, and
    score = Count/(Count-Used) + (2 - 1) with Used as it was before the take (:168-170,180).  An empty node gives
    2/2 + 1 = 2 through card 1; so does node 0 (card 1 is its emptier one).  Ties go to the highest row (our canonical
    rule), so node 10 takes the first pod on card 1; then its card 0 is the emptier one (2 again, slot 0); then both
    cards have one slot left, the stable sort keeps registration order and the walk meets card 1 first: 2/1 + 1 = 3,
    then card 0 likewise: 3.  Node 10 is full (Used = Count on both cards) and the next four pods do the same to
    node 9, and so on down to node 1.  Nodes that fit pod 4k+j: rows 0..10-k."""
    n = 11
    planes = np.zeros((8, n, 2), np.int32)
    USED, COUNT, USEDMEM, TOTALMEM, USEDCORES, TOTALCORE = 0, 1, 2, 3, 4, 5
    planes[COUNT], planes[TOTALMEM], planes[TOTALCORE] = 2, 1000, 100
    planes[USED, 0, 0], planes[USEDMEM, 0, 0], planes[USEDCORES, 0, 0] = 1, 300, 20
    snap = v.Snapshot(planes, np.full(n, 2, np.int32), ["NVIDIA-A100"])
    batch = v.PodBatch.from_pods([{"containers": [[(1, 400, 101, 30, "NVIDIA")]]} for _ in range(40)])
    want_rows = np.repeat(np.arange(10, 0, -1), 4).astype(np.int32)
    want_scores = np.tile(np.array([2.0, 2.0, 3.0, 3.0], np.float32), 10)
    want_slots = np.tile(np.array([1, 0, 1, 0], np.int32), 10)
    want_fit = np.repeat(np.arange(11, 1, -1), 4).astype(np.int32)
    with v.Context(max_nodes=n, max_devices=2, max_pods=40, tuning={"drain_chunk": chunk}) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
        assert ctx.plan().drain_chunks == {0: 1, 7: 6, -1: 0}[chunk]
    assert np.array_equal(got.best_row, want_rows)
    assert np.array_equal(got.score, want_scores)
    assert np.array_equal(got.fit_count, want_fit)
    assert np.array_equal(got.picks, np.stack([want_slots, np.full(40, 400), np.full(40, 30)], axis=1))
    end = planes.copy()
    end[USED, 1:], end[USEDMEM, 1:], end[USEDCORES, 1:] = 2, 800, 60
    assert np.array_equal(after, end)


@pytest.mark.parametrize("dmax,seed,n,p,chunk", [(8, 31, 260, 140, 0), (8, 32, 90, 200, 0), (12, 33, 400, 100, 0),
                                                  (4, 34, 150, 130, 9), (1, 35, 500, 70, 0), (7, 36, 40, 96, 33),
                                                  (8, 37, 3000, 160, 0)])
def test_drain_random_ragged(built, dmax, seed, n, p, chunk):
    """the chunked drain on random ragged snapshots (odd Count / Used / Totalcore values, negative numa, 0..dmax devices)
    with random multi-container pods (numa-bind, type filters, requests that cannot fit): few nodes, so rows are hit
    again and again, lists run dry, carried rows are hit by the next chunk"""
    rng = np.random.default_rng(seed)
    snap = _random_snapshot(rng, n, dmax)
    batch = _random_pods(rng, p, dmax)
    rows, scores, counts, picks, planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_devices=dmax, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum()),
                   tuning={"drain_chunk": chunk}) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
        assert ctx.plan().drain_chunks > 0
    assert np.array_equal(got.best_row, rows)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32))
    assert np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.picks, picks)
    assert np.array_equal(after, planes)


@pytest.mark.parametrize("seed,n,d,p", [(2050, 2500, 2, 190), (2051, 200, 4, 199), (2055, 700, 4, 142)])
def test_drain_tight_cluster_rescans(built, seed, n, d, p):
    """nodes that take one or two pods each, pods that differ little: rows fill up quickly, the candidate lists run dry
    chunk after chunk (dozens of re-scans, over rows carried in from the chunk before) -- still sequential Filters"""
    rng = np.random.default_rng(seed)
    planes = np.zeros((8, n, d), np.int32)
    planes[1] = rng.integers(1, 3, (n, d))
    planes[0] = np.where(rng.random((n, d)) < 0.8, planes[1], planes[1] - 1)
    planes[3], planes[5] = 16000, 100
    planes[2] = 1000 * planes[0]
    planes[6] = rng.integers(0, 2, (n, d))
    snap = v.Snapshot(planes, np.full(n, d, np.int32), ["NVIDIA-A100"])
    batch = v.PodBatch.from_pods([{"containers": [[(int(rng.choice([1, 1, 1, 2])), int(rng.choice([500, 1000])), 101,
                                                    int(rng.choice([0, 10])), "NVIDIA")]]} for _ in range(p)])
    rows, scores, counts, picks, want_planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_devices=d, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
        assert ctx.plan().drain_rescans >= 3  # (a re-scan refills the list of the pod and of its replicas: far fewer than pods)
    assert np.array_equal(got.best_row, rows) and np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32)) and np.array_equal(got.picks, picks)
    assert np.array_equal(after, want_planes)


def test_drain_gives_the_chunks_up_when_rescans_pile_up(built):
    """a tight cluster at a size where a re-scan costs more than placing a pod the plain way (40k nodes: 8 re-scans
    per chunk allowed): the chunked drain stops at the pod that would pass that, the rest goes pod by pod
    (vgs_plan.drain_chain_pods) -- same answers as sequential Filters throughout"""
    rng = np.random.default_rng(77)
    n, d, p = 40_000, 2, 200
    planes = np.zeros((8, n, d), np.int32)
    planes[1] = rng.integers(1, 3, (n, d))
    planes[0] = np.where(rng.random((n, d)) < 0.8, planes[1], planes[1] - 1)
    planes[3], planes[5] = 16000, 100
    planes[2] = 1000 * planes[0]
    snap = v.Snapshot(planes, np.full(n, d, np.int32), ["NVIDIA-A100"])
    batch = v.PodBatch.from_pods([{"containers": [[(int(rng.choice([1, 1, 1, 2])), int(rng.choice([500, 1000])), 101,
                                                    int(rng.choice([0, 10])), "NVIDIA")]]} for _ in range(p)])
    rows, scores, counts, picks, want_planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_devices=d, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        plan = ctx.plan()
        after, _ = ctx.download_snapshot()
        again = ctx.drain(batch)  # and a second drain on top of it starts in chunks again
    assert 0 < plan.drain_chain_pods < p and plan.drain_rescans <= 8 * plan.drain_chunks
    assert np.array_equal(got.best_row, rows) and np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32)) and np.array_equal(got.picks, picks)
    assert np.array_equal(after, want_planes)
    rows2, scores2, counts2, picks2, _ = _oracle_drain(v.Snapshot(want_planes, snap.ndev, snap.type_names), batch)
    assert np.array_equal(again.best_row, rows2) and np.array_equal(again.fit_count, counts2) and np.array_equal(again.picks, picks2)


@pytest.mark.parametrize("req", [(1, 3000, 101, 30), (2, 8000, 101, 0), (1, 0, 50, 100)])
def test_drain_replicas_share_the_refilled_list(built, req):
    """300 identical pods (one Deployment's replicas) and a few others in between: the pods in front use up the
    candidates all of them share; the re-scan that follows refills the list of its pod AND of the replicas behind it"""
    snap, other = v.workload.make_config("cfg3", 20_000, 12)
    pods = [{"containers": [[req + ("NVIDIA",)]]} for _ in range(300)]
    batch = v.PodBatch.from_pods(pods)
    rows, scores, counts, picks, want_planes = _oracle_drain(snap, batch)
    with v.Context(max_nodes=snap.n_nodes, max_pods=300, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        plan = ctx.plan()
        after, _ = ctx.download_snapshot()
    assert plan.drain_chain_pods == 0 and 0 < plan.drain_rescans < 40
    assert np.array_equal(got.best_row, rows) and np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32)) and np.array_equal(got.picks, picks)
    assert np.array_equal(after, want_planes)


def test_drain_candidate_list_runs_dry(built):
    """identical pods on nodes that take exactly one of them: inside a chunk every candidate of a pod's list is
    consumed by the pods before it, and no touched row fits any more -- the commit kernel has to scan the walk
    copy (vgs_plan.drain_rescans), and still places every pod where sequential Filters would"""
    n, p = 500, 120
    planes = np.zeros((8, n, 4), np.int32)
    USED, COUNT, TOTALMEM, TOTALCORE = 0, 1, 3, 5  # include/vgpuscore.h VGS_PLANE_*
    planes[COUNT] = 1
    planes[TOTALMEM] = 16000
    planes[TOTALCORE] = 100
    ndev = np.full(n, 4, np.int32)
    planes[USED, :, 1:] = 1  # one free card per node
    snap = v.Snapshot(planes, ndev, ["NVIDIA-A100"])
    batch = v.PodBatch.from_pods([{"containers": [[(1, 1000, 0, 10, "NVIDIA")]]} for _ in range(p)])
    rows, scores, counts, picks, after_want = _oracle_drain(snap, batch)
    assert len(set(rows.tolist())) == p and rows.max() == n - 1  # one node each, highest rows first
    with v.Context(max_nodes=n, max_devices=4, max_pods=p) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        plan = ctx.plan()
        after, _ = ctx.download_snapshot()
    assert np.array_equal(got.best_row, rows) and np.array_equal(got.fit_count, counts)
    assert np.array_equal(got.score.view(np.uint32), scores.view(np.uint32)) and np.array_equal(got.picks, picks)
    assert np.array_equal(after, after_want)
    assert plan.drain_chunks == 2 and plan.drain_rescans > 0


def test_drain_does_not_keep_the_batch_for_a_score_pass(built):
    """the drain builds none of the score pass's class tables for its batch: vgs_score_resident after it is a sequence
    error until the pods are uploaded again, and a score of the same batch afterwards sees the drained snapshot"""
    snap, batch = v.workload.make_config("cfg3", 3000, 700)
    rows, scores, counts, picks, after_want = _oracle_drain(snap, batch)
    with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum()), tuning={"slice_min_nodes": 0}) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        assert np.array_equal(got.best_row, rows) and np.array_equal(got.picks, picks)
        with pytest.raises(v.VgsError) as e:
            ctx.score_resident()
        assert e.value.code == v.capi.VGS_ESTATE
        res = ctx.score(batch)  # uploads the pods (with their tables) again
    want = oracle.score_packed(v.Snapshot(after_want, snap.ndev, snap.type_names), batch, full=False)
    assert_same(res, want, "score after drain")


@pytest.mark.parametrize("seed,n,dmax", [(1, 3000, 8), (2, 33, 7), (3, 1, 1), (4, 517, 12), (5, 64, 4), (6, 5000, 2)])
def test_narrow_planes_upload_like_int32_planes(built, seed, n, dmax):
    """vgs_snapshot_upload_narrow: planes of int8 / int16 / int32 elements (every plane in the narrowest type its values
    fit: numa -1 and Totalcore 200 of the random snapshots exercise sign extension and the 2-byte path, odd sizes the
    tails) leave exactly the device planes the int32 upload leaves, from pageable memory (staging copy) and from pinned
    memory (read over PCIe by the widening kernel), synchronously and queued -- and score the same"""
    import torch
    rng = np.random.default_rng(900 + seed)
    snap = _random_snapshot(rng, n, dmax)
    if seed == 6:
        snap.planes[0][7, 1] = 40_000  # Used beyond int16: that plane alone stays int32
    batch = _random_pods(rng, 60, dmax)
    want = oracle.score_packed(snap, batch)
    arrs = v.capi.narrow_planes(snap)
    widths = [a.dtype.itemsize for a in arrs]
    assert min(widths) == 1 and (seed != 6 or widths[0] == 4)
    with v.Context(max_nodes=n, max_devices=dmax, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum())) as ctx:
        ctx.upload_narrow(snap)  # pageable numpy arrays
        planes, ndev = ctx.download_snapshot()
        assert np.array_equal(planes, snap.planes) and np.array_equal(ndev, snap.ndev)
        assert_same(ctx.score(batch), want, "narrow, pageable")
        pinned = [torch.from_numpy(a.reshape(-1).copy()).pin_memory() for a in arrs]
        for asynchronous in (False, True):
            ctx.upload(v.Snapshot(np.zeros_like(snap.planes), np.zeros_like(snap.ndev), snap.type_names))  # wipe
            ctx.upload_narrow_ptrs(n, dmax, [t.data_ptr() for t in pinned[:8]], pinned[8].data_ptr(), widths,
                                   asynchronous=asynchronous)
            planes, ndev = ctx.download_snapshot()
            assert np.array_equal(planes, snap.planes) and np.array_equal(ndev, snap.ndev), asynchronous
            assert_same(ctx.score(batch), want, f"narrow, pinned, async={asynchronous}")


def test_narrow_upload_of_the_usage_planes_alone(built):
    """planes left out (NULL) keep what is resident: a Filter whose node registry has not changed uploads Used / Usedmem /
    Usedcores only (what getNodesUsage recomputes from the pod list, scheduler.go:280-297); without a resident snapshot
    of the same shape that is a sequence error"""
    rng = np.random.default_rng(4242)
    snap = _random_snapshot(rng, 2100, 8)
    batch = _random_pods(rng, 80, 8)
    later = v.Snapshot(snap.planes.copy(), snap.ndev, snap.type_names)
    later.planes[0] = np.minimum(snap.planes[1], snap.planes[0] + rng.integers(0, 2, snap.planes[0].shape))  # more in use
    later.planes[2] = np.minimum(snap.planes[3], snap.planes[2] + 100 * (later.planes[0] - snap.planes[0]))
    later.planes[4] = np.minimum(snap.planes[5], snap.planes[4] + 5 * (later.planes[0] - snap.planes[0]))
    arrs = v.capi.narrow_planes(later)
    widths = [a.dtype.itemsize for a in arrs]
    usage = [arrs[f].ctypes.data if f in (0, 2, 4) else None for f in range(8)]
    with v.Context(max_nodes=snap.n_nodes, max_devices=8, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum())) as ctx:
        with pytest.raises(v.VgsError) as e:
            ctx.upload_narrow_ptrs(snap.n_nodes, 8, usage, None, widths)  # nothing resident
        assert e.value.code == v.capi.VGS_ESTATE
        ctx.upload(snap)
        assert_same(ctx.score(batch), oracle.score_packed(snap, batch), "before")
        for asynchronous in (False, True):
            ctx.upload_narrow_ptrs(snap.n_nodes, 8, usage, None, widths, asynchronous=asynchronous)
            planes, ndev = ctx.download_snapshot()
            assert np.array_equal(planes, later.planes) and np.array_equal(ndev, snap.ndev)
            assert_same(ctx.score(batch), oracle.score_packed(later, batch), "usage planes replaced")
        with pytest.raises(v.VgsError) as e:
            ctx.upload_narrow_ptrs(snap.n_nodes - 1, 8, usage, None, widths)  # another shape
        assert e.value.code == v.capi.VGS_ESTATE


def test_narrow_upload_errors(built):
    """a width other than 1, 2, 4 is VGS_EINVAL; a value out of range in a narrow plane is reported like in an int32 one
    (queued upload: by the next call that needs the snapshot)"""
    snap, batch = v.workload.make_config("cfg3", 200, 8)
    arrs = v.capi.narrow_planes(snap)
    widths = [a.dtype.itemsize for a in arrs]
    with v.Context(max_nodes=200, max_devices=snap.dmax, max_pods=8) as ctx:
        with pytest.raises(v.VgsError) as e:
            ctx.upload_narrow_ptrs(200, snap.dmax, [a.ctypes.data for a in arrs[:8]], arrs[8].ctypes.data, [3] + widths[1:])
        assert e.value.code == v.capi.VGS_EINVAL
        bad = [a.copy() for a in arrs]
        bad[8][5] = snap.dmax + 1  # len(node.Devices) beyond dmax
        ctx.set_type_flags([v.capi.host_type_flags(t) for t in snap.type_names])
        ctx.type_names = snap.type_names
        ctx.upload_narrow_ptrs(200, snap.dmax, [a.ctypes.data for a in bad[:8]], bad[8].ctypes.data, widths, asynchronous=True)
        with pytest.raises(v.VgsError):
            ctx.score(batch)
        ctx.upload_narrow(snap)
        assert_same(ctx.score(batch), oracle.score_packed(snap, batch), "after the bad upload")


def test_contexts_give_their_memory_back(built):
    """vgs_destroy releases everything a context took -- the planes, the arena, the verify matrices, the drain's score
    matrices (allocated at the first chunked drain): the device's free memory after 12 contexts that scored, drained
    and were closed is where it was after the first one (which pays for the lazily loaded kernels)"""
    import torch
    snap, batch = v.workload.make_config("cfg3", 20_000, 700)

    def cycle(verify):
        with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=batch.n_pods, max_reqs=batch.n_reqs,
                       max_picks=int(batch.nums.sum()), verify=verify) as ctx:
            ctx.upload(snap)
            ctx.score(batch)
            ctx.score(batch.select([3]))
            ctx.drain(batch)
            assert ctx.plan().drain_chunks > 0

    cycle(True)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for i in range(12):
        cycle(bool(i & 1))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 < (4 << 20), (free0, free1)  # one context of this size holds > 60 MB


def test_empty_batch_and_concurrent_callers(built):
    """n_pods = 0 is a no-op; calls on one context from several host threads are serialised by the
    library (net/http runs each /filter in its own goroutine) and all get the right answer"""
    import threading
    snap, batch = v.workload.make_config("cfg3", 2000, 96)
    want = oracle.score_packed(snap, batch)
    with v.Context(max_nodes=2000, max_pods=96) as ctx:
        ctx.upload(snap)
        r, p = ctx.score_packed(np.zeros(0, v.capi.POD_DTYPE), np.zeros(0, v.capi.REQ_DTYPE))
        assert len(r) == 0 and len(p) == 0
        pods, reqs = v.capi.pack_batch(batch, snap.type_names)
        out, errs = {}, []

        def worker(i):
            try:
                for _ in range(5):
                    res = np.zeros(len(pods), dtype=v.capi.RESULT_DTYPE)
                    pk = np.zeros(int(reqs["nums"].sum()), dtype=v.capi.PICK_DTYPE)
                    rc = ctx.lib.vgs_score_batch(ctx.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data,
                                                 res.ctypes.data, pk.ctypes.data)
                    assert rc == 0
                    out[i] = (res, pk)
            except Exception as e:  # pragma: no cover
                errs.append(e)

        ts = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errs
        for res, pk in out.values():
            assert np.array_equal(res["best_row"], want.best_row)
            assert np.array_equal(res["score"].view(np.uint32), want.score.view(np.uint32))
            assert np.array_equal(np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], 1), want.picks)
