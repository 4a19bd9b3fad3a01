"""The C++ mirror of Scheduler.Filter (host/scheduler.hpp -> C ABI -> GPU) against the reference
flow restated in Python around the oracle: same pods, same order, same cluster.  Compared per
Filter call: chosen node, the 4pd.io/vgpu-ids-new annotation string (util.EncodePodDevices),
failed nodes; usage carried from pod to pod (addPod/delPod, scheduler.go:280-297,394)."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v
from oracle.ref import encode_pod_devices


def make_cluster(rng, n_nodes):
    types = [("NVIDIA-Tesla V100-PCIE-32GB", 32510), ("NVIDIA-NVIDIA A100-SXM4-40GB", 40960), ("NVIDIA-Tesla T4", 15360)]
    cluster = {}
    for i in range(n_nodes):
        t, mem = types[int(rng.integers(0, len(types)))]
        nd = int(rng.choice([4, 8, 8, 8]))
        cluster[f"node-{i:03d}"] = [dict(ID=f"GPU-{i:03d}-{s}", Count=int(rng.choice([2, 4, 10])), Devmem=mem, Devcore=100,
                                         Type=t, Numa=(0 if s < nd // 2 else 1)) for s in range(nd)]
    return cluster


def make_pods(rng, n):
    pods = []
    for i in range(n):
        ctrs = []
        for _ in range(int(rng.choice([1, 1, 1, 2, 3]))):
            if rng.random() < 0.12:
                ctrs.append({})  # container without a GPU request
                continue
            c = {"gpu": int(rng.choice([1, 1, 1, 2, 4]))}
            r = rng.random()
            if r < 0.45:
                c["gpumem"] = int(rng.choice([1000, 4000, 12000, 30000]))
            elif r < 0.8:
                c["gpumem-percentage"] = int(rng.choice([10, 25, 50, 100]))
            if rng.random() < 0.7:
                c["gpucores"] = int(rng.choice([0, 10, 30, 50, 100]))
            ctrs.append(c)
        annos = {}
        r = rng.random()
        if r < 0.2:
            annos["nvidia.com/use-gputype"] = str(rng.choice(["a100", "V100,T4", "t4"]))
        elif r < 0.35:
            annos["nvidia.com/nouse-gputype"] = str(rng.choice(["a100", "T4"]))
        if rng.random() < 0.2:
            annos["nvidia.com/numa-bind"] = "true"
        pods.append((f"pod-{i:03d}", ctrs, annos))
    return pods


class ReferenceFlow:
    """Filter restated around the oracle (test-side only): getNodesUsage rebuild by UUID,
    Resourcereqs defaults, calcScore via oracle R1, canonical pick, EncodePodDevices, addPod."""

    def __init__(self, cluster):
        self.cluster, self.pods = cluster, {}

    def requests(self, ctrs):
        out = []
        for c in ctrs:
            one = []  # vendor order of the C++ mirror: Cambricon, Nvidia, Hygon (Go's map order is random)
            if c.get("mlunum", 0) > 0:  # cambricon/device.go:106-137
                one.append((c["mlunum"], c.get("mlumem", 0), 0, 0, "MLU"))
            if c.get("gpu", 0) > 0:
                mem, pct = c.get("gpumem", 0), c.get("gpumem-percentage", 101)
                if pct == 101 and mem == 0:
                    pct = 100  # DefaultMem == 0 (nvidia/device.go:147-153)
                one.append((c["gpu"], mem, pct, c.get("gpucores", 0), "NVIDIA"))
            if c.get("dcunum", 0) > 0:  # hygon/device.go:86-135
                mem = c.get("dcumem", 0)
                one.append((c["dcunum"], mem, 100 if mem == 0 else 0, c.get("dcucores", 0), "DCU"))
            out.append(one)
        return out

    def filter(self, uid, ctrs, annos, node_names):
        reqs = self.requests(ctrs)
        if sum(r[0] for c in reqs for r in c) == 0:
            return dict(node=None, ids=None, failed={}, passed=list(node_names))
        self.pods.pop(uid, None)
        failed = {n: "node unregisterd" for n in node_names if n not in self.cluster}
        order = [n for n in dict.fromkeys(node_names) if n in self.cluster]
        nodes = []
        for n in order:
            devs = [dict(Used=0, Count=d["Count"], Usedmem=0, Totalmem=d["Devmem"], Usedcores=0, Totalcore=d["Devcore"],
                         Numa=d["Numa"], Type=d["Type"]) for d in self.cluster[n]]
            for node, pd in self.pods.values():
                if node != n:
                    continue
                for cd in pd:
                    for uuid, _, mem, cores in cd:
                        for d, reg in zip(devs, self.cluster[n]):
                            if reg["ID"] == uuid:
                                d["Used"] += 1
                                d["Usedmem"] += mem
                                d["Usedcores"] += cores
            nodes.append(devs)
        snap = v.Snapshot.from_nodes(nodes, dmax=8)
        batch = v.PodBatch.from_pods([{"containers": reqs, "annos": annos}])
        r = oracle.score_faithful(snap, batch)
        if r.best_row[0] < 0:
            return dict(node=None, ids=None, failed=failed, passed=[])
        node = order[int(r.best_row[0])]
        pb, pd = batch.pick_begin, []
        for c in range(len(reqs)):
            cd = []
            for q in range(int(batch.req_begin[c]), int(batch.req_begin[c + 1])):
                for j in range(int(batch.nums[q])):
                    s, m, k = (int(x) for x in r.picks[int(pb[q]) + j])
                    cd.append((self.cluster[node][s]["ID"], batch.req_type[q], m, k))
            pd.append(cd)
        self.pods[uid] = (node, pd)
        return dict(node=node, ids=encode_pod_devices(pd), failed={}, passed=[node])


@pytest.mark.gpu
@pytest.mark.parametrize("incremental", [False, True])
def test_filter_sequence_matches_reference_flow(built, incremental):
    rng = np.random.default_rng(2024)
    cluster = make_cluster(rng, 48)
    pods = make_pods(rng, 90)
    ref = ReferenceFlow(cluster)
    sch = v.hostlib.Scheduler(max_nodes=64, incremental=incremental)
    for n, devs in cluster.items():
        sch.add_node(n, devs)
    names = list(cluster) + ["ghost-node"]  # an unregistered node must come back in FailedNodes
    placed = 0
    for i, (uid, ctrs, annos) in enumerate(pods):
        if i % 17 == 16:  # a scheduled pod terminates
            victim = pods[i - 9][0]
            sch.del_pod(victim)
            ref.pods.pop(victim, None)
        if i == 40:  # the candidate list changes -> rows change -> full upload even when incremental
            names = names[5:] + names[:5]
        got = sch.filter(uid, ctrs, annos, names)
        want = ref.filter(uid, ctrs, annos, names)
        assert got["node"] == want["node"], (uid, ctrs, annos, got, want)
        assert got["ids"] == want["ids"], (uid, got, want)
        assert got["passed"] == want["passed"], uid
        assert got["failed"] == want["failed"], uid
        placed += got["node"] is not None
    assert 30 < placed < len(pods)
    sch.close()


@pytest.mark.gpu
def test_incremental_filter_skips_the_rebuild_at_100k_nodes(built):
    """SURVEY 8(f) rank 1 on the host side: with the snapshot resident, a Filter must not redo getNodesUsage
    (scheduler.go:249-310).  100k registered nodes, a few hundred scheduled pods: the first Filter uploads, every
    later one takes the resident path (deltas only) and its host time is a small fraction of the rebuild's; both
    modes place the pods on the same nodes with the same devices."""
    N = 100_000
    rng = np.random.default_rng(5)
    names = [f"node-{i:06d}" for i in range(N)]
    devs = lambda i: [dict(ID=f"GPU-{i}-{s}", Count=10, Devmem=32768 if i % 3 else 16384, Devcore=100,
                           Type="NVIDIA-Tesla V100-PCIE-32GB" if i % 3 else "NVIDIA-Tesla T4", Numa=s // 4) for s in range(8)]
    out = {}
    for incremental in (False, True):
        sch = v.hostlib.Scheduler(max_nodes=N, incremental=incremental)
        for i, n in enumerate(names):
            sch.add_node(n, devs(i))
        placed, times = [], []
        for k in range(12):
            got = sch.filter(f"pod-{k}", [{"gpu": 1 + k % 2, "gpumem": 3000 + 500 * (k % 5), "gpucores": 10 * (k % 4)}], {}, names)
            total_us, gpu_us, fast = sch.last_timing()
            placed.append((got["node"], got["ids"]))
            times.append((total_us, gpu_us, fast))
            assert got["node"] is not None
        sch.close()
        out[incremental] = (placed, times)
    assert out[False][0] == out[True][0]                       # same placements either way
    assert not any(f for _, _, f in out[False][1])             # rebuild mode never takes the resident path
    assert [f for _, _, f in out[True][1]] == [False] + [True] * 11
    rebuild = np.median([t - g for t, g, _ in out[False][1][1:]])      # host time outside libvgpuscore
    resident = np.median([t - g for t, g, _ in out[True][1][1:]])
    print(f"host us per Filter at {N} nodes: rebuild {rebuild:.0f}, resident {resident:.0f}; "
          f"total rebuild {np.median([t for t, _, _ in out[False][1][1:]]):.0f}, "
          f"resident {np.median([t for t, _, _ in out[True][1][1:]]):.0f}")
    assert resident * 10 < rebuild


@pytest.mark.gpu
def test_filter_edge_behaviour(built):
    sch = v.hostlib.Scheduler(max_nodes=8)
    sch.add_node("n0", [dict(ID=f"U{s}", Count=10, Devmem=32768, Devcore=100, Type="NVIDIA-Tesla V100", Numa=0) for s in range(8)])
    # no vGPU requested: every node passes untouched (scheduler.go:357-370)
    r = sch.filter("p0", [{}], {}, ["n0", "nX"])
    assert r == dict(node=None, ids=None, failed={}, passed=["n0", "nX"])
    # KAT 1 of SURVEY 8(c) through the whole host path
    r = sch.filter("p1", [{"gpu": 1, "gpumem": 3000, "gpucores": 30}], {}, ["n0"])
    assert r["node"] == "n0" and r["ids"] == "U7,NVIDIA,3000,30:"
    # same pod again: Filter first drops its previous placement (scheduler.go:372)
    r = sch.filter("p1", [{"gpu": 1, "gpumem": 3000, "gpucores": 30}], {}, ["n0"])
    assert r["ids"] == "U7,NVIDIA,3000,30:"
    # nothing fits: FailedNodes only, no error (scheduler.go:381-385)
    r = sch.filter("p2", [{"gpu": 9}], {}, ["n0", "gone"])
    assert r["node"] is None and r["failed"] == {"gone": "node unregisterd"} and r["passed"] == []
    # two containers, second sees the first's usage; default memory = 100 % (device.go:147-153)
    r = sch.filter("p3", [{"gpu": 1, "gpucores": 100}, {}, {"gpu": 2, "gpumem": 1000}], {}, ["n0"])
    assert r["ids"] == "U6,NVIDIA,32768,100:;;U5,NVIDIA,1000,0:U4,NVIDIA,1000,0:"
    sch.close()


@pytest.mark.gpu
@pytest.mark.parametrize("incremental", [False, True])
def test_filter_mixed_vendors(built, incremental):
    """Cambricon MLU, Hygon DCU and NVIDIA nodes in one cluster; a sequence of Filters with usage
    carried over: an MLU370 that holds one memory-less pod is closed to the next one
    (cambricon/device.go:97-99), while pods that ask for mlumem keep sharing it"""
    rng = np.random.default_rng(99)
    kinds = [("MLU370-X8", 24576), ("MLU290", 32768), ("DCU-Z100", 16384), ("NVIDIA-Tesla T4", 15360)]
    cluster = {}
    for i in range(24):
        t, mem = kinds[i % 4]
        cluster[f"node-{i:02d}"] = [dict(ID=f"DEV-{i:02d}-{s}", Count=4, Devmem=mem, Devcore=100, Type=t, Numa=s // 2)
                                    for s in range(4)]
    sch = v.hostlib.Scheduler(max_nodes=64, incremental=incremental)
    ref = ReferenceFlow(cluster)
    for n, devs in cluster.items():
        sch.add_node(n, devs)
    names = list(cluster)
    placed = {"MLU": 0, "DCU": 0, "NVIDIA": 0}
    for i in range(150):
        r = rng.random()
        if r < 0.55:
            c = {"mlunum": int(rng.choice([1, 1, 2]))}
            if rng.random() < 0.5:
                c["mlumem"] = int(rng.choice([1000, 4000]))
            kind = "MLU"
        elif r < 0.8:
            c = {"dcunum": 1, "dcucores": int(rng.choice([0, 30, 60]))}
            if rng.random() < 0.5:
                c["dcumem"] = 2000
            kind = "DCU"
        else:
            c, kind = {"gpu": 1, "gpumem": 3000}, "NVIDIA"
        ctrs = [c] if rng.random() < 0.8 else [c, dict(c)]
        annos = {}
        if i % 6 == 0:
            annos["cambricon.com/nouse-mlutype"] = "mlu290"
        if i % 7 == 0:
            annos["hygon.com/use-dcutype"] = "z100"
        uid = f"pod-{i:03d}"
        if i % 13 == 12:
            victim = f"pod-{i - 5:03d}"
            sch.del_pod(victim)
            ref.pods.pop(victim, None)
        got = sch.filter(uid, ctrs, annos, names)
        want = ref.filter(uid, ctrs, annos, names)
        assert got["node"] == want["node"], (uid, ctrs, annos, got, want)
        assert got["ids"] == want["ids"], (uid, got, want)
        if got["node"] is not None:
            placed[kind] += 1
            assert kinds[int(got["node"][-2:]) % 4][0].startswith(kind)
    assert all(n > 5 for n in placed.values()), placed
    sch.close()


def _encode_node_devices(devs):
    """util.EncodeNodeDevices (util.go:101-108): what the device plugin writes into the register annotation"""
    return "".join(f"{d['ID']},{d['Count']},{d['Devmem']},{d['Devcore']},{d['Type']},{d['Numa']},true:" for d in devs)


@pytest.mark.gpu
def test_restarted_extender_rebuilds_state_from_annotations(built):
    """A second extender that has only seen the cluster's ANNOTATIONS -- node register strings
    (DecodeNodeDevices, scheduler.go:152-222) and the scheduled pods' vgpu-node / vgpu-ids-new
    (DecodePodDevices via onAddPod, scheduler.go:72-92) -- answers every further Filter like the one
    that made the placements"""
    rng = np.random.default_rng(21)
    cluster = make_cluster(rng, 30)
    pods = make_pods(rng, 90)
    first = v.hostlib.Scheduler(max_nodes=64)
    for n, devs in cluster.items():
        first.add_node(n, devs)
    names = list(cluster)
    placed = {}
    for uid, ctrs, annos in pods[:50]:
        r = first.filter(uid, ctrs, annos, names)
        if r["node"] is not None:
            placed[uid] = (r["node"], r["ids"])
    assert len(placed) > 20
    second = v.hostlib.Scheduler(max_nodes=64, incremental=True)
    for n, devs in cluster.items():
        half = len(devs) // 2  # two register annotations per node (two vendors' plugins): the node GAINS devices
        assert second.register_node(n, _encode_node_devices(devs[:half])) == half
        assert second.register_node(n, _encode_node_devices(devs[half:])) == len(devs) - half
        assert second.register_node(n, _encode_node_devices(devs[half:])) == 0  # seen again 15 s later: nothing new
    second.register_node("bogus", "no-colon-here")  # rejected by DecodeNodeDevices, not registered
    for uid, (node, ids) in placed.items():
        second.on_add_pod(uid, node, ids)
    second.on_add_pod("half-annotated", "node-000", None)  # no ids annotation: ignored (:82-85)
    gone = list(placed)[3]
    second.on_add_pod(gone, *placed[gone], terminated=True)  # Succeeded/Failed pod: its usage is released (:86-89)
    first.del_pod(gone)
    for uid, ctrs, annos in pods[50:]:
        a = first.filter(uid, ctrs, annos, names + ["bogus"])
        b = second.filter(uid, ctrs, annos, names + ["bogus"])
        assert a == b, (uid, a, b)
    # a vendor's devices leave a node (rmNodeDevice, nodes.go:76-101)
    n0 = names[0]
    for sch in (first, second):
        sch.rm_node_devices(n0, _encode_node_devices(cluster[n0][:2]))
    a = first.filter("after-rm", [{"gpu": len(cluster[n0]) - 1}], {}, [n0])
    b = second.filter("after-rm", [{"gpu": len(cluster[n0]) - 1}], {}, [n0])
    assert a == b and a["node"] is None  # only len-2 devices are left
    first.close()
    second.close()


def test_host_mirror_without_gpu(built):
    """CPU box: the C++ host mirror loads; a pod that asks for no vGPU passes every node without
    touching the GPU (scheduler.go:357-370); a pod that does ask fails LOUDLY -- no CPU fallback"""
    import torch
    sch = v.hostlib.Scheduler(max_nodes=8)
    sch.add_node("n0", [dict(ID=f"U{s}", Count=10, Devmem=32768, Devcore=100, Type="NVIDIA-Tesla V100", Numa=0) for s in range(8)])
    r = sch.filter("p0", [{}, {}], {"nvidia.com/use-gputype": "a100"}, ["n0", "n1"])
    assert r == dict(node=None, ids=None, failed={}, passed=["n0", "n1"])
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no usable GPU"):
            sch.filter("p1", [{"gpu": 1}], {}, ["n0"])
    sch.close()
