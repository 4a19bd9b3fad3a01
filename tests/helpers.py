"""shared test helpers: KAT loading and result comparison (oracle = checker only)"""
import json
import os

import numpy as np

import vgpu_b200 as v

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_kats():
    with open(os.path.join(GOLDEN, "kat.json")) as f:
        return json.load(f)


def kat_inputs(k):
    snap = v.Snapshot.from_nodes([k["devices"]] * k["n_nodes"])
    ctrs = [[tuple(r) for r in c] for c in k["containers"]]
    batch = v.PodBatch.from_pods([{"containers": ctrs, "annos": k["annos"]}])
    return snap, batch


def check_kat(k, res, batch):
    """res: BatchResult-like for the single pod of the KAT"""
    if not k["fit"]:
        assert res.best_row[0] == -1, k["name"]
        assert res.fit_count[0] == 0, k["name"]
        return
    assert res.best_row[0] == k["best_row"], k["name"]
    assert res.fit_count[0] == k["n_nodes"], k["name"]
    assert int(res.score.view(np.uint32)[0]) == int(k["score_bits"], 16), (k["name"], res.score)
    pb = batch.pick_begin
    got = []
    for c in range(int(batch.ctr_begin[0]), int(batch.ctr_begin[1])):
        devs = []
        for r in range(int(batch.req_begin[c]), int(batch.req_begin[c + 1])):
            for j in range(int(batch.nums[r])):
                devs.append([int(x) for x in res.picks[int(pb[r]) + j]])
        got.append(devs)
    assert got == [[list(p) for p in c] for c in k["picks"]], (k["name"], got)


def assert_same(a, b, what=""):
    """bit-exact comparison of two BatchResult-likes (winner, score bits, counts, picks)"""
    assert np.array_equal(a.best_row, b.best_row), what + " best_row"
    assert np.array_equal(a.score.view(np.uint32), b.score.view(np.uint32)), what + " score bits"
    assert np.array_equal(a.fit_count, b.fit_count), what + " fit_count"
    assert np.array_equal(a.n_picks, b.n_picks), what + " n_picks"
    # picks of pods without a winner are unspecified (zero in both implementations)
    assert np.array_equal(a.picks, b.picks), what + " picks"
    if a.fit is not None and b.fit is not None:
        assert np.array_equal(a.fit, b.fit), what + " fit matrix"
        assert np.array_equal(a.scores.view(np.uint32), b.scores.view(np.uint32)), what + " score matrix"


def cambricon_cluster(rng, n_nodes, n_pods):
    """MLU370 / MLU290 / NVIDIA nodes (a few hold two card types) and pods of 1-3 containers asking
    for MLUs with and without memory, some with cambricon.com/{use,nouse}-mlutype, some NVIDIA"""
    kinds = ["MLU370-X8", "MLU290", "MLU370-X4", "NVIDIA-Tesla V100"]
    nodes = []
    for i in range(n_nodes):
        t = kinds[i % 4]
        devs = []
        for s in range(8):
            tt = kinds[(i + s // 4) % 4] if i % 9 == 0 else t
            used = int(rng.choice([0, 0, 0, 1, 2, 4]))
            devs.append(dict(Used=used, Count=4, Usedmem=int(rng.integers(0, 6000)) * (used > 0), Totalmem=24576,
                             Usedcores=int(rng.choice([0, 0, 50, 100])) * (used > 0), Totalcore=100, Numa=s // 4, Type=tt))
        nodes.append(devs[: int(rng.choice([8, 8, 8, 5, 2]))])
    pods = []
    for i in range(n_pods):
        annos = {}
        if i % 5 == 1:
            annos["cambricon.com/use-mlutype"] = str(rng.choice(["x8", "MLU290", "x4,x8", "370"]))
        elif i % 5 == 2:
            annos["cambricon.com/nouse-mlutype"] = str(rng.choice(["x8", "mlu290,x4"]))
        if i % 6 == 0:
            annos["nvidia.com/numa-bind"] = "true"  # Cambricon never asserts numa (cambricon/device.go:36-38)
        ctrs = []
        for _ in range(int(rng.choice([1, 1, 2, 3]))):
            if rng.random() < 0.2:
                ctrs.append([(int(rng.choice([1, 2])), 2000, 101, 10, "NVIDIA")])
            else:
                ctrs.append([(int(rng.choice([1, 1, 2, 3])), int(rng.choice([0, 0, 1000, 8000])), 0, 0, "MLU")])
        if rng.random() < 0.1:  # both vendors inside one container
            ctrs[0] = [(1, 0, 0, 0, "MLU"), (1, 1000, 101, 0, "NVIDIA")]
        pods.append({"containers": ctrs, "annos": annos})
    return v.Snapshot.from_nodes(nodes, dmax=8), v.PodBatch.from_pods(pods)
