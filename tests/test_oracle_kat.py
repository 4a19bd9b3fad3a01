"""The oracle against the hand-derived known-answer vectors -- CPU only.

The reference has no test for pkg/scheduler; these vectors are the oracle's only pin."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v
from helpers import assert_same, cambricon_cluster, check_kat, kat_inputs, load_kats

K = load_kats()


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_kat_faithful(k):
    snap, batch = kat_inputs(k)
    check_kat(k, oracle.score_faithful(snap, batch), batch)


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_kat_packed(k):
    snap, batch = kat_inputs(k)
    check_kat(k, oracle.score_packed(snap, batch, threads=2), batch)


def test_policy_truth_table():
    for annos, card, want in K["policy"]:
        m, _ = oracle.typemask(annos, "NVIDIA", [card])
        assert bool(m & 1) == want, (annos, card)
    # d.Type must contain the request type (score.go:73); unknown vendors never pass
    assert oracle.typemask({}, "NVIDIA", ["MLU370"])[0] == 0
    assert oracle.typemask({}, "TPU", ["TPU-v4"])[0] == 0
    # Cambricon (cambricon/device.go:93-104), by hand: (mask over [MLU370-X8, MLU290, NVIDIA-x], flags)
    names = ["MLU370-X8", "MLU290", "NVIDIA-x"]
    assert oracle.typemask({}, "MLU", names, memreq=0) == (0b011, 2)       # flags bit 1: 370 in use fails
    assert oracle.typemask({}, "MLU", names, memreq=1000) == (0b001, 0)    # :94-96 drops the 290
    assert oracle.typemask({"cambricon.com/use-mlutype": "mlu290"}, "MLU", names, memreq=0) == (0b010, 2)
    assert oracle.typemask({"cambricon.com/nouse-mlutype": "x8,t4"}, "MLU", names, memreq=0) == (0b010, 2)
    assert oracle.typemask({"nvidia.com/use-gputype": "none"}, "MLU", names, memreq=0) == (0b011, 2)


def test_numa_parse_bool():
    for val, want in K["numa"]:
        annos = {} if val is None else {"nvidia.com/numa-bind": val}
        assert oracle.typemask(annos, "NVIDIA", ["NVIDIA-x"])[1] == want, val


def test_encode_golden():
    from oracle.ref import encode_pod_devices
    for ctrs, want in K["encode"]:
        assert encode_pod_devices([[tuple(d) for d in c] for c in ctrs]) == want


@pytest.mark.parametrize("cfg,n,p", [("cfg2", 600, 48), ("cfg3", 700, 96), ("cfg4", 500, 64)])
def test_faithful_equals_packed(cfg, n, p):
    """the string/pointer restatement and the packed multi-threaded arm agree bit for bit"""
    snap, batch = v.workload.make_config(cfg, n, p)
    a = oracle.score_faithful(snap, batch, full=True)
    b = oracle.score_packed(snap, batch, full=True, threads=3)
    from helpers import assert_same
    assert_same(a, b, cfg)
    assert a.fit.any() and not a.fit.all()


def test_ragged_and_empty():
    # nodes with 0, 3 and 8 devices; dmax 8
    devs = lambda n: [dict(Used=0, Count=10, Usedmem=0, Totalmem=32768, Usedcores=0, Totalcore=100, Numa=0)] * n
    snap = v.Snapshot.from_nodes([devs(0), devs(3), devs(8)], dmax=8)
    batch = v.PodBatch.from_pods([{"containers": [[(4, 100, 101, 0)]]}, {"containers": [[(1, 100, 101, 0)]]}])
    r = oracle.score_faithful(snap, batch, full=True)
    assert r.fit.tolist() == [[False, False, True], [False, True, True]]
    # 1 device of 8 -> 10/10 + 7 = 8 ; 1 of 3 -> 1 + 2 = 3
    assert r.scores[1].tolist() == [0.0, 3.0, 8.0] and r.best_row.tolist() == [2, 2]
    empty = v.Snapshot(np.zeros((8, 0, 8), np.int32), np.zeros(0, np.int32), ["NVIDIA-x"])
    r = oracle.score_packed(empty, batch)
    assert r.best_row.tolist() == [-1, -1]


def test_negative_scores_keep_their_order():
    """two requests in one container make len(devices)-sums negative (score.go:180); the
    best-node pick must still be the numeric maximum in both oracle arms"""
    rng = np.random.default_rng(11)
    n = 300
    planes = np.zeros((8, n, 8), np.int32)
    planes[1] = 10
    planes[0] = rng.integers(0, 10, (n, 8))
    planes[3] = 32768
    planes[5] = 100
    snap = v.Snapshot(planes, np.full(n, 8, np.int32), ["NVIDIA-Tesla V100"])
    batch = v.PodBatch.from_pods([{"containers": [[(5, 100, 101, 0), (5, 100, 101, 0)]]}])
    a = oracle.score_faithful(snap, batch, full=True)
    b = oracle.score_packed(snap, batch, full=True, threads=2)
    from helpers import assert_same
    assert_same(a, b, "negative")
    fs = a.scores[0][a.fit[0]]
    assert fs.min() < 0 and a.score[0] == fs.max()


def test_cambricon_faithful_equals_packed():
    """the string arm evaluates cambricon/device.go:93-104 per device with the CURRENT Used; the packed
    arm splits it into a static mask + an in-use rule: same fit/score matrices"""
    rng = np.random.default_rng(370)
    snap, batch = cambricon_cluster(rng, 500, 200)
    a = oracle.score_faithful(snap, batch, full=True)
    b = oracle.score_packed(snap, batch, full=True)
    assert_same(a, b, "mlu R1 vs R2")
    mlu_only = [p for p in range(batch.n_pods) if all(t == "MLU" for t in batch.req_type[slice(*batch.pod_req_range(p))])]
    assert a.fit[mlu_only].any() and not a.fit[mlu_only].all()
