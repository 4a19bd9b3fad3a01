"""Three restatements of /root/reference/pkg/scheduler/score.go have to agree bit for bit: R1 (C, structs and strings),
R2 (C, packed planes and type masks) and R0 (oracle/score_py.py: plain Python objects, written from the Go again and
sharing no code with the C file).  R0 is also held against the hand-derived known answers.  The oracle stays unpinned
by anything the reference itself produced (it cannot be run here); this closes the gap "one C file by one author" as far
as that can be done without Go.  Small cases only: R0 is Python loops.  No GPU."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v
from helpers import assert_same, check_kat, kat_inputs, load_kats
from oracle.score_py import score_python

K = load_kats()

TYPES = ["NVIDIA-Tesla V100-PCIE-32GB", "NVIDIA-NVIDIA A100-SXM4-40GB", "NVIDIA-Tesla T4", "DCU-Z100", "MLU370-X8", "MLU290",
         "NVIDIA-карта Ж1", "NVIDIA-straße"]


def _snapshot(rng, n, dmax, mixed_types, n_types=len(TYPES), friendly=False):
    planes = np.zeros((8, n, dmax), np.int32)
    count = rng.integers(1, 12, (n, dmax))
    planes[1] = count
    planes[0] = np.minimum(rng.integers(0, 14, (n, dmax)), count + rng.integers(0, 2, (n, dmax)))  # sometimes Used > Count
    totalmem = rng.choice([8000, 16000, 32768, 81920, 30_000_000], (n, dmax))  # the last: Totalmem * pct wraps in int32
    planes[3] = totalmem
    planes[2] = (np.minimum(totalmem, 81920) * rng.random((n, dmax)) * (planes[0] > 0)).astype(np.int32)
    totalcore = rng.choice([100, 100, 100, 200, 0], (n, dmax))
    planes[5] = totalcore
    planes[4] = np.minimum(totalcore, rng.choice([0, 0, 10, 50, 100, 200], (n, dmax))) * (planes[0] > 0)
    planes[6] = rng.integers(0, 3, (n, dmax)) * 3 - 1  # numa -1, 2, 5
    if mixed_types:
        planes[7] = rng.integers(0, n_types, (n, dmax))
    else:
        planes[7] = rng.integers(0, n_types, (n, 1))
    ndev = rng.integers(0, dmax + 1, n).astype(np.int32)
    if friendly:  # room everywhere: most requests find their devices, so the takes and re-sorts carry the comparison
        planes[0] = rng.integers(0, count // 2 + 1, (n, dmax))
        planes[2] = (np.minimum(totalmem, 81920) * 0.3 * rng.random((n, dmax)) * (planes[0] > 0)).astype(np.int32)
        planes[5] = 100
        planes[4] = rng.choice([0, 10, 30], (n, dmax)) * (planes[0] > 0)
        ndev = rng.integers((dmax + 1) // 2, dmax + 1, n).astype(np.int32)
    return v.Snapshot(planes, ndev, TYPES)


def _pods(rng, p, dmax, friendly=False):
    """friendly: NVIDIA requests of modest size (most nodes fit, many takes and re-sorts); else the hard mix of vendors,
    Nums up to dmax + 1, Coresreq beyond 100, negative values"""
    pods = []
    for _ in range(p):
        ctrs = []
        for _ in range(int(rng.integers(1, 4))):
            if rng.random() < 0.15:
                ctrs.append([])  # a container that asks for nothing (score.go:194-197)
                continue
            ctr = []
            vendors = ["NVIDIA"] if friendly else rng.permutation(["NVIDIA", "DCU", "MLU"])[:int(rng.choice([1, 1, 1, 2, 3]))]
            for vendor in vendors:
                nums = int(rng.choice([1, 1, 2, 3] if friendly else [1, 1, 2, 3, dmax, dmax + 1]))
                u = rng.random()
                if u < 0.4:
                    mem, pct = int(rng.choice([0, 500, 4000, 20000])), 101
                elif u < 0.8:
                    mem, pct = 0, int(rng.choice([0, 10, 50, 100, 101, 90]))
                else:
                    mem, pct = int(rng.choice([500, 4000, -5, -4000])), int(rng.choice([10, 50, 100, 0]))
                cores = int(rng.choice([0, 0, 10, 30] if friendly else [0, 0, 10, 50, 100, 101, -10]))
                ctr.append((nums, mem, pct, cores, str(vendor)))
            ctrs.append(ctr)
        if not any(ctrs):
            ctrs.append([(1, 0, 101, 0, "NVIDIA")])
        annos = {}
        r = rng.random()
        if r < 0.25:
            annos["nvidia.com/use-gputype"] = str(rng.choice(["a100", "V100,t4", "h100,", "mlu", "карта", "STRASSE", "straße,t4", "teſla"]))
        elif r < 0.45:
            annos["nvidia.com/nouse-gputype"] = str(rng.choice(["a100", "V100,t4", "T4", "ж1", ""]))
        if rng.random() < 0.35:
            annos["nvidia.com/numa-bind"] = str(rng.choice(["true", "1", "false", "T", "yes"]))
        if rng.random() < 0.3:
            annos["hygon.com/use-dcutype"] = str(rng.choice(["z100", "k100", "Z100,k100"]))
        if rng.random() < 0.3:
            annos[str(rng.choice(["cambricon.com/use-mlutype", "cambricon.com/nouse-mlutype"]))] = str(rng.choice(["370", "mlu290", "x8,290"]))
        pods.append({"containers": ctrs, "annos": annos})
    return v.PodBatch.from_pods(pods)


@pytest.mark.parametrize("k", K["kats"], ids=[k["name"] for k in K["kats"]])
def test_kat_python(k):
    snap, batch = kat_inputs(k)
    check_kat(k, score_python(snap, batch), batch)


@pytest.mark.parametrize("seed,n,dmax,p,mixed,friendly", [
    (1, 120, 8, 100, True, False), (2, 60, 12, 60, True, False), (3, 150, 4, 80, False, False), (4, 80, 1, 50, True, False),
    (5, 90, 7, 80, False, False), (6, 100, 8, 100, True, True), (7, 80, 12, 80, False, True), (8, 120, 3, 100, True, True)])
def test_three_restatements_agree(built, seed, n, dmax, p, mixed, friendly):
    rng = np.random.default_rng(7000 + seed)
    snap = _snapshot(rng, n, dmax, mixed, n_types=3 if friendly else len(TYPES), friendly=friendly)  # (types 0-2: NVIDIA cards)
    batch = _pods(rng, p, dmax, friendly)
    r0 = score_python(snap, batch)
    assert (r0.best_row >= 0).sum() > 0 and r0.fit.sum() < r0.fit.size  # the case is not trivial
    assert not friendly or r0.fit.mean() > 0.2
    assert_same(oracle.score_faithful(snap, batch, full=True), r0, "R1 vs R0")
    assert_same(oracle.score_packed(snap, batch, full=True), r0, "R2 vs R0")


def test_workload_generators_agree(built):
    """the bench's own generators at a size Python gets through"""
    for cfg in ("cfg3", "cfg4"):
        snap, batch = v.workload.make_config(cfg, 48, 24)
        r0 = score_python(snap, batch)
        assert_same(oracle.score_faithful(snap, batch, full=True), r0, cfg + " R1 vs R0")
        assert_same(oracle.score_packed(snap, batch, full=True), r0, cfg + " R2 vs R0")
