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
