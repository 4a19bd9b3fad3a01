"""vgs_group_*: node sharding inside the library (one process, one host thread, one context per shard).
The 1-shard / 1-GPU result is the oracle for the sharded one; the CPU oracle checks both."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v
from helpers import assert_same

pytestmark = pytest.mark.gpu


def _single(snap, batch):
    with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=batch.n_pods,
                   max_reqs=max(batch.n_reqs, batch.n_pods), max_picks=max(int(batch.nums.sum()), batch.n_reqs)) as ctx:
        ctx.upload(snap)
        return ctx.score(batch)


def _group(devices, exchange, snap, batch, passes=2, **kw):
    out = None
    with v.Group(devices, max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=batch.n_pods,
                 max_reqs=max(batch.n_reqs, batch.n_pods), max_picks=max(int(batch.nums.sum()), batch.n_reqs),
                 exchange=exchange, **kw) as g:
        g.upload(snap)
        for _ in range(passes):  # twice: every shard's candidate block must come back clean
            out = g.score(batch)
        total_us, exchange_us = g.timing()
        assert total_us > 0
    return out


@pytest.mark.parametrize("cfg,n,p,shards", [("cfg3", 9001, 700, 2), ("cfg4", 5000, 96, 3), ("cfg3", 40_000, 2100, 8),
                                            ("cfg2", 10_000, 256, 4), ("cfg3", 5, 33, 8)])
def test_group_peer_exchange_equals_single_context(built, cfg, n, p, shards):
    """several shards on ONE GPU through vgs_group_* (peer-store exchange): winner, score bits, summed fit counts
    and the winner's picks equal one context over the whole node set, and the oracle"""
    snap, batch = v.workload.make_config(cfg, n, p)
    want = oracle.score_packed(snap, batch)
    assert_same(_single(snap, batch), want, "single")
    assert_same(_group([0] * shards, v.capi.EXCHANGE_PEER, snap, batch), want, f"group x{shards}")


def test_group_deltas_reach_the_owning_shard(built):
    snap, batch = v.workload.make_config("cfg3", 6000, 300)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    with v.Group([0, 0, 0], max_nodes=snap.n_nodes, max_pods=batch.n_pods, exchange=v.capi.EXCHANGE_PEER) as g:
        g.upload(snap)
        r0, pk0 = g.score_packed(pods, reqs)
        # commit every winner (global rows), score again: equals the oracle on the updated planes
        pod_of_pick = np.repeat(np.arange(len(pods)), reqs["nums"])  # one request per pod in cfg3
        rows = r0["best_row"][pod_of_pick]
        ok = rows >= 0
        d = np.zeros(int(ok.sum()), dtype=v.capi.DELTA_DTYPE)
        d["row"], d["slot"], d["d_used"] = rows[ok], pk0["slot"][ok], 1
        d["d_usedmem"], d["d_usedcores"] = pk0["usedmem"][ok], pk0["usedcores"][ok]
        g.delta(d)
        got = g.score(batch)
    planes = snap.planes.copy()
    np.add.at(planes[0], (d["row"], d["slot"]), 1)
    np.add.at(planes[2], (d["row"], d["slot"]), d["d_usedmem"])
    np.add.at(planes[4], (d["row"], d["slot"]), d["d_usedcores"])
    assert_same(got, oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), batch), "after deltas")


def test_group_nccl_exchange(built):
    """the NCCL all-gather exchange: one rank per GPU, so it needs >= 2 GPUs (gpurun --gpus 2); with one GPU the
    single-shard NCCL group must still work and duplicate devices must be refused"""
    import torch
    snap, batch = v.workload.make_config("cfg3", 20_000, 1500)
    want = oracle.score_packed(snap, batch)
    assert_same(_group([0], v.capi.EXCHANGE_NCCL, snap, batch), want, "nccl x1")
    with pytest.raises(v.VgsError) as e:
        v.Group([0, 0], max_nodes=100, exchange=v.capi.EXCHANGE_NCCL)
    assert e.value.code == v.capi.VGS_EINVAL
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible: the multi-rank NCCL exchange runs under gpurun --gpus 2")
    devs = list(range(min(n, 8)))
    assert_same(_group(devs, v.capi.EXCHANGE_NCCL, snap, batch), want, f"nccl x{len(devs)}")
    assert_same(_group(devs, v.capi.EXCHANGE_PEER, snap, batch), want, f"peer x{len(devs)}")
    snap4, batch4 = v.workload.make_config("cfg4", 30_000, 200)
    want4 = oracle.score_packed(snap4, batch4)
    assert_same(_group(devs, v.capi.EXCHANGE_NCCL, snap4, batch4), want4, "nccl cfg4")
    assert_same(_group(devs, v.capi.EXCHANGE_PEER, snap4, batch4), want4, "peer cfg4")
