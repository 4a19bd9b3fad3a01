"""vgs_score_batch_noalloc -- the entry point for callers that may not allocate GPU memory
(memory-capped vGPU containers, processes whose device memory belongs to another pool).

The caller owns every device buffer (here: torch tensors); the library works out of its
fixed built-in scratch array.  Its answers must be identical to the allocating entry point
(vgs_score_batch) -- winner row, score bits, fit count, picks -- on two calls in a row with
different sizes, so stale scratch contents from the first call cannot leak into the second.
Collected last on purpose (file name), runs alone or after the other GPU tests."""
import numpy as np
import pytest

import oracle
import vgpu_b200 as v

pytestmark = pytest.mark.gpu


def _via_context(snap, batch):
    with v.Context(max_nodes=snap.n_nodes, max_devices=snap.dmax, max_pods=batch.n_pods,
                   max_reqs=max(batch.n_reqs, batch.n_pods), max_picks=max(int(batch.nums.sum()), batch.n_reqs)) as ctx:
        ctx.upload(snap)
        pods, reqs = v.capi.pack_batch(batch, snap.type_names)
        return ctx.score_packed(pods, reqs)


def _via_noalloc(snap, batch, torch, dev):
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    planes = torch.from_numpy(snap.planes).to(dev)  # the caller's own device memory
    ndev = torch.from_numpy(snap.ndev).to(dev)
    n_picks = int(reqs["nums"].sum())
    out_res = torch.full((len(pods) * 4,), 0x5A5A5A5A, dtype=torch.int32, device=dev)  # garbage in
    out_picks = torch.full((max(n_picks, 1) * 3,), 0x5A5A5A5A, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    rc = v.capi.score_batch_noalloc(dev.index, 0, snap.n_nodes, snap.dmax, 0,
                                    [planes[f].data_ptr() for f in range(8)], ndev.data_ptr(), pods, reqs,
                                    out_res.data_ptr(), out_picks.data_ptr())
    assert rc == 0, rc
    res = out_res.cpu().numpy().view(v.capi.RESULT_DTYPE)
    picks = out_picks.cpu().numpy()[: n_picks * 3].view(v.capi.PICK_DTYPE)
    return res, picks


def test_noalloc_matches_allocating_entry_point_twice(built):
    import torch
    dev = torch.device("cuda", 0)
    cases = [v.workload.make_config("cfg4", 9000, 200),   # multi-container, numa-bind
             v.workload.make_config("cfg3", 2500, 77)]    # smaller second call: stale scratch must not matter
    for i, (snap, batch) in enumerate(cases):
        want_res, want_picks = _via_context(snap, batch)
        got_res, got_picks = _via_noalloc(snap, batch, torch, dev)
        assert np.array_equal(got_res["best_row"], want_res["best_row"]), f"call {i}"
        assert np.array_equal(got_res["score"].view(np.uint32), want_res["score"].view(np.uint32)), f"call {i}"
        assert np.array_equal(got_res["fit_count"], want_res["fit_count"]), f"call {i}"
        assert np.array_equal(got_res["n_picks"], want_res["n_picks"]), f"call {i}"
        assert np.array_equal(got_picks, want_picks), f"call {i}"
        assert (want_res["best_row"] >= 0).any()
        # ... and a parity test in its own right: the same answers as the CPU oracle
        ref = oracle.score_packed(snap, batch)
        assert np.array_equal(got_res["best_row"], ref.best_row), f"call {i} vs oracle"
        assert np.array_equal(got_res["score"].view(np.uint32), ref.score.view(np.uint32)), f"call {i} vs oracle"
        assert np.array_equal(got_res["fit_count"], ref.fit_count), f"call {i} vs oracle"
        assert np.array_equal(np.stack([got_picks["slot"], got_picks["usedmem"], got_picks["usedcores"]], axis=1),
                              ref.picks), f"call {i} vs oracle"


def test_noalloc_rejects_sizes_beyond_its_limits(built):
    import torch
    dev = torch.device("cuda", 0)
    snap, batch = v.workload.make_config("cfg2", 64, 4)
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    planes = torch.from_numpy(snap.planes).to(dev)
    ndev = torch.from_numpy(snap.ndev).to(dev)
    out_res = torch.zeros(64, dtype=torch.int32, device=dev)
    out_picks = torch.zeros(64, dtype=torch.int32, device=dev)
    args = ([planes[f].data_ptr() for f in range(8)], ndev.data_ptr(), pods, reqs, out_res.data_ptr(),
            out_picks.data_ptr())
    assert v.capi.score_batch_noalloc(0, 0, v.capi.NOALLOC_MAX_NODES + 1, 8, 0, *args) == v.capi.VGS_ELIMIT
    assert v.capi.score_batch_noalloc(0, 0, 64, v.capi.NOALLOC_MAX_DEVICES + 1, 0, *args) == v.capi.VGS_ELIMIT
    many = np.zeros(v.capi.NOALLOC_MAX_PODS + 1, dtype=v.capi.POD_DTYPE)
    assert v.capi.score_batch_noalloc(0, 0, 64, 8, 0, args[0], args[1], many, reqs, args[4], args[5]) == v.capi.VGS_ELIMIT
    assert v.load_library().vgs_strerror(v.capi.VGS_ELIMIT).startswith(b"larger than the fixed")
    assert v.capi.score_batch_noalloc(0, 0, 64, 8, 0, *args) == 0  # still usable afterwards
