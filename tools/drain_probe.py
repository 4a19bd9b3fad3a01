"""time vgs_drain_batch: the batch-scale drain (chunks) against the per-pod chain, results compared with each other
usage: python tools/drain_probe.py [cfg] [nodes] [pods] [chunk ...]"""
import os, sys, time, importlib, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
P = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
chunks = [int(x) for x in sys.argv[4:]] or [-1, 0, 32]
snap, batch = v.workload.make_config(cfg, N, P)
ref = None
for ch in chunks:
    with v.Context(max_nodes=N, max_devices=snap.dmax, max_pods=P, max_reqs=batch.n_reqs,
                   max_picks=int(batch.nums.sum()), tuning={"drain_chunk": ch}) as ctx:
        best = None
        pods, reqs = v.capi.pack_batch(batch, snap.type_names)
        results = np.zeros(len(pods), dtype=v.capi.RESULT_DTYPE)
        picks = np.zeros(int(reqs["nums"].sum()), dtype=v.capi.PICK_DTYPE)
        for rep in range(5):
            ctx.upload(snap)
            t0 = time.perf_counter()
            rc = ctx.lib.vgs_drain_batch(ctx.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data,
                                         results.ctypes.data, picks.ctypes.data)  # the C call alone
            t1 = time.perf_counter()
            assert rc == 0, rc
            best = min(best or 1e9, t1 - t0)
        r = ctx._to_result(results, picks)
        plan = ctx.plan()
        t = ctx.timing()
        same = ""
        if ref is None:
            ref = r
        else:
            same = " same=" + str(bool(np.array_equal(ref.best_row, r.best_row) and np.array_equal(ref.fit_count, r.fit_count)
                                        and np.array_equal(ref.score.view(np.uint32), r.score.view(np.uint32))
                                        and np.array_equal(ref.picks, r.picks)))
        print(f"{cfg} {P} pods x {N} nodes chunk={ch}: {1e3*best:.2f} ms = {1e6*best/P:.2f} us/pod (device {t.score_ms:.2f} ms), "
              f"chunks {plan.drain_chunks} rescans {plan.drain_rescans}, placed {(r.best_row>=0).sum()}, "
              f"distinct nodes {len(set(r.best_row.tolist()))}{same}", flush=True)
