"""the drain with 4096 identical pods on the cfg3 snapshot (replicas of one Deployment)"""
import os, sys, time, importlib, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
snap, _ = v.workload.make_config("cfg3", 100000, 1)
for req in ((1, 3000, 101, 30), (2, 8000, 101, 0), (1, 0, 50, 100)):
    p = 4096
    batch = v.PodBatch.from_pods([{"containers": [[req + ("NVIDIA",)]]} for _ in range(p)])
    pods, reqs = v.capi.pack_batch(batch, snap.type_names)
    out = []
    for ch in (0, -1):
        with v.Context(max_nodes=snap.n_nodes, max_pods=p, max_reqs=len(reqs), max_picks=int(reqs["nums"].sum()), tuning={"drain_chunk": ch}) as ctx:
            res = np.zeros(len(pods), dtype=v.capi.RESULT_DTYPE); pk = np.zeros(int(reqs["nums"].sum()), dtype=v.capi.PICK_DTYPE)
            best = 1e9
            for rep in range(2):
                ctx.upload(snap)
                t0 = time.perf_counter()
                rc = ctx.lib.vgs_drain_batch(ctx.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data, res.ctypes.data, pk.ctypes.data)
                best = min(best, time.perf_counter() - t0); assert rc == 0
            pl = ctx.plan()
            out.append((best, pl.drain_chunks, pl.drain_rescans, pl.drain_chain_pods, res.copy(), pk.copy()))
    same = bool(np.array_equal(out[0][4], out[1][4]) and np.array_equal(out[0][5], out[1][5]))
    print(f"req {req}: chunked {1e3*out[0][0]:.2f} ms (rescans {out[0][2]}, chain pods {out[0][3]}) | per-pod chain {1e3*out[1][0]:.2f} ms | same={same} placed {(out[0][4]['best_row']>=0).sum()} distinct rows {len(set(out[0][4]['best_row'].tolist()))}", flush=True)
