"""wall clock of the resident vgs_score_batch call for a big batch on small snapshots, default tuning against
"slice only when pods x nodes >= slice_small_pairs" (slice_min out of reach)"""
import os, sys, time, importlib, statistics, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
for N in (64, 512, 2048, 4096, 8192, 16384):
    snap = v.workload.make_snapshot(N, v.workload.SEED + 3)
    for P in (1024, 4096):
        _, batch = v.workload.make_config("cfg3", 8, P)
        pods, reqs = v.capi.pack_batch(batch, snap.type_names)
        out = []
        for tune in (None, {"slice_min": 10**9}):
            with v.Context(max_nodes=N, max_pods=P, max_reqs=P, max_picks=4 * P, tuning=tune) as ctx:
                ctx.upload(snap)
                ts = []
                for r in range(25):
                    t0 = time.perf_counter(); ctx.score_packed(pods, reqs); ts.append(time.perf_counter() - t0)
                out.append((1e6 * statistics.median(ts[5:]), 1e3 * ctx.timing().score_ms))
        print(f"N={N:6d} P={P:5d}: default {out[0][0]:7.1f} us (score {out[0][1]:6.1f}) | pairs rule {out[1][0]:7.1f} us (score {out[1][1]:6.1f})", flush=True)
