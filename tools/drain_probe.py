"""time vgs_drain_batch against P single-pod vgs_score_batch + vgs_snapshot_delta round trips"""
import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
N, P = int(sys.argv[1]) if len(sys.argv) > 1 else 100000, int(sys.argv[2]) if len(sys.argv) > 2 else 1024
snap, batch = v.workload.make_config("cfg3", N, P)
with v.Context(max_nodes=N, max_pods=P, max_reqs=P, max_picks=2 * P) as ctx:
    ctx.upload(snap)
    t0 = time.perf_counter(); r = ctx.drain(batch); t1 = time.perf_counter()
    print(f"drain {P} pods x {N} nodes: {1e3*(t1-t0):.1f} ms = {1e6*(t1-t0)/P:.1f} us/pod, placed {(r.best_row>=0).sum()}, distinct nodes {len(set(r.best_row.tolist()))}")
    ctx.upload(snap)
    pb = batch.pick_begin
    t0 = time.perf_counter()
    for p in range(min(P, 256)):
        one = batch.select([p]); rr = ctx.score(one, snap.type_names)
        if rr.best_row[0] >= 0:
            d = np.array([(int(rr.best_row[0]), int(s), 1, int(m), int(c)) for s, m, c in rr.picks], dtype=v.capi.DELTA_DTYPE)
            ctx.delta(d)
    t1 = time.perf_counter()
    print(f"host loop (score P=1 + delta): {1e6*(t1-t0)/min(P,256):.1f} us/pod")
