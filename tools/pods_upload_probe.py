import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
snap, batch = v.workload.make_config("cfg3", 1000, 4096)
pods, reqs = v.capi.pack_batch(batch, snap.type_names)
for smin in ("512", "100000"):
    os.environ["VGS_SLICE_MIN"] = smin
    ctx = v.Context(max_nodes=1000, max_pods=4096, max_reqs=4096, max_picks=8192)
    ctx.upload(snap)
    ts = []
    for i in range(40):
        t0 = time.perf_counter(); ctx.pods_upload(pods, reqs); ctx.sync(); ts.append(time.perf_counter() - t0)
    print("slice_min", smin, "pods_upload+sync median us", 1e6 * np.median(ts[5:]))
    ctx.close()
