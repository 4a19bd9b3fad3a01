"""K0 (snapshot preparation) time per refresh, CUDA events"""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
snap, batch = v.workload.make_config("cfg3")
with v.Context(max_nodes=snap.n_nodes, max_pods=8) as ctx:
    ctx.upload(snap)
    ts = []
    for i in range(50):
        ctx.refresh(); ts.append(ctx.timing().derive_ms)
    print("K0 derive_ms median", float(np.median(ts[10:])), "min", min(ts))
