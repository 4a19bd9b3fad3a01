"""single pod against a 1M-node resident snapshot: the HBM-bound regime"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
n_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 40
snap, batch = v.workload.make_config("cfg3", 100_000, 64)
big = v.Snapshot(np.tile(snap.planes, (1, 10, 1)), np.tile(snap.ndev, 10), snap.type_names)
with v.Context(max_nodes=big.n_nodes, max_pods=1) as c:
    c.upload(big)
    for which in (0, 1, 3):  # pods of different request shapes
        p1, r1 = v.capi.pack_batch(batch.select([which]), snap.type_names)
        ks = []
        for i in range(n_iter):
            c.score_packed(p1, r1)
            ks.append(c.timing().score_ms)
        k = float(np.median(ks[n_iter // 4:]))
        print(f"pod {which} req={r1[0]} score phase {1e3 * k:.1f} us  {260.000128 / k / 1e3:.2f} TB/s algorithmic")
