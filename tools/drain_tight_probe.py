"""the drain on a tight cluster at scale (nodes that take one or two pods, near-identical pods): chunked path against
the per-pod chain.  usage: python tools/drain_tight_probe.py [nodes] [pods]"""
import os, sys, time, importlib, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
p = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
rng = np.random.default_rng(7)
d = 4
planes = np.zeros((8, n, d), np.int32)
planes[1] = rng.integers(1, 3, (n, d))
planes[0] = np.where(rng.random((n, d)) < 0.8, planes[1], planes[1] - 1)
planes[3], planes[5] = 16000, 100
planes[2] = 1000 * planes[0]
planes[6] = rng.integers(0, 2, (n, d))
snap = v.Snapshot(planes, np.full(n, d, np.int32), ["NVIDIA-A100"])
batch = v.PodBatch.from_pods([{"containers": [[(int(rng.choice([1, 1, 1, 2])), int(rng.choice([500, 1000])), 101,
                                                int(rng.choice([0, 10])), "NVIDIA")]]} for _ in range(p)])
pods, reqs = v.capi.pack_batch(batch, snap.type_names)
ref = None
for ch in (0, -1):
    with v.Context(max_nodes=n, max_devices=d, max_pods=p, max_reqs=len(reqs), max_picks=int(reqs["nums"].sum()), tuning={"drain_chunk": ch}) as ctx:
        res = np.zeros(len(pods), dtype=v.capi.RESULT_DTYPE); pk = np.zeros(int(reqs["nums"].sum()), dtype=v.capi.PICK_DTYPE)
        best = 1e9
        for rep in range(3):
            ctx.upload(snap)
            t0 = time.perf_counter()
            rc = ctx.lib.vgs_drain_batch(ctx.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data, res.ctypes.data, pk.ctypes.data)
            best = min(best, time.perf_counter() - t0); assert rc == 0
        pl = ctx.plan()
        same = "" if ref is None else f" same={bool(np.array_equal(ref[0], res) and np.array_equal(ref[1], pk))}"
        ref = ref or (res.copy(), pk.copy())
        print(f"tight {p} pods x {n} nodes chunk={ch}: {1e3*best:.2f} ms, chunks {pl.drain_chunks} rescans {pl.drain_rescans}{same}", flush=True)
