"""one-off fuzz of the chunked drain against sequential oracle Filters (the generators of tests/test_gpu_parity.py)
usage: python tools/drain_fuzz.py [first_seed] [n_cases]"""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import test_gpu_parity as T
v = T.v
s0, cases = int(sys.argv[1]) if len(sys.argv) > 1 else 1000, int(sys.argv[2]) if len(sys.argv) > 2 else 60
bad = 0
for seed in range(s0, s0 + cases):
    rng = np.random.default_rng(seed)
    dmax = int(rng.choice([1, 2, 4, 7, 8, 8, 8, 12]))
    n = int(rng.choice([20, 64, 200, 700, 2500]))
    p = int(rng.integers(32, 260))
    chunk = int(rng.choice([0, 0, 5, 17, 33, 64]))
    snap = T._random_snapshot(rng, n, dmax) if rng.random() < 0.7 else v.workload.make_snapshot(n, seed, dmax=min(dmax, 8) if dmax != 12 else 12)
    batch = T._random_pods(rng, p, snap.dmax) if rng.random() < 0.6 else (v.workload.pods_mixed(p, seed) if rng.random() < 0.5 else v.workload.pods_multi(p, seed))
    if len(sys.argv) > 3 and sys.argv[3] == "tight":
        # nodes that take one or two pods each and pods that differ little: rows fill up quickly, candidate lists run dry
        d = int(rng.choice([2, 4, 8]))
        planes_ = np.zeros((8, n, d), np.int32)
        planes_[1] = rng.integers(1, 3, (n, d))
        planes_[0] = np.where(rng.random((n, d)) < 0.8, planes_[1], planes_[1] - 1)
        planes_[3], planes_[5] = 16000, 100
        planes_[2] = 1000 * planes_[0]
        planes_[6] = rng.integers(0, 2, (n, d))
        snap = v.Snapshot(planes_, np.full(n, d, np.int32), ["NVIDIA-A100"])
        batch = v.PodBatch.from_pods([{"containers": [[(int(rng.choice([1, 1, 1, 2])), int(rng.choice([500, 1000])), 101, int(rng.choice([0, 10])), "NVIDIA")]]}
                                      for _ in range(p)])
    rows, scores, counts, picks, planes = T._oracle_drain(snap, batch)
    with v.Context(max_nodes=n, max_devices=snap.dmax, max_pods=p, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum()),
                   tuning={"drain_chunk": chunk}) as ctx:
        ctx.upload(snap)
        got = ctx.drain(batch)
        after, _ = ctx.download_snapshot()
        pl = ctx.plan()
    ok = (np.array_equal(got.best_row, rows) and np.array_equal(got.score.view(np.uint32), scores.view(np.uint32))
          and np.array_equal(got.fit_count, counts) and np.array_equal(got.picks, picks) and np.array_equal(after, planes))
    bad += not ok
    print(f"seed {seed}: dmax {snap.dmax} n {n} p {p} chunk {chunk} chunks {pl.drain_chunks} rescans {pl.drain_rescans} placed {(rows >= 0).sum()} {'ok' if ok else 'MISMATCH'}", flush=True)
print("mismatches:", bad)
