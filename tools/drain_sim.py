"""CPU simulation of the chunked batch-scale drain (design study, not product code): how often would pod q need a
full re-scan if, per chunk of C pods, every pod had only its K best rows of the chunk-start state plus the live
scores of the rows touched inside the chunk?
usage: python tools/drain_sim.py cfg3 20000 1024 [C ...]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
from oracle import ref as oracle

cfg, n, p = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
Cs = [int(x) for x in sys.argv[4:]] or [64, 256]
Ks = [1, 2, 4, 8, 16]
snap, batch = v.workload.make_config(cfg, n, p)


def keys_of(r, j=0):
    sc = r.scores[j].view(np.uint32).astype(np.uint64)
    k = (sc << np.uint64(32)) | np.arange(n, dtype=np.uint64)
    return np.where(r.fit[j], k, np.uint64(0))


for C in Cs:
    planes = snap.planes.copy()
    stats = {K: dict(fallback=0, list=0, touched=0, none=0) for K in Ks}
    tmax, tsum, nch = 0, 0, 0
    for c0 in range(0, p, C):
        idx = list(range(c0, min(p, c0 + C)))
        start = oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), batch.select(idx), full=True)
        K0 = [keys_of(start, j) for j in range(len(idx))]
        order = [np.argsort(k)[::-1][:max(Ks)] for k in K0]   # best rows of the chunk-start state
        T = []
        for j, q in enumerate(idx):
            one = batch.select([q])
            r = oracle.score_packed(v.Snapshot(planes, snap.ndev, snap.type_names), one, full=True)
            cur = keys_of(r)
            win = int(r.best_row[0])
            tb = max([cur[t] for t in T], default=np.uint64(0))
            for K in Ks:
                lst = [int(x) for x in order[j][:K] if K0[j][x] != 0]
                un = [x for x in lst if x not in T]
                lk = K0[j][un[0]] if un else np.uint64(0)
                exhausted = not un and len(lst) == K
                s = stats[K]
                if exhausted and tb < K0[j][lst[-1]]:
                    s["fallback"] += 1
                else:
                    best = max(tb, lk)
                    want = cur[win] if win >= 0 else np.uint64(0)
                    assert best == want, (q, K, best, want)
                    if win < 0: s["none"] += 1
                    elif win in T: s["touched"] += 1
                    else: s["list"] += 1
            if win >= 0:
                for sl, m, cc in r.picks:
                    planes[0, win, sl] += 1
                    planes[2, win, sl] += m
                    planes[4, win, sl] += cc
                if win not in T:
                    T.append(win)
        tmax, tsum, nch = max(tmax, len(T)), tsum + len(T), nch + 1
    print(f"{cfg} n={n} p={p} C={C}: touched rows per chunk max {tmax} mean {tsum / nch:.1f}")
    for K in Ks:
        print(f"   K={K}: {stats[K]}")
