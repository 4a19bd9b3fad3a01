"""where the end-to-end step (bench.py `e2e`) spends its wall clock: per call and per device phase"""
import sys, time, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
snap, batch = v.workload.make_config("cfg3")
pods, reqs = v.capi.pack_batch(batch, snap.type_names)
ctx = v.Context(max_nodes=snap.n_nodes, max_pods=4096, max_reqs=4096, max_picks=8192)
pin = torch.from_numpy(snap.planes).pin_memory(); pnd = torch.from_numpy(snap.ndev).pin_memory()
ptrs = [pin[f].data_ptr() for f in range(8)]
for asyn in (False, True):
    acc = {}
    def t(name, fn):
        t0 = time.perf_counter(); r = fn(); acc.setdefault(name, []).append(time.perf_counter() - t0); return r
    tot = []
    for i in range(25):
        t0 = time.perf_counter()
        t("upload", lambda: ctx.upload_ptrs(snap.n_nodes, 8, ptrs, pnd.data_ptr(), device=False, asynchronous=asyn))
        t("pods_upload", lambda: ctx.pods_upload(pods, reqs))
        t("score", lambda: ctx.score_resident())
        t("resolve", lambda: ctx.resolve_resident())
        t("download", lambda: ctx.results_download())
        tot.append(time.perf_counter() - t0)
        tm = ctx.timing()
    print("async upload" if asyn else "blocking upload", f"total {1e3*np.median(tot[5:]):.3f} ms")
    for k, vls in acc.items():
        print(f"  {k:12s} {1e3*np.median(vls[5:]):.3f} ms")
    print("  device phases:", {f: round(getattr(tm, f), 4) for f, _ in tm._fields_})
# resident: pods_upload + score + resolve + download only
tot = []
for i in range(25):
    t0 = time.perf_counter(); ctx.pods_upload(pods, reqs); t1 = time.perf_counter(); ctx.score_resident(); ctx.resolve_resident(); t2 = time.perf_counter(); ctx.results_download(); tot.append((t1 - t0, t2 - t1, time.perf_counter() - t2))
a = np.median(np.array(tot[5:]), axis=0) * 1e3
print(f"resident: pods_upload {a[0]:.3f}  score+resolve (queue) {a[1]:.3f}  download(sync) {a[2]:.3f}  total {a.sum():.3f} ms")
