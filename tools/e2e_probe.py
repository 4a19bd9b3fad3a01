import sys, time, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
snap, batch = v.workload.make_config("cfg3")
pods, reqs = v.capi.pack_batch(batch, snap.type_names)
ctx = v.Context(max_nodes=snap.n_nodes, max_pods=4096, max_reqs=4096, max_picks=8192)
pin = torch.from_numpy(snap.planes).pin_memory(); pnd = torch.from_numpy(snap.ndev).pin_memory()
ptrs = [pin[f].data_ptr() for f in range(8)]
acc = {}
def t(name, fn):
    t0 = time.perf_counter(); r = fn(); acc.setdefault(name, []).append(time.perf_counter() - t0); return r
for i in range(25):
    t("upload", lambda: ctx.upload_ptrs(snap.n_nodes, 8, ptrs, pnd.data_ptr(), device=False))
    t("pods_upload", lambda: ctx.pods_upload(pods, reqs))
    t("score", lambda: ctx.score_resident())
    t("resolve", lambda: ctx.resolve_resident())
    t("download", lambda: ctx.results_download())
    tm = ctx.timing()
for k, vls in acc.items():
    print(f"{k:12s} {1e3*np.median(vls[5:]):.3f} ms")
print("timing:", {f: round(getattr(tm, f), 4) for f, _ in tm._fields_})
