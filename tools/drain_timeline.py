"""Real (overlapped) kernel timeline of one vgs_drain_batch through torch.profiler (CUPTI)
usage: python tools/drain_timeline.py [cfg] [nodes] [pods] [first_n_events]"""
import json, os, sys, tempfile, importlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
P = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
show = int(sys.argv[4]) if len(sys.argv) > 4 else 60
snap, batch = v.workload.make_config(cfg, N, P)
with v.Context(max_nodes=N, max_devices=snap.dmax, max_pods=P, max_reqs=batch.n_reqs, max_picks=int(batch.nums.sum())) as ctx:
    ctx.upload(snap); ctx.drain(batch); ctx.upload(snap)
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        ctx.drain(batch)
        torch.cuda.synchronize()
    f = tempfile.mktemp(suffix=".json"); prof.export_chrome_trace(f)
ev = [e for e in json.load(open(f))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
for e in ev[:show]:
    print(f"{e['ts']-t0:9.1f} +{e['dur']:7.1f}us  stream {e['args'].get('stream','?'):>3}  {e['name'][:60]}")
print("span", ev[-1]["ts"] + ev[-1]["dur"] - t0, "events", len(ev))
import collections
d = collections.defaultdict(list)
for e in ev: d[e["name"][:40]].append(e["dur"])
for k, x in d.items(): print(f"{k:42s} n={len(x):4d} mean {sum(x)/len(x):8.1f} us")
