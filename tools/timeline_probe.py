"""Real (overlapped) kernel timeline of one cfg3 step through torch.profiler (CUPTI): start / duration / stream of
every kernel between the K0 launch and K3's end."""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import importlib
v = importlib.import_module("4paradigm-k8s-vgpu-scheduler_b200")
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
tuning = dict((k, int(x)) for k, x in (kv.split("=") for kv in sys.argv[2].split(","))) if len(sys.argv) > 2 else None
snap, batch = v.workload.make_config(cfg)
pods, reqs = v.capi.pack_batch(batch, snap.type_names)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
with v.Context(max_nodes=snap.n_nodes, max_pods=len(pods), max_reqs=max(len(reqs), len(pods)), max_picks=int(reqs["nums"].sum()), tuning=tuning) as ctx:
    ctx.upload(snap); ctx.pods_upload(pods, reqs)
    for _ in range(5):
        ctx.refresh(); ctx.score_resident(); ctx.resolve_resident(); ctx.sync()
    flush.zero_(); torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        ctx.refresh(); ctx.score_resident(); ctx.resolve_resident(); ctx.sync()
        torch.cuda.synchronize()
    f = tempfile.mktemp(suffix=".json"); prof.export_chrome_trace(f)
ev = [e for e in json.load(open(f))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
for e in ev:
    print(f"{e['ts']-t0:9.1f} +{e['dur']:7.1f}us  stream {e['args'].get('stream','?'):>3}  {e['name'][:70]}")
print("span", ev[-1]["ts"] + ev[-1]["dur"] - t0)
