"""profiles/rNN_k1_profile.json from an ncu launch list of `python bench.py --steps 1 ...`:
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
      --clock-control none -c N --csv --log-file launches.csv python bench.py --steps 1 --warmup 3 --no-cpu ...
Picks the LAST full K1 phase of the headline leg (the kernels between the last k0_prepare of the batch legs and
the k3_resolve that follows) -- here: sums per kernel name over one step, identified by --step-kernels.
usage: make_k1_profile.py launches.csv out.json <tree> [p1_launches.csv]"""
import csv, json, subprocess, sys
from collections import OrderedDict

def load(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H = rows[hi]
    launches = OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) < len(H):
            continue
        d = dict(zip(H, r))
        lid = int(d["ID"])
        e = launches.setdefault(lid, {"name": d["Kernel Name"], "grid": d.get("Grid Size", "")})
        e[d["Metric Name"]] = float(d["Metric Value"].replace(",", ""))
    return list(launches.values())

def short(n):
    n = n.replace("void ", "").replace("vgs::", "")
    return n.split("(")[0]

L = load(sys.argv[1])
# steps are delimited by k0_prepare launches over the whole snapshot; take the last complete step that is
# followed by a k3_resolve and contains sliced kernels (the headline leg)
k0 = [i for i, l in enumerate(L) if short(l["name"]).startswith("k0_prepare")]
best = None
for a, b in zip(k0, k0[1:] + [len(L)]):
    seg = L[a:b]
    names = [short(l["name"]) for l in seg]
    if any(n.startswith("k1_score_sliced") for n in names) and any(n.startswith("k3_resolve") for n in names):
        k3 = [i for i, n in enumerate(names) if n.startswith("k3_resolve")][0]
        best = seg[: k3 + 1]
        break  # the first such step after warm-up is as good as any: every step launches the same kernels
if best is None:
    raise SystemExit("no headline step found")
k1 = [l for l in best if not short(l["name"]).startswith(("k0_prepare", "k3_resolve")) and "vgs" in l["name"]]
per = OrderedDict()
for l in k1:
    e = per.setdefault(short(l["name"]) + " grid " + l["grid"], {"launches": 0, "us": 0.0, "warp_inst": 0.0, "dram_bytes": 0.0})
    e["launches"] += 1
    e["us"] += l.get("gpu__time_duration.sum", 0.0) / 1e3
    e["warp_inst"] += l.get("smsp__inst_executed.sum", 0.0)
    e["dram_bytes"] += l.get("dram__bytes_read.sum", 0.0) + l.get("dram__bytes_write.sum", 0.0)
out = {
    "tree": sys.argv[3],
    "source": "ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum "
              "--clock-control none (per-launch, serialised, cold cache) of `python bench.py --steps 1`: one cfg3 step",
    "k1_warp_instructions_per_step": sum(e["warp_inst"] for e in per.values()),
    "k1_dram_bytes_per_step": sum(e["dram_bytes"] for e in per.values()),
    "k1_serial_kernel_us": sum(e["us"] for e in per.values()),
    "k0_us": sum(l.get("gpu__time_duration.sum", 0) / 1e3 for l in best if short(l["name"]).startswith("k0_prepare")),
    "k3_us": sum(l.get("gpu__time_duration.sum", 0) / 1e3 for l in best if short(l["name"]).startswith("k3_resolve")),
    "kernels": per,
}
if len(sys.argv) > 4:
    P1 = [l for l in load(sys.argv[4]) if short(l["name"]).startswith("k1_score_single")]
    if P1:
        l = P1[-1]
        out["p1_dram_bytes_per_launch"] = l.get("dram__bytes_read.sum", 0.0) + l.get("dram__bytes_write.sum", 0.0)
        out["p1_kernel_us_under_ncu"] = l.get("gpu__time_duration.sum", 0.0) / 1e3
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "kernels"}, indent=1))
