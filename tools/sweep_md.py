"""profiles/*_sweep.md from the JSON lines of `python bench.py --sweep`
usage: sweep_md.py sweep.jsonl "<title line>" > sweep.md"""
import json, sys
def t(s):
    return f"{s*1e6:.0f} µs" if s < 1e-3 else f"{s*1e3:.2f} ms" if s < 1 else f"{s:.1f} s"
def x(r):
    return f"{r:.2g}x" if r < 10 else f"{r:.0f}x" if r < 1000 else f"{r:.1e}x"
cells = [json.loads(l) for l in open(sys.argv[1]) if l.strip()]
print(f"# {sys.argv[2]}\n")
print("`python bench.py --sweep` — GPU = wall clock of the C ABI calls (median); CPU arms = oracle restatements on bounded\n"
      "samples scaled to the cell (`cpu_sample_pods` in the `.jsonl`); the reference Go cannot run here. Config-3 pod mix.\n")
print("| nodes | pods | GPU resident (C ABI call) | GPU incl. snapshot upload | CPU R1 (1 thread, no log I/O) | CPU R1-io (with the reference's log lines) "
      f"| CPU R2 ({cells[0]['cpu_threads']} threads, packed) | GPU resident vs R1 | vs R1-io | vs R2 |")
print("|---|---|---|---|---|---|---|---|---|---|")
for c in cells:
    g = c["gpu_resident_s"]
    print(f"| {c['nodes']} | {c['pods']} | {t(g)} | {t(c['gpu_with_upload_s'])} | {t(c['cpu_r1_s'])} | {t(c['cpu_r1io_s'])} | {t(c['cpu_r2_s'])} "
          f"| {x(c['cpu_r1_s']/g)} | {x(c['cpu_r1io_s']/g)} | {x(c['cpu_r2_s']/g)} |")
