"""Per-source-line instruction counts of one kernel: aligns the SASS rows of an `ncu --page source --csv`
dump with `nvdisasm -g` line markers of the same cubin (same instruction order).
usage: ncu_lines.py <ncu_source.csv> <nvdisasm_with_lines.txt> [divisor]"""
import csv, re, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
H = rows[1]
ia = H.index("Instructions Executed")
counts = [int(r[ia]) for r in rows[2:] if len(r) > ia]
div = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
cur = None
lines = []
for ln in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"^\s*/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
assert len(lines) == len(counts), (len(lines), len(counts))
agg = defaultdict(int)
for l, c in zip(lines, counts):
    agg[l] += c
tot = sum(counts)
print("total", tot, "per unit", tot / div)
for l, c in sorted(agg.items(), key=lambda x: (x[0] or ("", 0))):
    if c > tot * 0.003:
        print(f"{100*c/tot:5.1f}% {c/div:8.1f}  {l}")
