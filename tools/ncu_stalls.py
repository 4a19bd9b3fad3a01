"""Per-source-line warp stall samples of one kernel (barrier waits left out): aligns the SASS rows of an
`ncu --page source --csv` dump with the `nvdisasm -g` listing of the same function.
usage: ncu_stalls.py <ncu_source.csv> <nvdisasm_function.txt> [top]"""
import csv, re, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
H = rows[1]
idx = {h: i for i, h in enumerate(H)}
data = rows[2:]
cur = None
lines = []
for ln in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    if re.match(r"^\s*/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
assert len(lines) == len(data), (len(lines), len(data))
stalls = [h for h in H if h.startswith('stall_') and 'Not Issued' not in h]
agg = defaultdict(lambda: defaultdict(int))
inst = defaultdict(int)
tot = defaultdict(int)
for l, r in zip(lines, data):
    for st in stalls:
        v = int(r[idx[st]] or 0)
        agg[l][st] += v
        tot[st] += v
    inst[l] += int(r[idx['Instructions Executed']] or 0)
print({k: v for k, v in tot.items() if v})
ex = [(sum(v for k, v in d.items() if k != 'stall_barrier'), l) for l, d in agg.items()]
T = sum(x for x, _ in ex)
print('non-barrier samples', T, 'instructions', sum(inst.values()))
for x, l in sorted(ex, reverse=True)[:int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    d = agg[l]
    top = sorted(((v, k) for k, v in d.items() if k != 'stall_barrier' and v), reverse=True)[:3]
    print(f"{100*x/T:5.1f}% {x:6d} inst {inst[l]:7d} {l} {top}")
