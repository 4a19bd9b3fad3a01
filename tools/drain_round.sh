#!/bin/bash
# drain tests + timing of the batch-scale drain and of its two kernels (ncu launch list)
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "drain" 2>&1 | tail -5
VGS_DRAIN_DEBUG=1 timeout 200 python tools/drain_probe.py cfg3 100000 4096 0 -1 2>&1 | tail -4
timeout 200 python tools/drain_probe.py cfg4 100000 4096 0 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_drain -c 192 --csv --log-file gpurun_out/drain_launches.csv python tools/drain_probe.py cfg3 100000 4096 0 > gpurun_out/drain_ncu.log 2>&1
python - <<EOF
import csv, collections
rows=list(csv.reader(open("gpurun_out/drain_launches.csv")))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
H=rows[hi]
d=collections.defaultdict(list)
for r in rows[hi+1:]:
    x=dict(zip(H,r)); d[x["Kernel Name"][:30]].append(float(x["Metric Value"].replace(",","")))
for k,v in d.items(): print(k, len(v), sum(v)/len(v), min(v), max(v), x["Metric Unit"])
EOF
