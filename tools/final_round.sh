#!/bin/bash
# last check of a tree on the GPU box: whole GPU suite, smoke, the drain's figures, the bench line
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for c in cfg3 cfg4; do timeout 200 python tools/drain_probe.py $c 100000 4096 0 2>&1 | tail -1; done
timeout 200 python tools/drain_same_probe.py 2>&1 | tail -3
timeout 300 python bench.py > gpurun_out/r02z_bench_1gpu.json 2> gpurun_out/bench.err; tail -c 400 gpurun_out/bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02z_bench_1gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], json.dumps(d.get('drain'))[-330:])
P
