#!/bin/bash
# last check of a tree on the GPU box: whole GPU suite, smoke, the bench line
timeout 500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/r02z_bench_1gpu.json 2> gpurun_out/bench.err; tail -c 400 gpurun_out/bench.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02z_bench_1gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], {k: (d[k]['ms_per_step'], d[k].get('h2d_bytes_per_step')) for k in ('e2e','e2e_int32_planes','e2e_usage_planes','e2e_delta')}, d['drain']['ms'])
P
