#!/bin/bash
# the configs leg under several tunings, alternating on ONE box: tools/ab_cfgs.sh <rounds> "<tuning>" ...
R=$1; shift
for i in $(seq 1 $R); do
  for X in "$@"; do
    python bench.py --tuning "$X" --no-cpu --no-e2e --no-hbm-leg --no-latency --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$X', round(d['ms_per_step'],4), {k: round(v['ms_per_step'],4) for k,v in d['configs'].items()})"
  done
done
