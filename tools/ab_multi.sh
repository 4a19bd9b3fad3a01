#!/bin/bash
# several tuning variants alternating on ONE box: tools/ab_multi.sh <rounds> "<tuning A>" "<tuning B>" ...
R=$1; shift
for i in $(seq 1 $R); do
  for X in "$@"; do
    python bench.py --tuning "$X" --no-cpu --no-e2e --no-hbm-leg --no-configs --no-latency --steps 30 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$X', round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4), d['check']['bit_exact_vs_oracle'])"
  done
done
