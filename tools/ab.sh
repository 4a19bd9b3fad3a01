#!/bin/bash
# A/B on ONE box: alternate `bench.py $COMMON $A` and `bench.py $COMMON $B`, 3 rounds; prints ms_per_step and K1 ms
# usage: tools/ab.sh "<common args>" "<args A>" "<args B>" [rounds]
R=${4:-3}
for i in $(seq 1 $R); do
  for v in A B; do
    if [ $v = A ]; then X="$2"; else X="$3"; fi
    python bench.py $1 $X --no-cpu --no-e2e --no-hbm-leg --no-configs 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', '$X', round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4), d['check']['bit_exact_vs_oracle'])"
  done
done
