#!/bin/bash
for kb in 0 16 64 128 256 512 768 1024 1280 1536 2048 3072 4096 8192; do
  python bench.py --tuning alloc_pad_kb=$kb --no-cpu --no-e2e --no-hbm-leg --no-configs --no-latency --steps 30 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($kb, round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4))"
done
