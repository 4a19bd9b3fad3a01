#!/bin/bash
# ncu --set full on one batch launch of a kernel (regex $1) of `bench.py --config $2`; report into gpurun_out/$3
ncu --set full --import-source on --clock-control none -k regex:$1 -s ${4:-1} -c 1 -o gpurun_out/$3 python bench.py --config $2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg > gpurun_out/ncu.log 2>&1
