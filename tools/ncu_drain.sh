#!/bin/bash
# ncu captures of the batch-scale drain (called by tools/ncu_round.sh; alone: tools/ncu_drain.sh <tag>)
T=$1
M=gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
set -x
# 5. the batch-scale drain: launch list of a 1024-pod drain at 100k nodes, --set full of one launch of each kernel
python tools/drain_probe.py cfg3 100000 1024 0 > gpurun_out/${T}_drain_plain.log 2>&1 || exit 1
ncu --metrics $M --clock-control none -k regex:k_drain -c 130 --csv --log-file gpurun_out/${T}_drain_launches.csv \
    python tools/drain_probe.py cfg3 100000 1024 0 > gpurun_out/${T}_ncu6.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_drain -s 12 -c 4 -o gpurun_out/${T}_drain \
    python tools/drain_probe.py cfg3 100000 1024 0 > gpurun_out/${T}_ncu7.log 2>&1
ls -la gpurun_out/${T}_*
