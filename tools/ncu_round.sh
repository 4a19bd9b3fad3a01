#!/bin/bash
# The ncu captures of a round, in one gpurun call (each bench command has exited 0 without ncu in the same call).
# usage: tools/ncu_round.sh <tag>       -> gpurun_out/<tag>_*
T=$1
M=gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
set -x
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err || exit 1
# 1. launch list of one cfg3 step (the headline leg)
ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_ncu1.log 2>&1
# 2. the single-pod 1M-node score kernel (HBM-bound regime): last launches of k1_score_single
ncu --metrics $M --clock-control none -k regex:k1_score_single -c 60 --csv --log-file gpurun_out/${T}_p1_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-configs --no-latency > gpurun_out/${T}_ncu2.log 2>&1
# 3. cfg4: launch list, and --set full of the walk kernel
ncu --metrics $M --clock-control none -c 80 --csv --log-file gpurun_out/${T}_cfg4_launches.csv \
    python bench.py --config cfg4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_ncu3.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k1_score_walk -s 1 -c 1 -o gpurun_out/${T}_walk \
    python bench.py --config cfg4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_ncu4.log 2>&1
# 4. --set full of the sliced launches of the last step: the two pod-sliced mains and the node-sliced launches (scouts, odd words)
ncu --set full --import-source on --clock-control none -k regex:k1_score_sliced -s 6 -c 2 -o gpurun_out/${T}_sliced \
    python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_ncu5.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k1_score_nodesliced -s 9 -c 3 -o gpurun_out/${T}_nodesliced \
    python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-hbm-leg --no-configs > gpurun_out/${T}_ncu5b.log 2>&1
ls -la gpurun_out/${T}_*
# 5. the batch-scale drain
bash tools/ncu_drain.sh $T
