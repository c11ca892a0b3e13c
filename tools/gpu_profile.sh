#!/bin/bash
# usage: tools/gpu_profile.sh <tag> [env...]  (GPU box): launch list + full ncu capture of the walk kernel at the bench configuration
T=$1; shift; O=gpurun_out; mkdir -p $O
B="env $* python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline --microbench-rays 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/${T}_launches.csv $B > $O/${T}_ncu_launches.log 2>&1
python tools/launch_breakdown.py $O/${T}_launches.csv > $O/${T}_breakdown.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_trace_persistent -s 40 -c 1 -o $O/${T}_walk -f $B > $O/${T}_ncu_walk.log 2>&1
cat $O/${T}_breakdown.txt
