#!/bin/bash
# round-2 GPU session 1: tests, bench line, sort probe, ncu launch list + full capture at the bench configuration
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt
nproc >> $O/gpu.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2a_pytest.log
timeout 900 python bench.py > $O/r2a_bench.json 2> $O/r2a_bench.err
echo "bench rc=$?" >> $O/r2a_bench.err
timeout 600 python tools/sort_probe.py diningroom 3 8 > $O/r2a_sort_probe.log 2>&1
timeout 600 python tools/sort_probe.py diningroom 9 8 >> $O/r2a_sort_probe.log 2>&1
B="python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline --microbench-rays 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2a_launches.csv $B > $O/r2a_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_trace_persistent -s 40 -c 1 -o $O/r2a_walk -f $B > $O/r2a_ncu_walk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_shade -s 40 -c 1 -o $O/r2a_shade -f $B > $O/r2a_ncu_shade.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sectors_op_read.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,dram__bytes_read.sum --clock-control none -k regex:k_read_sweep -c 12 --csv --log-file $O/r2a_l2probe.csv python -c "
from montecarlopathtracing_b200 import device
print(device.measure_read_bandwidth(32<<20, iters=40))" > $O/r2a_l2probe.log 2>&1
ls -la $O
