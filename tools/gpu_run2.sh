#!/bin/bash
# round-2 GPU session: parity tests, A/B of the sorted bounce queues, sort probe, ncu profiles (sort off / on), L2 probe line
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt; nproc >> $O/gpu.txt
( time timeout 1500 python -m pytest tests -m gpu -q --durations=8 --deselect tests/test_gpu_tolerance.py::test_converged_cbox_every_pixel_within_4_sigma_and_mre_below_1_percent ) > $O/r2f_pytest.log 2>&1
tail -4 $O/r2f_pytest.log
bash tools/ab_env.sh MCPT_SORT=0 MCPT_SORT=1 2>&1 | tee $O/r2f_ab_sort.log
PROBE_SET=2 timeout 300 python tools/sort_probe.py diningroom 6 8 > $O/r2f_sort_probe.log 2>&1
PROBE_SET=2 timeout 300 python tools/sort_probe.py veach_mis 2 8 >> $O/r2f_sort_probe.log 2>&1
PROBE_SET=2 timeout 300 python tools/sort_probe.py cbox 3 8 >> $O/r2f_sort_probe.log 2>&1
bash tools/gpu_profile.sh r2f_unsorted MCPT_SORT=0
bash tools/gpu_profile.sh r2f_sorted MCPT_SORT=1
timeout 300 ncu --metrics gpu__time_duration.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sectors_op_read.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,dram__bytes_read.sum --clock-control none -k regex:k_read_sweep -c 8 --csv --log-file $O/r2f_l2probe.csv python -c "
from montecarlopathtracing_b200 import device
print(device.measure_read_bandwidth(32<<20, iters=40))" > $O/r2f_l2probe.log 2>&1
