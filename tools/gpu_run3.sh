#!/bin/bash
# round-2 GPU session 3: full GPU test suite + default bench line (state check after the sort / drop-in commits)
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt; nproc >> $O/gpu.txt
( time timeout 1700 python -m pytest tests -m gpu -x -q --durations=10 ) > $O/r2h_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2h_pytest.log
tail -5 $O/r2h_pytest.log
timeout 900 python bench.py > $O/r2h_bench.json 2> $O/r2h_bench.err
echo "bench rc=$?" >> $O/r2h_bench.err
tail -c 3000 $O/r2h_bench.json
