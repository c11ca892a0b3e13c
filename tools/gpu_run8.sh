#!/bin/bash
# round-2 GPU session 8: full GPU suite after the hot-file clean-up + refill threshold 16, bench line, launch list, full ncu captures (walk + shade) at the bench configuration
O=gpurun_out; mkdir -p $O
( time timeout 1700 python -m pytest tests -m gpu -x -q --durations=8 ) > $O/r2m_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2m_pytest.log
tail -14 $O/r2m_pytest.log
timeout 900 python bench.py > $O/r2m_bench.json 2> $O/r2m_bench.err; echo "bench rc=$?" >> $O/r2m_bench.err
python -c "
import json; d=json.loads(open('$O/r2m_bench.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ('value','ms_per_step')}, d['e2e'], d['roofline'].get('frac'), d.get('reference_seeding',{}).get('chain'))"
B="python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline --microbench-rays 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/r2m_launches.csv $B > $O/r2m_ncu_launches.log 2>&1
python tools/launch_breakdown.py $O/r2m_launches.csv > $O/r2m_breakdown.txt; cat $O/r2m_breakdown.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_trace_persistent -s 40 -c 1 -o $O/r2m_walk -f $B > $O/r2m_ncu_walk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_shade -s 40 -c 1 -o $O/r2m_shade -f $B > $O/r2m_ncu_shade.log 2>&1
ls -la $O | grep r2m
