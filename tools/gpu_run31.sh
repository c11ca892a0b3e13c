#!/bin/bash
# round-3 GPU session 25: final state (shader divisions through double-precision reciprocals, 32 shade blocks per SM) - full suite, smoke, both bench arms at the defaults (128 samples in flight), launch list + full ncu captures
O=gpurun_out; mkdir -p $O
( time timeout 1700 python -m pytest tests -m gpu -x -q --durations=6 ) > $O/r3y_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r3y_pytest.log; tail -12 $O/r3y_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > $O/r3y_bench.json 2> $O/r3y_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r3y_bench_reference.json 2> $O/r3y_bench_reference.err; echo "reference arm rc=$?"; tail -c 400 $O/r3y_bench_reference.json
python -c "
import json; d=json.loads(open('$O/r3y_bench.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d.get('scene_load'))"
B="python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline --microbench-rays 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r3y_launches.csv $B > $O/r3y_ncu_launches.log 2>&1
python tools/launch_breakdown.py $O/r3y_launches.csv > $O/r3y_breakdown.txt; cat $O/r3y_breakdown.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_trace_persistent -s 24 -c 1 -o $O/r3y_walk -f $B > $O/r3y_ncu_walk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_shade -s 24 -c 1 -o $O/r3y_shade -f $B > $O/r3y_ncu_shade.log 2>&1
ls -la $O | grep r3y
python tools/scene_rates.py in-tree 2>&1 | tail -1 | tee $O/r3y_scene_rates.log
