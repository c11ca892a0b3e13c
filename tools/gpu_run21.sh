#!/bin/bash
# round-2 GPU session 21: final state - full suite, smoke, both bench arms at the defaults (128 samples in flight), launch list + full ncu captures
O=gpurun_out; mkdir -p $O
( time timeout 1700 python -m pytest tests -m gpu -x -q --durations=6 ) > $O/r3i_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r3i_pytest.log; tail -12 $O/r3i_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > $O/r3i_bench.json 2> $O/r3i_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r3i_bench_reference.json 2> $O/r3i_bench_reference.err; echo "reference arm rc=$?"; tail -c 400 $O/r3i_bench_reference.json
python -c "
import json; d=json.loads(open('$O/r3i_bench.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d.get('scene_load'))"
B="python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline --microbench-rays 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r3i_launches.csv $B > $O/r3i_ncu_launches.log 2>&1
python tools/launch_breakdown.py $O/r3i_launches.csv > $O/r3i_breakdown.txt; cat $O/r3i_breakdown.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_trace_persistent -s 24 -c 1 -o $O/r3i_walk -f $B > $O/r3i_ncu_walk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_shade -s 24 -c 1 -o $O/r3i_shade -f $B > $O/r3i_ncu_shade.log 2>&1
ls -la $O | grep r3i
