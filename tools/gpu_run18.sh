#!/bin/bash
# round-2 GPU session 18: device fast-tree builder in the upload path - full suite, tree probe, bench line
O=gpurun_out; mkdir -p $O
( time timeout 1700 python -m pytest tests -m gpu -x -q --durations=5 ) > $O/r3a_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r3a_pytest.log; tail -10 $O/r3a_pytest.log
python tools/tree_probe.py 2>&1 | tee $O/r3a_tree_probe.log | tail -12
timeout 900 python bench.py --no-cpu-baseline > $O/r3a_bench.json 2> $O/r3a_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('$O/r3a_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['cold_upload_s'], json.dumps(d.get('scene_load')))"
