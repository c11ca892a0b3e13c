#!/bin/bash
# round-2 GPU session 9: scaled throughput division (bit-exactness tests + A/B against the unscaled build)
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_gpu_fullsize.py -x -q ) > $O/r2n_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2n_pytest.log; tail -4 $O/r2n_pytest.log
bash tools/ab_lib.sh "" build/ab/libmcpt_noscale.so 2>&1 | tee $O/r2n_ab_scaled_div.log
timeout 600 python bench.py --no-cpu-baseline --steps 5 > $O/r2n_bench.json 2>$O/r2n_bench.err
python -c "
import json; d=json.loads(open('$O/r2n_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['scenes'], d['reference_seeding'])"
