#!/bin/bash
# round-2 GPU session 4: new tests (scene cache / CLI / tolerance), compute-sanitizer sweep
O=gpurun_out; mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_cli.py tests/test_gpu_edges.py tests/test_gpu_tolerance.py -q --durations=10 ) > $O/r2i_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2i_pytest.log
tail -30 $O/r2i_pytest.log
( time bash tools/sanitize.sh $O/sanitizer ) > $O/r2i_sanitize.log 2>&1
tail -40 $O/r2i_sanitize.log
