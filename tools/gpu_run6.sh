#!/bin/bash
# round-2 GPU session 6: checked build over every kernel family + determinism stress; MIS modes probe; MIS tests
O=gpurun_out; mkdir -p $O
( time bash tools/checked_run.sh $O/checked_r2.txt ) > $O/r2k_checked.log 2>&1
tail -25 $O/r2k_checked.log
( timeout 900 python tools/mis_accept_probe.py veach_mis 18 10 16; timeout 600 python tools/mis_accept_probe.py cbox 17 9 16 128 128 8 ) > $O/r2k_mis_probe.log 2>&1
cat $O/r2k_mis_probe.log
( time timeout 900 python -m pytest tests/test_gpu_mis.py tests/test_abi.py -q ) > $O/r2k_pytest.log 2>&1
tail -5 $O/r2k_pytest.log
