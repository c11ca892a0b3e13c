#!/bin/bash
# round-2 GPU session 5: CLI tests again, MIS with the glossy-lobe NEE (tests + acceptance probe)
O=gpurun_out; mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_cli.py tests/test_gpu_mis.py -q --durations=10 ) > $O/r2j_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2j_pytest.log
tail -30 $O/r2j_pytest.log
( timeout 600 python tools/mis_accept_probe.py veach_mis 19 14; timeout 600 python tools/mis_accept_probe.py veach_mis 20 15 160 90 8; timeout 600 python tools/mis_accept_probe.py cbox 17 12 128 128 8; timeout 300 python tools/mis_probe.py veach_mis 32768 2048 ) > $O/r2j_mis_probe.log 2>&1
cat $O/r2j_mis_probe.log
