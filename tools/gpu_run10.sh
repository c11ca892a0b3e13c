#!/bin/bash
# round-2 GPU session 10: throughput division detours (scaled / double) - bit-exactness tests + A/B
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline.py -x -q ) > $O/r2p_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2p_pytest.log; tail -4 $O/r2p_pytest.log
bash tools/ab_lib.sh "" build/ab/libmcpt_scale1.so build/ab/libmcpt_noscale.so 2>&1 | tee $O/r2p_ab_div.log
