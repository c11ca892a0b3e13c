#!/bin/bash
# round-2 GPU session 7: A/B of walk variants (top-of-stack register, cache hints, bins, refill threshold, small shared stack) + new MIS tests
O=gpurun_out; mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_mis.py -q ) > $O/r2l_pytest.log 2>&1
tail -5 $O/r2l_pytest.log
bash tools/ab_lib.sh "" build/ab/libmcpt_topcache.so build/ab/libmcpt_leafna.so build/ab/libmcpt_nodeel.so build/ab/libmcpt_bins32.so build/ab/libmcpt_refill16.so build/ab/libmcpt_refill26.so build/ab/libmcpt_refill30.so build/ab/libmcpt_sstack8.so 2>&1 | tee $O/r2l_ab_walk.log
