#!/bin/bash
# shade kernel visiting the queue in walk order (MCPT_OPT=128) against queue order
O=gpurun_out; mkdir -p $O
bash tools/ab_env.sh MCPT_OPT=0 MCPT_OPT=128 2>&1 | tee $O/r2v_ab_shade_order.log
MCPT_OPT=128 timeout 600 python -m pytest tests/test_gpu_headline.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -3
