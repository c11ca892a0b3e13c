#!/bin/bash
# per-scene A/B of the shader's division variants (MCPT_DRCP_DIV bit 0 normalisations, bit 1 division by 2 pi, bit 2 the same for tiny operands only), the library before the round-3 shade changes, and the deferred pows
O=gpurun_out; mkdir -p $O; : > $O/r3q_scene_rates.log
for rep in 1 2; do
for v in old drcp0 drcp1 drcp2 drcp3 drcp4 drcp5; do MCPT_LIB=$PWD/build/ab/libmcpt_$v.so python tools/scene_rates.py $v 2>&1 | tail -1 | tee -a $O/r3q_scene_rates.log; done
MCPT_OPT=128 MCPT_LIB=$PWD/build/ab/libmcpt_drcp3.so python tools/scene_rates.py "drcp3 pows in place" 2>&1 | tail -1 | tee -a $O/r3q_scene_rates.log
MCPT_OPT=128 MCPT_LIB=$PWD/build/ab/libmcpt_drcp0.so python tools/scene_rates.py "drcp0 pows in place" 2>&1 | tail -1 | tee -a $O/r3q_scene_rates.log
done
