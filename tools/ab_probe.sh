#!/bin/bash
# development A/B runs of the probe under different MCPT_OPT / MCPT_BLOCK settings
for cfg in ${AB_CONFIGS:-"0,128 2,128"}; do
  opt=${cfg%,*}; blk=${cfg#*,}
  echo "=== MCPT_OPT=$opt MCPT_BLOCK=$blk"
  MCPT_OPT=$opt MCPT_BLOCK=$blk timeout 300 python tools/gpu_probe.py ${AB_SCENES:-cbox diningroom} 2>&1 | grep -E "fast:|render|vs reference|Error|error"
done
