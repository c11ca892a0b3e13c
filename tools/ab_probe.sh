#!/bin/bash
# development A/B runs of the probe under different MCPT_OPT / MCPT_BLOCK settings
for cfg in "0 128" "1 128" "0 64" "0 256" "1 64"; do
  set -- $cfg
  echo "=== MCPT_OPT=$1 MCPT_BLOCK=$2"
  MCPT_OPT=$1 MCPT_BLOCK=$2 python tools/gpu_probe.py cbox diningroom 2>&1 | grep -E "fast:|render"
done
