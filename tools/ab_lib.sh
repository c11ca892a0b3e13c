#!/bin/bash
# development A/B of alternative builds of the CUDA library (build/ab/*.so) with bench.py
for lib in "" build/ab/libmcpt_mb10.so build/ab/libmcpt_mb12.so; do
  echo "=== MCPT_LIB=$lib"
  MCPT_LIB=${lib:+$PWD/$lib} timeout 300 python tools/gpu_probe.py diningroom 2>&1 | grep -E "fast:|render"
  MCPT_LIB=${lib:+$PWD/$lib} timeout 300 python bench.py --no-cpu-baseline --steps 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench', d['value'], d['e2e']['value'])"
done
