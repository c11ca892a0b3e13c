#!/bin/bash
# development A/B of alternative builds of the CUDA library with bench.py, interleaved twice for a noise estimate
# usage (on the GPU box): tools/ab_lib.sh "" build/ab/libmcpt_x.so ...     ("" = the in-tree library)
for rep in 1 2; do
for lib in "$@"; do
  MCPT_LIB=${lib:+$PWD/$lib} timeout 300 python bench.py --no-cpu-baseline --no-extras --steps 4 --microbench-rays 4194304 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); m=d['microbench']
print('%-34s bench %7.1f e2e %7.1f coherent %7.1f random %7.1f' % ('${lib:-in-tree}', d['value'], d['e2e']['value'], m['coherent']['mrays_per_s'], m['random']['mrays_per_s']))"
done; done
