#!/bin/bash
# development A/B of environment switches with bench.py, interleaved twice: tools/ab_env.sh "MCPT_OPT=0" "MCPT_OPT=4" ...
for rep in 1 2; do
for e in "$@"; do
  env $e timeout 300 python bench.py --no-cpu-baseline --no-extras --steps 4 --microbench-rays 4194304 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); m=d['microbench']
print('%-34s bench %7.1f e2e %7.1f coherent %7.1f random %7.1f' % ('$e', d['value'], d['e2e']['value'], m['coherent']['mrays_per_s'], m['random']['mrays_per_s']))"
done; done
