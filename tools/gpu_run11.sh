#!/bin/bash
# wave size sweep: samples in flight 8 / 16 / 32 / 64 (denser ray population per sort cell)
for s in 8 16 32 64; do
  timeout 600 python bench.py --no-cpu-baseline --no-extras --steps 3 --microbench-rays 0 --spp-per-step $s --samples-in-flight $s | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']
print('in flight %3d: bench %7.1f e2e %7.1f ms/step %7.1f walk ms/step %7.1f launches %d' % ($s, d['value'], d['e2e']['value'], d['ms_per_step'], r['kernel_ms_per_step'], d['gpu_launches']))"
done 2>&1 | tee gpurun_out/r2q_wave_size.log
