#!/bin/bash
# round-2 8-GPU session: the north-star target run (diningroom 1080p x 4096 spp) - one process per device (--procs 8,
# cache miss then hit) against one process with a thread per device (--gpus 8); bench.py under torchrun at N = 8.
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader > $O/r2r_gpus.txt; nproc >> $O/r2r_gpus.txt
rm -f assets/diningroom/*.mcptcache
L=$O/r2r_target_8gpu.log; : > $L
run() { echo "### $*" >> $L; local t0=$(date +%s%N); "$@" >> $L 2>&1; local rc=$?; echo "wall $(( ($(date +%s%N) - t0) / 1000000 )) ms (rc $rc)" >> $L; }
run build/mcpt --configid 3 --procs 8 --spp 4095 --out $O/r2r_procs_miss
run build/mcpt --configid 3 --procs 8 --spp 4095 --out $O/r2r_procs_hit
run build/mcpt --configid 3 --gpus 8 --spp 4095 --out $O/r2r_threads
run build/mcpt --configid 3 --gpus 8 --spp 4095 --overlap-init --out $O/r2r_threads_overlap
run build/mcpt --configid 3 --procs 8 --spp 4095 --mis-ratio --out $O/r2r_procs_misratio
run build/mcpt --configid 3 --order-independent-seeds --spp 511 --out $O/r2r_one_gpu_512
cmp $O/r2r_procs_miss.hdr $O/r2r_procs_hit.hdr && echo "procs miss == procs hit (bytes)" >> $L
python tools/compare_hdr.py $O/r2r_procs_hit.hdr $O/r2r_threads.hdr >> $L 2>&1
grep -E "^###|^\{|wall|==|differ|pixels" $L
rm -f $O/*.hdr            # 8 MB each: keep the PNGs of two of them only
ls $O/*.png | tail -n +3 | xargs rm -f
[ -n "$SKIP_BENCH" ] || timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 3 --warmup 3 > $O/r2r_bench_n8.json 2> $O/r2r_bench_n8.err
[ -n "$SKIP_BENCH" ] || python -c "
import json; d=json.loads(open('$O/r2r_bench_n8.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d.get('strong_scaling'))"
