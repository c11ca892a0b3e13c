#!/bin/bash
# 8-GPU box: staggered one-process-per-device start (--procs 8 --overlap-init) against the static modes
O=gpurun_out; mkdir -p $O
L=$O/r2u_target_8gpu.log; : > $L
run() { echo "### $*" >> $L; local t0=$(date +%s%N); "$@" >> $L 2>&1; local rc=$?; echo "wall $(( ($(date +%s%N) - t0) / 1000000 )) ms (rc $rc)" >> $L; }
run build/mcpt --configid 3 --procs 8 --overlap-init --spp 4095 --out $O/r2u_stag_miss
run build/mcpt --configid 3 --procs 8 --overlap-init --spp 4095 --out $O/r2u_stag_hit
run build/mcpt --configid 3 --procs 8 --overlap-init --spp 4095 --out $O/r2u_stag_hit2
run build/mcpt --configid 3 --gpus 8 --spp 4095 --out $O/r2u_threads
python tools/compare_hdr.py $O/r2u_stag_hit.hdr $O/r2u_threads.hdr >> $L 2>&1
python tools/compare_hdr.py $O/r2u_stag_hit.hdr $O/r2u_stag_hit2.hdr >> $L 2>&1
grep -E "^###|^\{|wall|==" $L
rm -f $O/r2u_*.hdr $O/r2u_*.png
