#!/bin/bash
# 8-GPU box: CUDA start-up breakdown of a process (all devices visible / one visible / eight single-device processes at
# once), then the CLI target runs again with create / upload timing and the non-blocking worker collection.
O=gpurun_out; mkdir -p $O
P=$O/r2s_startup_probe.log; : > $P
echo "# all devices visible" >> $P; build/startup_probe >> $P 2>&1
echo "# CUDA_VISIBLE_DEVICES=0" >> $P; CUDA_VISIBLE_DEVICES=0 build/startup_probe >> $P 2>&1
echo "# again (driver state warm?)" >> $P; CUDA_VISIBLE_DEVICES=0 build/startup_probe >> $P 2>&1
echo "# eight single-device processes at once" >> $P
for g in 0 1 2 3 4 5 6 7; do CUDA_VISIBLE_DEVICES=$g build/startup_probe >> $P 2>&1 & done; wait
nvidia-smi -q | grep -i -m3 "persistence" >> $P
cat $P
L=$O/r2s_target_8gpu.log; : > $L
run() { echo "### $*" >> $L; local t0=$(date +%s%N); "$@" >> $L 2>&1; local rc=$?; echo "wall $(( ($(date +%s%N) - t0) / 1000000 )) ms (rc $rc)" >> $L; }
run build/mcpt --configid 3 --procs 8 --spp 4095 --out $O/r2s_procs
run build/mcpt --configid 3 --gpus 8 --spp 4095 --out $O/r2s_threads
run build/mcpt --configid 3 --gpus 1 --order-independent-seeds --spp 63 --out $O/r2s_one
cmp $O/r2s_procs.hdr $O/r2s_threads.hdr && echo "procs == threads (bytes)" >> $L
grep -E "^###|^\{|wall|==" $L
rm -f $O/r2s_*.hdr $O/r2s_*.png
