#!/bin/bash
# compute-sanitizer over small invocations of every kernel family (GPU box): memcheck (out-of-bounds / misaligned
# accesses, leaks of device memory), racecheck (shared-memory hazards: the block-level queues and the treelet warps'
# staging buffers), synccheck (divergent barriers) and initcheck (reads of never-written global memory).
# Cross-SM ordering through global memory (treelet arrival counters + __threadfence, the persistent kernels' work
# cursors) is outside what racecheck models; those are covered by the byte-equality of every target with its
# sequential counterpart, here under the sanitizer's changed timing and in the regular GPU tests.
# usage: tools/sanitize.sh [out_dir] ; summary -> <out_dir>/sanitizer_summary.txt (copy to profiles/)
O=${1:-gpurun_out/sanitizer}; mkdir -p $O
S=$O/sanitizer_summary.txt
echo "compute-sanitizer $(compute-sanitizer --version | tail -1)" > $S
nvidia-smi --query-gpu=name --format=csv,noheader >> $S
run() {  # tool target scale [extra sanitizer flags]
	local tool=$1 target=$2 scale=$3; shift 3
	local log=$O/${tool}_${target}.log
	local t0=$(date +%s)
	timeout 900 compute-sanitizer --tool $tool "$@" --error-exitcode 9 python tools/sanitize_targets.py $target $scale > $log 2>&1
	local rc=$?
	local t1=$(date +%s)
	printf "%-10s %-8s scale %-5s rc=%d %4ds  %s | %s\n" $tool $target $scale $rc $((t1 - t0)) "$(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1)" "$(grep -c 'target .* ok' $log) ok" >> $S
}
for t in treelet lbvh fasttree wave chain mis merge stage metrics; do run memcheck $t 1.0 --leak-check full; done
for t in treelet lbvh fasttree wave chain mis merge stage metrics; do run racecheck $t 0.25; done
for t in treelet wave chain mis; do run synccheck $t 0.25; done
for t in treelet lbvh fasttree wave chain mis merge; do run initcheck $t 0.25; done
cat $S
