#!/bin/bash
# Bounds / protocol evidence in place of compute-sanitizer (closed on this GPU pool: "runs under it have left GPUs
# needing a reset"): builds the library with -DMCPT_CHECKED=1 (every node / leaf / stack / permutation / queue / path
# index compared with its bound before use, treelet arrival protocol verified) and runs every kernel family
# (tools/sanitize_targets.py: each target also compares its result with its sequential counterpart), then a
# determinism stress of the cross-SM protocols in the PRODUCT build (tools/stress_determinism.py).
# usage: tools/checked_run.sh [out_file]     (GPU box; the checked library is built here by nvcc)
O=${1:-gpurun_out/checked_r2.txt}; mkdir -p "$(dirname $O)"
[ -f build/ab/libmcpt_checked.so ] || bash tools/build_variant.sh checked -DMCPT_CHECKED=1 > /dev/null || exit 1
echo "checked build (MCPT_CHECKED=1) on $(nvidia-smi --query-gpu=name --format=csv,noheader | head -1)" > $O
for t in treelet lbvh fasttree wave sorted chain mis merge stage metrics; do
	MCPT_LIB=build/ab/libmcpt_checked.so timeout 600 python tools/sanitize_targets.py $t 1.0 >> $O 2>&1 || echo "target $t FAILED rc=$?" >> $O
done
echo "--- determinism stress (product build)" >> $O
timeout 900 python tools/stress_determinism.py >> $O 2>&1 || echo "stress FAILED rc=$?" >> $O
cat $O
