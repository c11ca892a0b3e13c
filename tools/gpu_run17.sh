#!/bin/bash
# per-warp runs of consecutive work items in the persistent walk (MCPT_WARP_CHUNK) against the global cursor
O=gpurun_out; mkdir -p $O
bash tools/ab_lib.sh "" build/ab/libmcpt_chunk256.so build/ab/libmcpt_chunk512.so build/ab/libmcpt_chunk1024.so build/ab/libmcpt_chunk2048.so 2>&1 | tee $O/r2x_ab_warp_chunk.log
MCPT_LIB=$PWD/build/ab/libmcpt_chunk512.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
