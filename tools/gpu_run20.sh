#!/bin/bash
O=gpurun_out; mkdir -p $O
( timeout 1500 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_headline.py tests/test_gpu_cli.py -x -q ) 2>&1 | tail -4
python tools/tree_probe.py 2>&1 | tee $O/r3d_tree_probe.log | tail -11
timeout 600 python tools/stress_large_scene.py 3000000 2>&1 | tee $O/r3d_stress_3M.log | tail -4
