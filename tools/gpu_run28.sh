#!/bin/bash
# shade kernel: resident blocks per SM (launch bounds), grid size of the grid-stride launch, w-lane conversions skipped
O=gpurun_out; mkdir -p $O; : > $O/r3v_shade_variants.log
run() { python tools/scene_rates.py "$1" 2>&1 | tail -1 | tee -a $O/r3v_shade_variants.log; }
for g in 4 8 16 32; do MCPT_SHADE_GRID=$g run "in-tree grid $g/SM"; done
for g in 5 10 20; do MCPT_SHADE_GRID=$g MCPT_LIB=$PWD/build/ab/libmcpt_lb5.so run "5 blocks/SM (48 regs) grid $g/SM"; done
for g in 6 12 24; do MCPT_SHADE_GRID=$g MCPT_LIB=$PWD/build/ab/libmcpt_lb6.so run "6 blocks/SM (40 regs) grid $g/SM"; done
MCPT_LIB=$PWD/build/ab/libmcpt_w11.so run "w lanes skipped, grid 8/SM"
run "in-tree grid 8/SM again"
