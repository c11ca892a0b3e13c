#!/bin/bash
# shade kernel: grid size of the grid-stride launch, continued
O=gpurun_out; mkdir -p $O; : > $O/r3w_shade_grid.log
run() { python tools/scene_rates.py "$1" 2>&1 | tail -1 | tee -a $O/r3w_shade_grid.log; }
for g in 8 24 32 48 64 128 512 65536; do MCPT_SHADE_GRID=$g run "shade grid $g/SM"; done
