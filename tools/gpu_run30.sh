#!/bin/bash
# NEE + MIS shade kernels: grid of the grid-stride launches 8 against 32 blocks per SM; then the parity tests of the final library
O=gpurun_out; mkdir -p $O; : > $O/r3x_mis_shade_grid.log
for g in 8 32; do MCPT_SHADE_GRID=$g python - <<'PY' 2>&1 | tee -a $O/r3x_mis_shade_grid.log
import os, time, sys
sys.path.insert(0, '.')
from montecarlopathtracing_b200 import device, scenes
for name, w, h, depth, spp in (("veach_mis", 1280, 720, 16, 256), ("cbox", 1024, 1024, 8, 256), ("diningroom", 1920, 1080, 16, 64)):
    sc = scenes.load(name)
    with device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3, sampling=1) as c:
        c.upload_scene(sc.triangles, sc.materials, sc.nodes); c.set_camera(sc.camera)
        for k in range(4):
            c.render(spp // 2)
        c.reset_stats()
        best = 1e9
        for rep in range(3):
            t = time.time(); c.render(spp); best = min(best, time.time() - t)
        print("shade grid %s/SM, NEE + MIS %s: arrangement %d, %.0f Mrays/s" % (os.environ["MCPT_SHADE_GRID"], name, c.queue_sorting(), c.stats()["rays_traced"] / 3 / best / 1e6))
PY
done
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline.py tests/test_gpu_fullsize.py tests/test_gpu_mis.py tests/test_gpu_edges.py -x -q 2>&1 | tail -2
