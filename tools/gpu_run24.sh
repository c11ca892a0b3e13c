#!/bin/bash
# NEE + MIS: self-tuned queue sorting against both pinned arrangements
O=gpurun_out; mkdir -p $O; : > $O/r3k_mis_sort_tuned.log
for srt in tuned 0 3; do
if [ $srt = tuned ]; then unset MCPT_SORT; else export MCPT_SORT=$srt; fi
SRT=$srt python - <<'PY' 2>&1 | tee -a $O/r3k_mis_sort_tuned.log
import os, time, sys
sys.path.insert(0, '.')
from montecarlopathtracing_b200 import device, scenes
for name, w, h, depth, spp in (("veach_mis", 1280, 720, 16, 256), ("cbox", 1024, 1024, 8, 256), ("diningroom", 1920, 1080, 16, 64)):
    sc = scenes.load(name)
    for mode in (1, 2):
        with device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3, sampling=mode) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes); c.set_camera(sc.camera)
            for k in range(4):
                c.render(spp // 2)      # one wave each (128 in flight at <= 2 Mpixel is automatic: spp // 2 <= 128)
            c.reset_stats()
            best = 1e9
            for rep in range(3):
                t = time.time(); c.render(spp); best = min(best, time.time() - t)
            st = c.stats()
            print("sort=%s %s mode %d: arrangement %d, %d spp in %.3f s (best of 3), %.0f Mrays/s (shadow rays included), sums checksum %.6e" % (os.environ["SRT"], name, mode, c.queue_sorting(), spp, best, st["rays_traced"] / 3 / best / 1e6, float(c.read_sums().sum())))
PY
done
timeout 1200 python -m pytest tests/test_gpu_mis.py tests/test_gpu_edges.py -x -q 2>&1 | tail -3
