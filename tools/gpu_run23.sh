#!/bin/bash
# NEE + MIS: which of the two queues pays for its sort, per scene (MCPT_SORT bit 0 main queue, bit 1 shadow queue)
O=gpurun_out; mkdir -p $O; : > $O/r3k_mis_sort_mask.log
for srt in 0 1 2 3; do MCPT_SORT=$srt python - <<'PY' 2>&1 | tee -a $O/r3k_mis_sort_mask.log
import os, time, sys
sys.path.insert(0, '.')
from montecarlopathtracing_b200 import device, scenes
for name, w, h, depth, spp in (("veach_mis", 1280, 720, 16, 256), ("cbox", 1024, 1024, 8, 128), ("diningroom", 1920, 1080, 16, 32)):
    sc = scenes.load(name)
    for mode in (1,):
        with device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3, sampling=mode) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes); c.set_camera(sc.camera)
            c.render(spp // 4); c.reset_stats()
            best = 1e9
            for rep in range(3):
                t = time.time(); c.render(spp); best = min(best, time.time() - t)
            st = c.stats()
            print("MCPT_SORT=%s %s mode %d: %d spp in %.3f s (best of 3), %.0f Mrays/s (shadow rays included), sums checksum %.6e" % (os.environ.get("MCPT_SORT"), name, mode, spp, best, st["rays_traced"] / 3 / best / 1e6, float(c.read_sums().sum())))
PY
done
