#!/bin/bash
O=gpurun_out; mkdir -p $O
for srt in 0 1; do MCPT_SORT=$srt python tools/mis_probe.py veach_mis 16384 16384 2>&1 | grep -E "^ref:|^mis:|count planes identical|total flux" ; done | tee $O/r3j_mis_sorted.log
for srt in 0 1; do MCPT_SORT=$srt python - <<'PY' 2>&1 | tee -a $O/r3j_mis_sorted.log
import os, time, sys
sys.path.insert(0, '.')
from montecarlopathtracing_b200 import device, scenes
for name, w, h, depth, spp in (("veach_mis", 1280, 720, 16, 256), ("cbox", 1024, 1024, 8, 128), ("diningroom", 1920, 1080, 16, 32)):
    sc = scenes.load(name)
    for mode in (1, 2):
        with device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3, sampling=mode) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes); c.set_camera(sc.camera)
            c.render(spp // 4); c.reset_stats()
            t = time.time(); c.render(spp); dt = time.time() - t
            st = c.stats()
            print("MCPT_SORT=%s %s mode %d: %d spp in %.3f s, %.0f Mrays/s (shadow rays included), sums checksum %.6e" % (os.environ.get("MCPT_SORT"), name, mode, spp, dt, st["rays_traced"] / dt / 1e6, float(c.read_sums().sum())))
PY
done
timeout 900 python -m pytest tests/test_gpu_mis.py -x -q 2>&1 | tail -3
