#!/usr/bin/env python3
"""Mrays/s of the three benchmark scenes in the reference-faithful sampling mode (seed_mode 1 waves) and of the
seed_mode 0 pixel chains on diningroom, best of 3 - the per-scene figure of merit for A/B runs of library variants
(MCPT_LIB selects the library).  usage: scene_rates.py [label]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from montecarlopathtracing_b200 import device, scenes

label = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("MCPT_LIB", "in-tree")
out = []
for name, w, h, depth, spp in (("diningroom", 1920, 1080, 16, 128), ("cbox", 1024, 1024, 8, 128), ("veach_mis", 1280, 720, 16, 128)):
    sc = scenes.load(name)
    with device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3) as c:
        c.upload_scene(sc.triangles, sc.materials, sc.nodes)
        c.set_camera(sc.camera)
        c.render(spp)
        c.reset_stats()
        best = 1e9
        for rep in range(3):
            t = time.time(); c.render(spp); best = min(best, time.time() - t)
        out.append("%s %.0f" % (name, c.stats()["rays_traced"] / 3 / best / 1e6))
    if name == "diningroom":
        with device.Context(w, h, depth, 1 << 30, seed_mode=0, run_seed=3) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes)
            c.set_camera(sc.camera)
            c.render(4)
            c.reset_stats()
            best = 1e9
            for rep in range(3):
                t = time.time(); c.render(8); best = min(best, time.time() - t)
            out.append("chain %.0f" % (c.stats()["rays_traced"] / 3 / best / 1e6))
print("%-28s %s" % (label, "  ".join(out)))
