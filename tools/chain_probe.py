#!/usr/bin/env python3
"""seed_mode 0 pixel chains on diningroom 1080p (the reference-seeded drop-in path): one warm-up render and one of `spp` samples,
for ncu captures of k_pixel_chain.  usage: chain_probe.py [spp=4]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from montecarlopathtracing_b200 import device, scenes

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4
sc = scenes.load("diningroom")
with device.Context(1920, 1080, 16, 1 << 30, seed_mode=0, run_seed=3) as c:
    c.upload_scene(sc.triangles, sc.materials, sc.nodes)
    c.set_camera(sc.camera)
    c.render(2)
    c.reset_stats()
    t = time.time(); c.render(spp); dt = time.time() - t
    print("chain: %d spp in %.3f s, %.0f Mrays/s" % (spp, dt, c.stats()["rays_traced"] / dt / 1e6))
