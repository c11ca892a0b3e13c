#!/usr/bin/env python3
"""Do two independent wave pipelines on one GPU (two contexts, two streams, sample-sharded) fill each other's kernel tails?
Throughput of one context rendering 2N spp against two contexts rendering N spp each at the same time."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from montecarlopathtracing_b200 import device, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "diningroom"
w, h, depth = {"diningroom": (1920, 1080, 16), "cbox": (1024, 1024, 8), "veach_mis": (1280, 720, 16)}[name]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 256
sif = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sc = scenes.load(name)


def make(first):
    c = device.Context(w, h, depth, 1 << 30, seed_mode=1, run_seed=3, samples_in_flight=sif)
    c.upload_scene(sc.triangles, sc.materials, sc.nodes)
    c.set_camera(sc.camera)
    c.set_first_sample(first)
    return c


a = make(0)
a.render(128)
a.reset_stats()
best = 1e9
for rep in range(2):
    t = time.time(); a.render(2 * N); best = min(best, time.time() - t)
one = a.stats()["rays_traced"] / 2 / best / 1e6
b = make(1 << 20)
b.render(128)
a.reset_stats(); b.reset_stats()
best = 1e9
for rep in range(2):
    t = time.time()
    a.render_async(N); b.render_async(N)
    a.synchronize(); b.synchronize()
    best = min(best, time.time() - t)
two = (a.stats()["rays_traced"] + b.stats()["rays_traced"]) / 2 / best / 1e6
print("%s, %d samples in flight per context: one context %.0f Mrays/s, two concurrent contexts %.0f Mrays/s (%+.1f %%)" % (name, sif, one, two, 100 * (two / one - 1)))
a.close(); b.close()
