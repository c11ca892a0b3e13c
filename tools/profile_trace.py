"""Launches the intersection kernel once per traversal on random rays (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from montecarlopathtracing_b200 import device, scenes
from tests.util import random_rays

name = sys.argv[1] if len(sys.argv) > 1 else "diningroom"
travs = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "3,1").split(",")]
n = 1 << 20
sc = scenes.load(name)
with device.Context(64, 64, 8, 16) as ctx:
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    rays = random_rays(n, sc.nodes)
    d_rays = torch.from_numpy(rays.view(np.uint8)).cuda()
    d_hits = torch.empty(n * 48, dtype=torch.uint8, device="cuda")
    d_closest = torch.empty(n, dtype=torch.int32, device="cuda")
    for rep in range(2):
        for t in travs:
            ctx.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), d_closest.data_ptr(), t)
            ctx.synchronize()
print("ok")
