"""Development aid: magnitude of the path throughput (the colour buffer) bounce by bounce, through the stage-level ABI -
what the shader's division by 2 pi sees (the hardware divider takes a ~100-instruction slow path for operands far from
1).  usage: throughput_probe.py [scene] [w h depth]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from montecarlopathtracing_b200 import device, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "diningroom"
w, h, depth = (int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (960, 540, 16)
sc = scenes.load(name)
print("materials:", [(int(m["type"]), float(m["Ns"]), [round(float(x), 4) for x in m["kd"][:3]], [round(float(x), 3) for x in m["ka_ks"][:3]]) for m in sc.materials])
with device.Context(w, h, depth, 1 << 20, seed_mode=0, run_seed=5) as ctx:
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    ctx.set_camera(sc.camera)
    ctx.generate_rays()
    ctx.reset_color()
    for d in range(depth):
        ctx.intersect()
        ctx.shade()
        rays = ctx.read_rays()
        col = ctx.read_color()[:, :3]
        live = (rays["state"].view(np.uint32) & 0xFF000000) == 0
        c = np.abs(col[live])
        if len(c) == 0:
            break
        nz = c[c > 0]
        e = np.floor(np.log2(nz.astype(np.float64)))
        any_below = lambda t: float(((c < t) & (c > 0)).any(axis=1).mean())
        print("bounce %2d: live %7d | channels == 0: %.3f | median exp %4d, 1%% quantile %4d | rays with a channel in (0, 2^-40): %.3f, < 2^-100: %.3f, < 2^-120: %.3f, denormal: %.4f" % (
            d, live.sum(), float((c == 0).mean()), int(np.median(e)), int(np.quantile(e, 0.01)), any_below(2.0 ** -40), any_below(2.0 ** -100), any_below(2.0 ** -120), any_below(2.0 ** -126)))
